// Stub of the cmake-generated 4C_config.hpp (template: /root/reference/src/config/4C_config.hpp.in)
// so that the reference's header-only dense kernels compile standalone in oracle/_ref.
#ifndef FOUR_C_CONFIG_HPP
#define FOUR_C_CONFIG_HPP
#define FOUR_C_NAMESPACE_OPEN namespace FourC {
#define FOUR_C_NAMESPACE_CLOSE }
#endif
