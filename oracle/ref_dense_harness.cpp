// ORACLE — TEST INFRASTRUCTURE ONLY.
// Thin C ABI over the REFERENCE's own header-only dense kernels, compiled from the sources where
// they lie under /root/reference (never copied into this repo):
//   src/core/linalg/src/dense/4C_linalg_fixedsizematrix.hpp
//   src/core/linalg/src/dense/4C_linalg_fixedsizematrix_voigt_notation.hpp (inline members only)
// Output: oracle/_ref/libfc_ref_dense.so.  Used by tests/test_oracle_ref.py to pin the
// restatement (oracle/fc_oracle_dense.hpp) bit-exactly.
#include <stdexcept>
#include <string>

#include "4C_linalg_fixedsizematrix.hpp"
#include "4C_linalg_fixedsizematrix_voigt_notation.hpp"

// the only non-header symbol the headers need (real definition needs Boost+MPI:
// src/core/utils/src/exceptions/4C_utils_exceptions.cpp)
namespace FourC::Core::Internal
{
  [[noreturn]] void throw_error(const ErrorHelper& error_helper, const std::string& formatted_message)
  {
    throw std::runtime_error(std::string(error_helper.file_name) + ": " + formatted_message);
  }
}  // namespace FourC::Core::Internal

using namespace FourC;
namespace D = FourC::Core::LinAlg::DenseFunctions;

extern "C"
{
  void ref_multiply_nn_6_6_24(double* out, const double* l, const double* r) { D::multiply<double, 6, 6, 24>(out, l, r); }
  void ref_multiply_tn_24_6_24(double fac, double* out, const double* l, const double* r)
  {
    D::multiply_tn<double, 24, 6, 24>(1.0, out, fac, l, r);
  }
  void ref_multiply_nt_3_8_3(double* out, const double* l, const double* r) { D::multiply_nt<double, 3, 8, 3>(out, l, r); }
  void ref_multiply_tn_3_3_3(double* out, const double* l, const double* r) { D::multiply_tn<double, 3, 3, 3>(out, l, r); }
  double ref_invert3x3(const double* in, double* out) { return D::invert3x3<double>(out, in); }
  double ref_determinant3x3(const double* in) { return D::determinant<double, 3, 3>(in); }
  double ref_voigt_determinant(int strain, const double* t)
  {
    Core::LinAlg::Matrix<6, 1> v(t, false);
    return strain ? Core::LinAlg::Voigt::Strains::determinant(v) : Core::LinAlg::Voigt::Stresses::determinant(v);
  }
  void ref_voigt_invariants(int strain, const double* t, double* prinv)
  {
    Core::LinAlg::Matrix<6, 1> v(t, false);
    Core::LinAlg::Matrix<3, 1> p(prinv, true);
    if (strain)
      Core::LinAlg::Voigt::Strains::invariants_principal(p, v);
    else
      Core::LinAlg::Voigt::Stresses::invariants_principal(p, v);
  }
}
