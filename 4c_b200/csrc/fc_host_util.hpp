// Host-side helpers shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "../../include/fourc_b200.h"
#include "fc_context.hpp"
#include "fc_fem.cuh"

namespace fc_host
{
  inline thread_local std::string g_create_error;

  inline int fail(fc_ctx* c, int code, const std::string& msg)
  {
    if (c) c->last_error = msg;
    return code;
  }

#define FC_CUDA(call)                                                                          \
  do                                                                                           \
  {                                                                                            \
    cudaError_t err__ = (call);                                                                \
    if (err__ != cudaSuccess)                                                                  \
      return fail(ctx, FC_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(err__));     \
  } while (0)

  template <class T>
  cudaError_t dev_upload(T*& dptr, const T* h, size_t n, cudaStream_t s)
  {
    if (dptr) cudaFree(dptr);
    dptr = nullptr;
    if (n == 0) return cudaSuccess;
    cudaError_t e = cudaMalloc((void**)&dptr, n * sizeof(T));
    if (e != cudaSuccess) return e;
    return cudaMemcpyAsync(dptr, h, n * sizeof(T), cudaMemcpyHostToDevice, s);
  }

  template <class T>
  void dev_free(T*& p)
  {
    if (p) cudaFree(p);
    p = nullptr;
  }

  inline bool is_device_ptr(const void* p)
  {
    if (!p) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess)
    {
      cudaGetLastError();
      return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
  }

  inline cudaError_t ensure_cap(double*& p, size_t& cap, size_t n)
  {
    if (cap >= n) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaMalloc((void**)&p, n * sizeof(double));
    if (e == cudaSuccess) cap = n;
    return e;
  }

  template <int CT>
  void fill_tables(std::vector<double>& dN, std::vector<double>& w)
  {
    constexpr int NEN = fc::Cell<CT>::nen, NGP = fc::Cell<CT>::ngp;
    dN.assign((size_t)(NGP + NEN + 1) * NEN * 3, 0.0);
    w.assign(NGP, 0.0);
    for (int g = 0; g < NGP; ++g)
    {
      double xi[3];
      fc::gauss_point<CT>(g, xi, &w[g]);
      fc::shape_deriv1<CT>(xi, &dN[(size_t)g * NEN * 3]);
    }
    for (int p = 0; p < NEN; ++p)
    {
      double xi[3];
      fc::node_param_coords(CT, p, xi);
      fc::shape_deriv1<CT>(xi, &dN[(size_t)(NGP + p) * NEN * 3]);
    }
    double xi[3];
    fc::centroid_param(CT, xi);
    fc::shape_deriv1<CT>(xi, &dN[(size_t)(NGP + NEN) * NEN * 3]);
    // the Gauss-point part again in the two orders k_nh_mma_gen reads it in (fc_kernels_nh_mma_gen.cuh):
    //   node-major  [nen][ngp][3]            - a thread owns a (Gauss point, direction) and walks over the nodes
    //   fragment    [3][KS * MT][32 lanes]   - lane (a8, q) of block (ks, mt) holds node 8 mt + a8 at Gauss point 4 ks + q, 0 = padding
    constexpr int MT = (NEN + 7) / 8, KS = (NGP + 3) / 4;
    const size_t o1 = dN.size(), o2 = o1 + (size_t)NEN * NGP * 3;
    dN.resize(o2 + (size_t)3 * KS * MT * 32, 0.0);
    for (int a = 0; a < NEN; ++a)
      for (int g = 0; g < NGP; ++g)
        for (int d = 0; d < 3; ++d) dN[o1 + ((size_t)a * NGP + g) * 3 + d] = dN[((size_t)g * NEN + a) * 3 + d];
    for (int d = 0; d < 3; ++d)
      for (int blk = 0; blk < KS * MT; ++blk)
        for (int lane = 0; lane < 32; ++lane)
        {
          const int g = 4 * (blk / MT) + (lane & 3), a = 8 * (blk % MT) + (lane >> 2);
          if (g < NGP && a < NEN) dN[o2 + ((size_t)d * KS * MT + blk) * 32 + lane] = dN[((size_t)g * NEN + a) * 3 + d];
        }
  }
}  // namespace fc_host
using namespace fc_host;
