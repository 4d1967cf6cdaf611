// Consistent mass matrix in the CSR layout of the stiffness graph (one-off).
//   add_mass_matrix            src/solid_3D_ele/4C_solid_3D_ele_calc_lib.hpp:940-958
//   struct_calc_nlnstiffmass   src/solid_3D_ele/4C_solid_3D_ele_evaluate.cpp:107-119
//   mass rule != stiffness rule (tet10): second loop with the mean density  4C_solid_3D_ele_calc.cpp:201-240
// One thread per (element, row node a): m_ab = sum_gp N_a N_b rho w detJ for all b, added to the diagonal of
// the 3x3 node block (a,b) of the owned rows.
#pragma once
#include "fc_kernels_common.cuh"
#include "fc_nurbs.cuh"

namespace fc
{
  // tab: [ngp_mass] records of { w, N[nen], dN[nen][3] }
  template <int CT>
  __global__ void k_mass(const double* __restrict__ X, const int* __restrict__ conn, int n_ele, int n_row_nodes,
      const double* __restrict__ tab, const int64_t* __restrict__ node_rowbase, const int* __restrict__ node_rowlen,
      const unsigned char* __restrict__ slots, double density, double* __restrict__ M)
  {
    constexpr int NEN = Cell<CT>::nen, NGP = MassRule<CT>::ngp, REC = 1 + 4 * NEN;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= (long long)n_ele * NEN) return;
    const int e = (int)(tid / NEN), a = (int)(tid % NEN);
    const int* en = conn + (size_t)e * NEN;
    const int node = en[a];
    if (node >= n_row_nodes) return;
    double acc[NEN];
    for (int b = 0; b < NEN; ++b) acc[b] = 0.0;
    for (int g = 0; g < NGP; ++g)
    {
      const double* rec = tab + (size_t)g * REC;
      const double* N = rec + 1;
      const double* dN = rec + 1 + NEN;
      double J[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      for (int b = 0; b < NEN; ++b)
      {
        const double* x = X + 3 * (size_t)en[b];
        for (int d = 0; d < 3; ++d)
          for (int i = 0; i < 3; ++i) J[3 * d + i] += dN[3 * b + d] * x[i];
      }
      const double ifactor = N[a] * (rec[0] * det3(J)) * density;
      for (int b = 0; b < NEN; ++b) acc[b] += N[b] * ifactor;
    }
    const int64_t rowbase = node_rowbase[node];
    const int rowlen = node_rowlen[node];
    const unsigned char* sl = slots + ((size_t)e * NEN + a) * NEN;
    for (int b = 0; b < NEN; ++b)
      for (int i = 0; i < 3; ++i) red_add(M + rowbase + (int64_t)i * rowlen + 3 * (int)sl[b] + i, acc[b]);
  }

  // nurbs27: the same sum with N and dN of the element's own table (fc_nurbs.cuh), hex_27point weights gpw
  static __global__ void k_mass_nurbs(const double* __restrict__ X, const int* __restrict__ conn, int n_ele, int n_row_nodes,
      const double* __restrict__ tab, const double* __restrict__ gpw, const int64_t* __restrict__ node_rowbase,
      const int* __restrict__ node_rowlen, const unsigned char* __restrict__ slots, double density, double* __restrict__ M)
  {
    constexpr int NEN = 27, NGP = 27;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= (long long)n_ele * NEN) return;
    const int e = (int)(tid / NEN), a = (int)(tid % NEN);
    const int* en = conn + (size_t)e * NEN;
    const int node = en[a];
    if (node >= n_row_nodes) return;
    const double* T = tab + (size_t)e * NurbsTab::stride;
    double acc[NEN];
    for (int b = 0; b < NEN; ++b) acc[b] = 0.0;
    for (int g = 0; g < NGP; ++g)
    {
      const double* N = T + NurbsTab::oN + g * NEN;
      const double* dN = T + (size_t)g * NEN * 3;
      double J[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      for (int b = 0; b < NEN; ++b)
      {
        const double* x = X + 3 * (size_t)en[b];
        for (int d = 0; d < 3; ++d)
          for (int i = 0; i < 3; ++i) J[3 * d + i] += dN[3 * b + d] * x[i];
      }
      const double ifactor = N[a] * (gpw[g] * det3(J)) * density;
      for (int b = 0; b < NEN; ++b) acc[b] += N[b] * ifactor;
    }
    const int64_t rowbase = node_rowbase[node];
    const int rowlen = node_rowlen[node];
    const unsigned char* sl = slots + ((size_t)e * NEN + a) * NEN;
    for (int b = 0; b < NEN; ++b)
      for (int i = 0; i < 3; ++i) red_add(M + rowbase + (int64_t)i * rowlen + 3 * (int)sl[b] + i, acc[b]);
  }
}  // namespace fc
