// One-off setup kernels: node-Jacobian check, scatter slot map, halo pack.
#pragma once
#include "fc_kernels_common.cuh"
#include "fc_nurbs.cuh"

namespace fc
{
  // ensure_positive_jacobian_determinant_at_element_nodes
  // (src/solid_3D_ele/4C_solid_3D_ele_calc_lib.hpp:475-496): det J at every node's parameter
  // coordinates must be > 0.  One thread per element; err[0] counts failing elements, err[1] is
  // the lowest failing GID.
  template <int CT>
  __global__ void k_check_node_jacobians(const double* __restrict__ X, const int* __restrict__ conn,
      const int* __restrict__ ele_gid, int n_ele, const double* __restrict__ dNtab, unsigned char* ele_bad, int* err)
  {
    constexpr int NEN = Cell<CT>::nen, NGP = Cell<CT>::ngp;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_ele) return;
    bool bad = false;
    for (int p = 0; p < NEN; ++p)
    {
      const double* dN = ele_table<CT>(dNtab, e) + (size_t)(NGP + p) * NEN * 3;
      double J[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      for (int a = 0; a < NEN; ++a)
      {
        const int n = conn[(size_t)e * NEN + a];
        for (int d = 0; d < 3; ++d)
          for (int i = 0; i < 3; ++i) J[3 * d + i] += dN[3 * a + d] * X[3 * (size_t)n + i];
      }
      if (!(det3(J) > 0.0)) bad = true;
    }
    ele_bad[e] = bad ? 1 : 0;
    if (bad) report_error(err, ele_gid[e]);
  }

  // Geometry cache of the hex8 closed-form tensor-core kernel (fc_kernels_hex8nh_mma.cuh): per element
  //   geom[0..63]  G[a][b] = sum_gp w detJ grad N_a . grad N_b   (hex_8point, w = 1; row-major = accumulator layout)
  //   geom[64..71] detJ at the 8 Gauss points
  // evaluate_jacobian_mapping (4C_solid_3D_ele_calc_lib.hpp:435-448) on the reference coordinates: constant for a
  // mesh, computed once.  One block of 64 threads per element.
  static __global__ void k_hex8_geom_cache(const double* __restrict__ X, const int* __restrict__ conn, int n_ele, double* __restrict__ geom)
  {
    __shared__ double sX[8][3];
    __shared__ double sN[8][8][3];  // [gp][a] sqrt(detJ) grad N_a
    const int e = blockIdx.x, tid = threadIdx.x;
    if (e >= n_ele) return;
    if (tid < 24)
    {
      const int a = tid / 3, i = tid % 3;
      const int n = conn[(size_t)e * 8 + a], n0 = conn[(size_t)e * 8];
      sX[a][i] = X[3 * (size_t)n + i] - X[3 * (size_t)n0 + i];
    }
    __syncthreads();
    if (tid < 8)
    {
      const double p = 0.57735026918962576451;
      const double xi[3] = {hex_node_pos(tid, 0) > 0 ? p : -p, hex_node_pos(tid, 1) > 0 ? p : -p, hex_node_pos(tid, 2) > 0 ? p : -p};
      double dN[24], J[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, Ji[9];
      shape_deriv1<HEX8>(xi, dN);
      for (int a = 0; a < 8; ++a)
        for (int d = 0; d < 3; ++d)
          for (int i = 0; i < 3; ++i) J[3 * d + i] += dN[3 * a + d] * sX[a][i];
      const double detJ = inv3(J, Ji);
      const double sd = sqrt(detJ > 0.0 ? detJ : 0.0);
      for (int a = 0; a < 8; ++a)
        for (int d = 0; d < 3; ++d) sN[tid][a][d] = sd * (Ji[3 * d] * dN[3 * a] + Ji[3 * d + 1] * dN[3 * a + 1] + Ji[3 * d + 2] * dN[3 * a + 2]);
      geom[(size_t)e * 72 + 64 + tid] = detJ;
    }
    __syncthreads();
    {
      const int a = tid >> 3, b = tid & 7;
      double g = 0.0;
      for (int gp = 0; gp < 8; ++gp)
        for (int d = 0; d < 3; ++d) g += sN[gp][a][d] * sN[gp][b][d];
      geom[(size_t)e * 72 + tid] = g;
    }
  }

  // slots[e][a][b] = position of column node conn[e][b] in the (ascending) neighbour list of row
  // node conn[e][a]; 0xFF where a is not owned.  Replaces the per-element lower_bound searches of
  // SparseMatrix::assemble (src/core/linalg/src/sparse/4C_linalg_sparsematrix.cpp:512-543) by a
  // map computed once per graph.  err[0] counts pairs missing from the graph.
  static __global__ void k_build_slots(const int* __restrict__ conn, int n_ele, int nen, int n_row_nodes,
      const long long* __restrict__ nodeptr, const int* __restrict__ nodeadj, unsigned char* slots, int* err)
  {
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= (long long)n_ele * nen) return;
    const int e = (int)(tid / nen), a = (int)(tid % nen);
    const int na = conn[(size_t)e * nen + a];
    unsigned char* out = slots + ((size_t)e * nen + a) * nen;
    if (na >= n_row_nodes)
    {
      for (int b = 0; b < nen; ++b) out[b] = 0xFF;
      return;
    }
    const long long lo0 = nodeptr[na], hi0 = nodeptr[na + 1];
    for (int b = 0; b < nen; ++b)
    {
      const int nb = conn[(size_t)e * nen + b];
      long long lo = lo0, hi = hi0;
      while (lo < hi)
      {
        const long long mid = (lo + hi) >> 1;
        if (nodeadj[mid] < nb)
          lo = mid + 1;
        else
          hi = mid;
      }
      if (lo < hi0 && nodeadj[lo] == nb && lo - lo0 < 255)
        out[b] = (unsigned char)(lo - lo0);
      else
      {
        out[b] = 0xFF;
        atomicAdd(&err[0], 1);
      }
    }
  }

  // send buffer of the forward halo import (Epetra_Import pack step behind
  // Discretization::set_state, src/core/fem/src/discretization/4C_fem_discretization.cpp:542-553)
  static __global__ void k_halo_pack(const double* __restrict__ u, const int* __restrict__ send_lids, int n, double* out)
  {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = u[send_lids[i]];
  }
}  // namespace fc
