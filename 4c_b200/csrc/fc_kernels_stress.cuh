// struct_calc_stress: Gauss-point stresses and their extrapolation to the nodes.
//   calculate_stress                       src/solid_3D_ele/4C_solid_3D_ele_calc.cpp:347-390
//   assemble_stress_type_to_matrix_row     src/solid_3D_ele/4C_solid_3D_ele_calc_lib_io.hpp:167-190
//   pk2_to_cauchy  sigma = F S F^T / det F src/solid_3D_ele/4C_solid_3D_ele_utils.cpp:20-33   (F-bar: the modified F)
//   extrapolate_gp_quantity_to_nodes_and_assemble, nodal_average = true
//                                          src/core/fem/src/general/utils/4C_fem_general_utils_gauss_point_extrapolation.cpp:260-284, 418-427
// Post-processing, not on the per-iteration path: one thread per (element, Gauss point) / (element, node).
#pragma once
#include "fc_kernels_generic.cuh"

namespace fc
{
  template <int CT, int FORM>
  __global__ void k_gp_stress(const EvalArgs A, int stress_type, double* __restrict__ gp_stress)
  {
    constexpr int NEN = Cell<CT>::nen, NGP = Cell<CT>::ngp, ND = 3 * NEN;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= (long long)A.n_ele * NGP) return;
    const int e = (int)(tid / NGP), gp = (int)(tid % NGP);
    double Xs[ND], us[ND];
    for (int a = 0; a < NEN; ++a)
    {
      const int node = __ldg(A.conn + (size_t)e * NEN + a);
      for (int i = 0; i < 3; ++i)
      {
        Xs[3 * a + i] = __ldg(A.X + 3 * (size_t)node + i);
        us[3 * a + i] = __ldg(A.u + 3 * (size_t)node + i);
      }
    }
    double Nx[ND], F[9];
    jacobian_mapping<NEN>(A.dNtab + (size_t)gp * ND, Xs, Nx);
    if (FORM == FORM_MULF)
    {
      const double* hist = A.mulf + ((size_t)e * NGP + gp) * 18;
      double hl[18], delta[9];
      mulf_load(hist, hl);
      mulf_delta<NEN>(A.dNtab + (size_t)gp * ND, us, hl, delta);
      mat3_mul(delta, hl + 9, F);
    }
    else
      deformation_gradient<CT>(Xs, us, Nx, F);
    if (FORM == FORM_FBAR)
    {
      double N0[ND], F0[9];
      jacobian_mapping<NEN>(A.dNtab + (size_t)(NGP + NEN) * ND, Xs, N0);
      deformation_gradient<CT>(Xs, us, N0, F0);
      const double fbar = pow(det3(F0) / det3(F), 1.0 / 3.0);
      for (int i = 0; i < 9; ++i) F[i] *= fbar;
    }
    double C[6] = {F[0] * F[0] + F[3] * F[3] + F[6] * F[6], F[1] * F[1] + F[4] * F[4] + F[7] * F[7],
        F[2] * F[2] + F[5] * F[5] + F[8] * F[8], F[0] * F[1] + F[3] * F[4] + F[6] * F[7],
        F[1] * F[2] + F[4] * F[5] + F[7] * F[8], F[0] * F[2] + F[3] * F[5] + F[6] * F[8]};
    if (FORM == FORM_LINEAR)
    {
      // linear kinematics: strain = sym(grad u), the material law and the push-forward see F = I
      const double H[9] = {F[0] - 1.0, F[1], F[2], F[3], F[4] - 1.0, F[5], F[6], F[7], F[8] - 1.0};
      C[0] = 2.0 * H[0] + 1.0;
      C[1] = 2.0 * H[4] + 1.0;
      C[2] = 2.0 * H[8] + 1.0;
      C[3] = H[1] + H[3];
      C[4] = H[5] + H[7];
      C[5] = H[2] + H[6];
      for (int i = 0; i < 9; ++i) F[i] = (i % 4 == 0) ? 1.0 : 0.0;
    }
    double S[6], cmat[36];
    material_evaluate(A.mat, C, S, cmat, A.fiber ? A.fiber + (size_t)e * A.fiber_ele + gp * A.fiber_gp : nullptr);
    double* out = gp_stress + ((size_t)e * NGP + gp) * 6;
    if (stress_type == 0)
    {
      for (int v = 0; v < 6; ++v) out[v] = S[v];
      return;
    }
    // sigma = F S F^T / det F; S = (00,11,22,01,12,02)
    const double Sm[9] = {S[0], S[3], S[5], S[3], S[1], S[4], S[5], S[4], S[2]};
    double FS[9];
    for (int i = 0; i < 3; ++i)
      for (int J = 0; J < 3; ++J) FS[3 * i + J] = F[3 * i] * Sm[J] + F[3 * i + 1] * Sm[3 + J] + F[3 * i + 2] * Sm[6 + J];
    const double idet = 1.0 / det3(F);
    const int vi6[6] = {0, 1, 2, 0, 1, 0}, vj6[6] = {0, 1, 2, 1, 2, 2};
    for (int v = 0; v < 6; ++v)
    {
      const int i = vi6[v], j = vj6[v];
      out[v] = idet * (FS[3 * i] * F[3 * j] + FS[3 * i + 1] * F[3 * j + 1] + FS[3 * i + 2] * F[3 * j + 2]);
    }
  }

  // nodal[node][6] += (E gp_stress_e)[a] / (elements at the node), owned nodes only.  Emat: [nen][ngp]
  template <int CT>
  __global__ void k_extrapolate_to_nodes(const int* __restrict__ conn, int n_ele, int n_row_nodes, const double* __restrict__ Emat,
      const int* __restrict__ node_elecount, const double* __restrict__ gp_stress, double* __restrict__ nodal)
  {
    constexpr int NEN = Cell<CT>::nen, NGP = Cell<CT>::ngp;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= (long long)n_ele * NEN) return;
    const int e = (int)(tid / NEN), a = (int)(tid % NEN);
    const int node = conn[(size_t)e * NEN + a];
    if (node >= n_row_nodes) return;
    double v[6] = {0, 0, 0, 0, 0, 0};
    for (int g = 0; g < NGP; ++g)
    {
      const double w = Emat[a * NGP + g];
      const double* s = gp_stress + ((size_t)e * NGP + g) * 6;
      for (int c = 0; c < 6; ++c) v[c] += w * s[c];
    }
    const double inv = 1.0 / (double)node_elecount[node];
    for (int c = 0; c < 6; ++c) red_add(nodal + 6 * (size_t)node + c, v[c] * inv);
  }
}  // namespace fc
