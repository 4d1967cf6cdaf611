// struct_calc_stress: Gauss-point stresses and their extrapolation to the nodes.
//   calculate_stress                       src/solid_3D_ele/4C_solid_3D_ele_calc.cpp:347-390
//   assemble_stress_type_to_matrix_row     src/solid_3D_ele/4C_solid_3D_ele_calc_lib_io.hpp:167-190
//   pk2_to_cauchy  sigma = F S F^T / det F src/solid_3D_ele/4C_solid_3D_ele_utils.cpp:20-33   (F-bar: the modified F)
//   extrapolate_gp_quantity_to_nodes_and_assemble, nodal_average = true
//                                          src/core/fem/src/general/utils/4C_fem_general_utils_gauss_point_extrapolation.cpp:260-284, 418-427
// Strain output of the same action (assemble_strain_type_to_matrix_row, 4C_solid_3D_ele_calc_lib_io.hpp:113-153; Euler-Almansi
// and logarithmic strains: 4C_solid_3D_ele_utils.cpp:35-84).  The reference takes the 3x3 eigenproblems from LAPACK; here cyclic
// Jacobi rotations.
// Post-processing, not on the per-iteration path: one thread per (element, Gauss point) / (element, node).
#pragma once
#include "fc_nurbs.cuh"
#include "fc_kernels_generic.cuh"
#include "fc_kernels_hex8eas.cuh"
#include "fc_kernels_shell.cuh"

namespace fc
{
  // Gauss-point output kinds of k_gp_stress
  enum GpOutput
  {
    OUT_PK2 = 0,
    OUT_CAUCHY = 1,
    OUT_STRAIN_GL = 2,
    OUT_STRAIN_EA = 3,
    OUT_STRAIN_LOG = 4
  };

  // symmetric 3x3 (row-major): A = V diag(w) V^T, V[3 r + c] = component r of eigenvector c
  __device__ inline void sym_eig3(const double* Am, double* w, double* V)
  {
    double a[9];
    for (int k = 0; k < 9; ++k)
    {
      a[k] = Am[k];
      V[k] = (k % 4 == 0) ? 1.0 : 0.0;
    }
    for (int sweep = 0; sweep < 60; ++sweep)
    {
      const double off = a[1] * a[1] + a[2] * a[2] + a[5] * a[5];
      const double diag = a[0] * a[0] + a[4] * a[4] + a[8] * a[8];
      if (off <= 1.0e-34 * diag || off == 0.0) break;
      for (int p = 0; p < 2; ++p)
        for (int q = p + 1; q < 3; ++q)
        {
          const double apq = a[3 * p + q];
          if (apq == 0.0) continue;
          const double theta = (a[4 * q] - a[4 * p]) / (2.0 * apq);
          const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
          const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
          for (int k = 0; k < 3; ++k)
          {
            const double akp = a[3 * k + p], akq = a[3 * k + q];
            a[3 * k + p] = c * akp - s * akq;
            a[3 * k + q] = s * akp + c * akq;
          }
          for (int k = 0; k < 3; ++k)
          {
            const double apk = a[3 * p + k], aqk = a[3 * q + k];
            a[3 * p + k] = c * apk - s * aqk;
            a[3 * q + k] = s * apk + c * aqk;
          }
          for (int k = 0; k < 3; ++k)
          {
            const double vkp = V[3 * k + p], vkq = V[3 * k + q];
            V[3 * k + p] = c * vkp - s * vkq;
            V[3 * k + q] = s * vkp + c * vkq;
          }
        }
    }
    for (int k = 0; k < 3; ++k) w[k] = a[4 * k];
  }
  // out = V diag(f(w)) V^T with f = sqrt (mode 0) or log(sqrt(2 x + 1)) (mode 1)
  __device__ inline void sym_fn3(const double* Am, int mode, double* out)
  {
    double w[3], V[9];
    sym_eig3(Am, w, V);
    for (int k = 0; k < 3; ++k) w[k] = mode == 0 ? sqrt(w[k]) : log(sqrt(2.0 * w[k] + 1.0));
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) out[3 * r + c] = V[3 * r] * w[0] * V[3 * c] + V[3 * r + 1] * w[1] * V[3 * c + 1] + V[3 * r + 2] * w[2] * V[3 * c + 2];
  }

  // one Gauss point: S (2nd Piola-Kirchhoff), C (symmetric-6 right Cauchy-Green the material saw) and F (row-major, the
  // deformation gradient the reference hands to the output) -> out[6] in stress-like Voigt order xx yy zz xy yz xz
  __device__ inline void gp_output(int kind, const double* S, const double* C, const double* F, double* out)
  {
    const int vi6[6] = {0, 1, 2, 0, 1, 0}, vj6[6] = {0, 1, 2, 1, 2, 2};
    if (kind == OUT_PK2)
    {
      for (int v = 0; v < 6; ++v) out[v] = S[v];
      return;
    }
    if (kind == OUT_CAUCHY)
    {
      // sigma = F S F^T / det F; S = (00,11,22,01,12,02)
      const double Sm[9] = {S[0], S[3], S[5], S[3], S[1], S[4], S[5], S[4], S[2]};
      double FS[9];
      for (int i = 0; i < 3; ++i)
        for (int J = 0; J < 3; ++J) FS[3 * i + J] = F[3 * i] * Sm[J] + F[3 * i + 1] * Sm[3 + J] + F[3 * i + 2] * Sm[6 + J];
      const double idet = 1.0 / det3(F);
      for (int v = 0; v < 6; ++v)
      {
        const int i = vi6[v], j = vj6[v];
        out[v] = idet * (FS[3 * i] * F[3 * j] + FS[3 * i + 1] * F[3 * j + 1] + FS[3 * i + 2] * F[3 * j + 2]);
      }
      return;
    }
    const double Em[9] = {0.5 * (C[0] - 1.0), 0.5 * C[3], 0.5 * C[5], 0.5 * C[3], 0.5 * (C[1] - 1.0), 0.5 * C[4], 0.5 * C[5], 0.5 * C[4],
        0.5 * (C[2] - 1.0)};
    double R[9];
    if (kind == OUT_STRAIN_GL)
      for (int k = 0; k < 9; ++k) R[k] = Em[k];
    else if (kind == OUT_STRAIN_EA)
    {
      // e = F^-T E F^-1
      double iF[9], t[9];
      inv3(F, iF);
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) t[3 * i + j] = iF[i] * Em[j] + iF[3 + i] * Em[3 + j] + iF[6 + i] * Em[6 + j];
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) R[3 * i + j] = t[3 * i] * iF[j] + t[3 * i + 1] * iF[3 + j] + t[3 * i + 2] * iF[6 + j];
    }
    else
      sym_fn3(Em, 1, R);
    for (int v = 0; v < 6; ++v) out[v] = 0.5 * (R[3 * vi6[v] + vj6[v]] + R[3 * vj6[v] + vi6[v]]);
  }

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
    jacobian_mapping<NEN>(ele_table<CT>(A.dNtab, e) + (size_t)gp * ND, Xs, Nx);
    constexpr int NREC = mulf_records(NGP, FORM);
    if (form_mulf(FORM))
    {
      const double* hist = A.mulf + ((size_t)e * NREC + gp) * 18;
      double hl[18], delta[9];
      mulf_load(hist, hl);
      mulf_delta<NEN>(ele_table<CT>(A.dNtab, e) + (size_t)gp * ND, us, hl, delta);
      mat3_mul(delta, hl + 9, F);
    }
    else
      deformation_gradient<CT>(Xs, us, Nx, F);
    if (form_fbar(FORM))
    {
      double N0[ND], F0[9];
      if (FORM == FORM_MULF_FBAR)
      {
        double hl[18], delta[9];
        mulf_load(A.mulf + ((size_t)e * NREC + NGP) * 18, hl);
        mulf_delta<NEN>(ele_table<CT>(A.dNtab, e) + (size_t)(NGP + NEN) * ND, us, hl, delta);
        mat3_mul(delta, hl + 9, F0);
      }
      else
      {
        jacobian_mapping<NEN>(ele_table<CT>(A.dNtab, e) + (size_t)(NGP + NEN) * ND, Xs, N0);
        deformation_gradient<CT>(Xs, us, N0, F0);
      }
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
    gp_output(stress_type, S, C, F, gp_stress + ((size_t)e * NGP + gp) * 6);
  }

  // hex8 EAS: the enhanced Green-Lagrange strains of the element's alpha; the deformation gradient handed to the output is
  // F^enh = R U^enh with the rotation of the displacement-based F (compute_deformation_gradient_from_gl_strains,
  // 4C_solid_3D_ele_calc_lib.hpp:1257-1299).  The iteration data (Kaa^-1, Kda, s) is left alone, like calculate_stress.
  template <int NEAS, bool LINEAR = false>
  __global__ void k_gp_stress_eas(const EvalArgs A, int kind, double* __restrict__ gp_stress)
  {
    using Cfg = EasCfg<NEAS>;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= (long long)A.n_ele * 8) return;
    const int e = (int)(tid / 8), gp = (int)(tid % 8);
    double Xs[24], us[24];
    for (int a = 0; a < 8; ++a)
    {
      const int node = __ldg(A.conn + (size_t)e * 8 + a);
      for (int i = 0; i < 3; ++i)
      {
        Xs[3 * a + i] = __ldg(A.X + 3 * (size_t)node + i);
        us[3 * a + i] = __ldg(A.u + 3 * (size_t)node + i);
      }
    }
    double Tm[36], Nx[24], F[9], m[6], alpha[NEAS];
    const double detJ0 = eas_centroid_transformation(A.dNtab + (size_t)16 * 24, Xs, Tm);
    const double detJ = jacobian_mapping<8>(A.dNtab + (size_t)gp * 24, Xs, Nx);
    deformation_gradient<HEX8>(Xs, us, Nx, F);
    double Cd[6] = {F[0] * F[0] + F[3] * F[3] + F[6] * F[6], F[1] * F[1] + F[4] * F[4] + F[7] * F[7],
        F[2] * F[2] + F[5] * F[5] + F[8] * F[8], F[0] * F[1] + F[3] * F[4] + F[6] * F[7],
        F[1] * F[2] + F[4] * F[5] + F[7] * F[8], F[0] * F[2] + F[3] * F[5] + F[6] * F[8]};
    if (LINEAR)
    {
      const double H[9] = {F[0] - 1.0, F[1], F[2], F[3], F[4] - 1.0, F[5], F[6], F[7], F[8] - 1.0};
      Cd[0] = 2.0 * H[0] + 1.0;
      Cd[1] = 2.0 * H[4] + 1.0;
      Cd[2] = 2.0 * H[8] + 1.0;
      Cd[3] = H[1] + H[3];
      Cd[4] = H[5] + H[7];
      Cd[5] = H[2] + H[6];
    }
    double C[6] = {Cd[0], Cd[1], Cd[2], Cd[3], Cd[4], Cd[5]};
    eas_monomials(gp, detJ0 / detJ, m);
    for (int k = 0; k < NEAS; ++k) alpha[k] = A.eas[(size_t)e * Cfg::hist_doubles + Cfg::oAlpha + k];
    eas_enhance_cauchy_green<NEAS>(Tm, m, alpha, C);
    double S[6], cmat[36];
    material_evaluate(A.mat, C, S, cmat, A.fiber ? A.fiber + (size_t)e * A.fiber_ele + gp * A.fiber_gp : nullptr);
    // F^enh = F U_disp^-1 U_enh
    const double Cm[9] = {C[0], C[3], C[5], C[3], C[1], C[4], C[5], C[4], C[2]};
    const double Cdm[9] = {Cd[0], Cd[3], Cd[5], Cd[3], Cd[1], Cd[4], Cd[5], Cd[4], Cd[2]};
    double Ue[9], Ud[9], iUd[9], Rm[9], Fe[9];
    sym_fn3(Cm, 0, Ue);
    sym_fn3(Cdm, 0, Ud);
    inv3(Ud, iUd);
    mat3_mul(F, iUd, Rm);
    mat3_mul(Rm, Ue, Fe);
    if (LINEAR)  // the output sees the identity as deformation gradient
      for (int k = 0; k < 9; ++k) Fe[k] = (k % 4 == 0) ? 1.0 : 0.0;
    gp_output(kind, S, C, Fe, gp_stress + ((size_t)e * 8 + gp) * 6);
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

  // nurbs27: Gauss points -> control points.  The reference inverts the element's own basis at the Gauss points
  // (evaluate_gauss_points_to_nurbs_knots_extrapolation_matrix, 4C_fem_general_utils_gauss_point_extrapolation.cpp:386-420, with
  // evaluate_nurbs_base_shape_functions_at_gauss_points :116-146: M[gp][a] = N_a(xi_gp), base cell = the cell itself): one
  // 27 x 27 system with six right-hand sides per element.  One warp per element: Gauss-Jordan with partial pivoting on the
  // augmented matrix in shared memory, lane r owns row r; then the averaged assembly of the other cells.
  static __global__ void __launch_bounds__(128) k_extrapolate_nurbs27(const int* __restrict__ conn, int n_ele, int n_row_nodes,
      const double* __restrict__ tab, const int* __restrict__ node_elecount, const double* __restrict__ gp_stress, double* __restrict__ nodal)
  {
    constexpr int N = 27, W = N + 6;
    __shared__ double sA[4][N][W + 1];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int e = blockIdx.x * 4 + warp;
    if (e >= n_ele) return;
    double(*A)[W + 1] = sA[warp];
    const double* Ntab = tab + (size_t)e * NurbsTab::stride + NurbsTab::oN;  // [gp][a]
    if (lane < N)
    {
      for (int a = 0; a < N; ++a) A[lane][a] = Ntab[lane * N + a];
      for (int c = 0; c < 6; ++c) A[lane][N + c] = gp_stress[((size_t)e * N + lane) * 6 + c];
    }
    __syncwarp();
    for (int k = 0; k < N; ++k)
    {
      // pivot: the largest |A[r][k]| over r >= k
      double v = (lane >= k && lane < N) ? fabs(A[lane][k]) : -1.0;
      int idx = lane;
      for (int o = 16; o > 0; o >>= 1)
      {
        const double v2 = __shfl_xor_sync(0xFFFFFFFFu, v, o);
        const int i2 = __shfl_xor_sync(0xFFFFFFFFu, idx, o);
        if (v2 > v || (v2 == v && i2 < idx))
        {
          v = v2;
          idx = i2;
        }
      }
      if (idx != k)  // swap rows k and idx, one column per lane (33 columns: two passes)
        for (int c = lane; c < W; c += 32)
        {
          const double t = A[k][c];
          A[k][c] = A[idx][c];
          A[idx][c] = t;
        }
      __syncwarp();
      const double piv = 1.0 / A[k][k];
      const double fac = (lane < N && lane != k) ? A[lane][k] * piv : 0.0;
      __syncwarp();
      if (lane < N && lane != k)
        for (int c = k; c < W; ++c) A[lane][c] -= fac * A[k][c];
      __syncwarp();
    }
    if (lane < N)
    {
      const int node = conn[(size_t)e * N + lane];
      if (node < n_row_nodes)
      {
        const double inv = 1.0 / ((double)node_elecount[node] * A[lane][lane]);
        for (int c = 0; c < 6; ++c) red_add(nodal + 6 * (size_t)node + c, A[lane][N + c] * inv);
      }
    }
  }

  // solid shells (k_shell, fc_kernels_shell.cuh): ANS strains of the Gauss point, optionally enhanced with the element's alpha; the
  // deformation gradient handed to the output is the one consistent with those strains (R U), like the reference
  template <int CT, int NEAS>
  __global__ void k_gp_stress_shell(const EvalArgs A, int kind, double* __restrict__ gp_stress)
  {
    using Cfg = ShellCfg<CT, NEAS>;
    constexpr int NEN = Cfg::nen, NGP = Cfg::ngp, ND = Cfg::nd, NSP = Cfg::nsp;
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
    double spj[NSP * 18];  // Jref | Jcur of every sampling point
    for (int k = 0; k < NSP; ++k)
    {
      double xi[3], dN[ND];
      shell_sampling_point<CT>(k, xi);
      shape_deriv1<CT>(xi, dN);
      double* Jr = spj + 18 * k;
      for (int q = 0; q < 18; ++q) Jr[q] = 0.0;
      for (int a = 0; a < NEN; ++a)
        for (int d = 0; d < 3; ++d)
          for (int i = 0; i < 3; ++i)
          {
            Jr[3 * d + i] += dN[3 * a + d] * Xs[3 * a + i];
            Jr[9 + 3 * d + i] += dN[3 * a + d] * (Xs[3 * a + i] + us[3 * a + i]);
          }
    }
    double xi[3], wq, J[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, jc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, iJ[9], Nx[ND], F[9];
    gauss_point<CT>(gp, xi, &wq);
    const double* dNg = A.dNtab + (size_t)gp * ND;
    for (int a = 0; a < NEN; ++a)
      for (int d = 0; d < 3; ++d)
        for (int i = 0; i < 3; ++i)
        {
          J[3 * d + i] += __ldg(dNg + 3 * a + d) * Xs[3 * a + i];
          jc[3 * d + i] += __ldg(dNg + 3 * a + d) * (Xs[3 * a + i] + us[3 * a + i]);
        }
    const double detJ = inv3(J, iJ);
    for (int a = 0; a < NEN; ++a)
      for (int i = 0; i < 3; ++i) Nx[3 * a + i] = iJ[3 * i] * __ldg(dNg + 3 * a) + iJ[3 * i + 1] * __ldg(dNg + 3 * a + 1) + iJ[3 * i + 2] * __ldg(dNg + 3 * a + 2);
    deformation_gradient<CT>(Xs, us, Nx, F);
    double w[8], El[6], E[6];
    shell_weights<CT>(xi[0], xi[1], w);
    shell_local_strain<CT>(J, jc, w, spj, 18, 0, El);
    for (int r = 0; r < 6; ++r)
    {
      double v = 0.0;
      for (int q = 0; q < 6; ++q) v += voigt_transformation_entry(iJ, r, q) * El[q];
      E[r] = v;
    }
    double C[6] = {2.0 * E[0] + 1.0, 2.0 * E[1] + 1.0, 2.0 * E[2] + 1.0, E[3], E[4], E[5]};
    if (NEAS > 0)
    {
      double iJ0[9];
      const double detJ0 = eas_centroid_jacobian_n<NEN>(A.dNtab + (size_t)(NGP + NEN) * ND, Xs, iJ0);
      const double sc = detJ0 / detJ;
      const double m[6] = {sc * xi[0], sc * xi[1], sc * xi[2], sc * xi[0] * xi[1], sc * xi[0] * xi[2], sc * xi[1] * xi[2]};
      double ea[6] = {0, 0, 0, 0, 0, 0};
      for (int k = 0; k < Cfg::ne; ++k)
      {
        const double ma = m[shell_mono<CT, NEAS>(k)] * A.eas[(size_t)e * Cfg::hist_doubles + Cfg::oAlpha + k];
        for (int v = 0; v < 6; ++v) ea[v] += voigt_transformation_entry(iJ0, v, shell_row<CT, NEAS>(k)) * ma;
      }
      for (int v = 0; v < 3; ++v) C[v] += 2.0 * ea[v];
      for (int v = 3; v < 6; ++v) C[v] += ea[v];
    }
    double S[6], cmat[36];
    material_evaluate(A.mat, C, S, cmat, A.fiber ? A.fiber + (size_t)e * A.fiber_ele + gp * A.fiber_gp : nullptr);
    const double Cm[9] = {C[0], C[3], C[5], C[3], C[1], C[4], C[5], C[4], C[2]};
    double Cd[9], Ue[9], Ud[9], iUd[9], Rm[9], Fe[9];
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) Cd[3 * r + c] = F[r] * F[c] + F[3 + r] * F[3 + c] + F[6 + r] * F[6 + c];
    sym_fn3(Cm, 0, Ue);
    sym_fn3(Cd, 0, Ud);
    inv3(Ud, iUd);
    mat3_mul(F, iUd, Rm);
    mat3_mul(Rm, Ue, Fe);
    gp_output(kind, S, C, Fe, gp_stress + ((size_t)e * NGP + gp) * 6);
  }
}  // namespace fc
