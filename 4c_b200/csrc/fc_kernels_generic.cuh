// Generic element kernel: any cell type of the path x {displacement-based, F-bar} x any material,
// B-operator formulation  K_e = sum_gp w (B^T C B + K_geo [+ F-bar rank-1 term]),  f_e = sum_gp w B^T S.
// Follows the structure of SolidEleCalc::evaluate_nonlinear_force_stiffness_mass
// (src/solid_3D_ele/4C_solid_3D_ele_calc.cpp:110-240) with the library functions of
// src/solid_3D_ele/4C_solid_3D_ele_calc_lib.hpp (:435-448 Jacobian, :585-602 F, :708-760 B,
// :851-927 force/stiffness) and src/solid_3D_ele/4C_solid_3D_ele_calc_lib_fbar.hpp (:49-154).
//
// Work decomposition (not the reference's): a CTA owns EPB elements, nen threads per element.
//   phase 0  thread (e,a) gathers X_a, u_a of node a into shared memory
//   phase 1  thread (e,t) evaluates Gauss points t, t+nen, ...: F, N_XYZ, S, cmat, factors -> smem
//   phase 2  thread (e,a) accumulates the 3x3 blocks K_ab for a chunk of BCH column nodes b over
//            all Gauss points in FP64 registers, and f_a
//   phase 3  thread (e,a) adds its blocks into the CSR value array (owned rows only,
//            4C_linalg_sparsematrix.cpp:474) with fire-and-forget FP64 reductions
#pragma once
#include "fc_kernels_common.cuh"
#include "fc_nurbs.cuh"

namespace fc
{
  enum
  {
    FORM_DISP = 0,
    FORM_FBAR = 1,
    FORM_LINEAR = 2,  // KINEM linear: B-operator at F = I, strain = B u, no geometric stiffness
                      // (src/solid_3D_ele/4C_solid_3D_ele_calc_displacement_based_linear_kinematics.hpp:36-155)
    FORM_MULF = 3,    // PRESTRESS_TECH mulf: F = (I + u N_xyz_hist^T) F_hist with the history of the prestress updates
                      // (src/solid_3D_ele/4C_solid_3D_ele_calc_mulf.hpp:40-196, ..._calc_lib_mulf.hpp:24-84)
    FORM_MULF_FBAR = 6  // both (hex8, pyramid5; 4C_solid_3D_ele_calc_mulf_fbar.hpp:33-300): F-bar with the MULF deformation gradients
                        // of the Gauss points and of the centroid, whose history is one more record behind the Gauss points'
  };
  __host__ __device__ constexpr bool form_fbar(int form) { return form == FORM_FBAR || form == FORM_MULF_FBAR; }
  __host__ __device__ constexpr bool form_mulf(int form) { return form == FORM_MULF || form == FORM_MULF_FBAR; }
  // MULF history records of one element
  __host__ __device__ constexpr int mulf_records(int ngp, int form) { return form == FORM_MULF_FBAR ? ngp + 1 : ngp; }

  // MULF: deformation gradient update delta = I + u (invJ_hist dN)^T of one Gauss point.  dN [nen][3], us [nen][3],
  // invJh row-major; delta row-major [3 i + I]
  template <int NEN>
  __device__ __forceinline__ void mulf_delta(const double* __restrict__ dN, const double* us, const double* __restrict__ invJh, double* delta)
  {
    for (int i = 0; i < 9; ++i) delta[i] = (i % 4 == 0) ? 1.0 : 0.0;
    for (int a = 0; a < NEN; ++a)
    {
      const double d0 = __ldg(dN + 3 * a), d1 = __ldg(dN + 3 * a + 1), d2 = __ldg(dN + 3 * a + 2);
      double nh[3];
      for (int I = 0; I < 3; ++I) nh[I] = invJh[3 * I] * d0 + invJh[3 * I + 1] * d1 + invJh[3 * I + 2] * d2;
      for (int i = 0; i < 3; ++i)
        for (int I = 0; I < 3; ++I) delta[3 * i + I] += us[3 * a + i] * nh[I];
    }
  }
  // history record in memory: inverse Jacobian | deformation gradient, both column-major like the Core::LinAlg::Matrix<3, 3>
  // members of MulfHistoryData; the kernels work on row-major copies
  __device__ __forceinline__ void mulf_load(const double* __restrict__ hist, double* hl)
  {
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c)
      {
        hl[3 * r + c] = __ldg(hist + r + 3 * c);
        hl[9 + 3 * r + c] = __ldg(hist + 9 + r + 3 * c);
      }
  }
  __device__ __forceinline__ void mat3_mul(const double* A_, const double* B_, double* C_)
  {
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) C_[3 * i + j] = A_[3 * i] * B_[j] + A_[3 * i + 1] * B_[3 + j] + A_[3 * i + 2] * B_[6 + j];
  }

  template <int CT, int FORM>
  struct GenCfg
  {
    static constexpr int nen = Cell<CT>::nen;
    static constexpr int ngp = Cell<CT>::ngp;
    // elements per CTA and column nodes per register chunk, indexed by cell type (hex8, hex27, tet10, hex20, tet4, wedge6, pyramid5, hex18, nurbs27)
    static constexpr int epb_tab[9] = {8, 4, 12, 4, 32, 16, 16, 4, 4}, bch_tab[9] = {4, 9, 5, 5, 4, 3, 5, 6, 9};
    static constexpr int epb = epb_tab[CT], bch = bch_tab[CT];
    static constexpr int threads = epb * nen;
    // per-Gauss-point record: F(9) N_XYZ(3 nen) cmat(36) S(6) fac_mat fac_geo fac_force pad [H(3 nen) v(6)]
    static constexpr int oF = 0, oN = 9, oC = 9 + 3 * nen, oS = oC + 36, oFac = oS + 6, oH = oFac + 4,
                         oV = oH + 3 * nen;
    static constexpr int gpd = form_fbar(FORM) ? oV + 6 : oH;
    static constexpr int ele_doubles = 6 * nen + ngp * gpd;  // X, u, gp records
    static constexpr size_t smem_bytes = sizeof(double) * (size_t)epb * ele_doubles + 64;
  };

  // MULF history of every (element, Gauss point), hist [n_ele][ngp][18] = inverse Jacobian | deformation gradient.
  //   u == nullptr: the state before the first prestress update - inverse Jacobian of the reference mapping (what
  //                 MulfFormulation::evaluate stores at its first call, calc_mulf.hpp:59-63) and F = I
  //   u != nullptr: MulfFormulation::update_prestress (calc_mulf.hpp:150-171) with the displacements of the finished step:
  //                 F_hist <- (I + u N_xyz_hist^T) F_hist,  invJ_hist <- (I + u N_xyz_hist^T)^-T invJ_hist
  // nrec = ngp, or ngp + 1 with the centroid's record last (FORM_MULF_FBAR)
  template <int CT>
  __global__ void k_mulf_history(const double* __restrict__ X, const int* __restrict__ conn, int n_ele, const double* __restrict__ dNtab,
      const double* __restrict__ u, double* hist, int nrec)
  {
    constexpr int NEN = Cell<CT>::nen, NGP = Cell<CT>::ngp;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= (long long)n_ele * nrec) return;
    const int e = (int)(tid / nrec), gp = (int)(tid % nrec);
    const double* dN = ele_table<CT>(dNtab, e) + (size_t)(gp < NGP ? gp : NGP + NEN) * NEN * 3;
    double* H = hist + (size_t)tid * 18;
    if (!u)
    {
      double J[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, iJ[9];
      for (int a = 0; a < NEN; ++a)
      {
        const int n = conn[(size_t)e * NEN + a];
        for (int d = 0; d < 3; ++d)
          for (int i = 0; i < 3; ++i) J[3 * d + i] += dN[3 * a + d] * X[3 * (size_t)n + i];
      }
      inv3(J, iJ);
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c)
        {
          H[r + 3 * c] = iJ[3 * r + c];
          H[9 + r + 3 * c] = r == c ? 1.0 : 0.0;
        }
      return;
    }
    double us[3 * NEN], hl[18], delta[9], Ft[9], id[9];
    for (int a = 0; a < NEN; ++a)
    {
      const int n = conn[(size_t)e * NEN + a];
      for (int i = 0; i < 3; ++i) us[3 * a + i] = u[3 * (size_t)n + i];
    }
    for (int r = 0; r < 3; ++r)  // plain loads: the record is rewritten below
      for (int c = 0; c < 3; ++c)
      {
        hl[3 * r + c] = H[r + 3 * c];
        hl[9 + 3 * r + c] = H[9 + r + 3 * c];
      }
    for (int i = 0; i < 9; ++i) delta[i] = (i % 4 == 0) ? 1.0 : 0.0;
    for (int a = 0; a < NEN; ++a)
    {
      double nh[3];
      for (int I = 0; I < 3; ++I) nh[I] = hl[3 * I] * dN[3 * a] + hl[3 * I + 1] * dN[3 * a + 1] + hl[3 * I + 2] * dN[3 * a + 2];
      for (int i = 0; i < 3; ++i)
        for (int I = 0; I < 3; ++I) delta[3 * i + I] += us[3 * a + i] * nh[I];
    }
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) Ft[3 * i + j] = delta[3 * i] * hl[9 + j] + delta[3 * i + 1] * hl[12 + j] + delta[3 * i + 2] * hl[15 + j];
    inv3(delta, id);
    for (int i = 0; i < 3; ++i)
      for (int d = 0; d < 3; ++d) H[i + 3 * d] = id[i] * hl[d] + id[3 + i] * hl[3 + d] + id[6 + i] * hl[6 + d];  // delta^-T invJ
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) H[9 + r + 3 * c] = Ft[3 * r + c];
  }

  // J = dN X^T, N_XYZ = J^-1 dN.  dN, Nxyz: [nen][3]; Xs: [nen][3].  returns det J.
  template <int NEN>
  __device__ __forceinline__ double jacobian_mapping(const double* __restrict__ dN, const double* Xs, double* Nxyz)
  {
    double J[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int a = 0; a < NEN; ++a)
    {
      const double d0 = __ldg(dN + 3 * a), d1 = __ldg(dN + 3 * a + 1), d2 = __ldg(dN + 3 * a + 2);
      const double x0 = Xs[3 * a], x1 = Xs[3 * a + 1], x2 = Xs[3 * a + 2];
      J[0] += d0 * x0; J[1] += d0 * x1; J[2] += d0 * x2;
      J[3] += d1 * x0; J[4] += d1 * x1; J[5] += d1 * x2;
      J[6] += d2 * x0; J[7] += d2 * x1; J[8] += d2 * x2;
    }
    double iJ[9];
    const double det = inv3(J, iJ);
    for (int a = 0; a < NEN; ++a)
    {
      const double d0 = __ldg(dN + 3 * a), d1 = __ldg(dN + 3 * a + 1), d2 = __ldg(dN + 3 * a + 2);
      Nxyz[3 * a + 0] = iJ[0] * d0 + iJ[1] * d1 + iJ[2] * d2;
      Nxyz[3 * a + 1] = iJ[3] * d0 + iJ[4] * d1 + iJ[5] * d2;
      Nxyz[3 * a + 2] = iJ[6] * d0 + iJ[7] * d1 + iJ[8] * d2;
    }
    return det;
  }

  // F[3*i+I]: hex8 from current coordinates, others from displacements (calc_lib.hpp:585-602)
  template <int CT>
  __device__ __forceinline__ void deformation_gradient(const double* Xs, const double* us, const double* Nxyz, double* F)
  {
    constexpr int NEN = Cell<CT>::nen;
    for (int i = 0; i < 9; ++i) F[i] = 0.0;
    for (int a = 0; a < NEN; ++a)
    {
      double v[3];
      for (int i = 0; i < 3; ++i) v[i] = CT == HEX8 ? Xs[3 * a + i] + us[3 * a + i] : us[3 * a + i];
      for (int i = 0; i < 3; ++i)
        for (int I = 0; I < 3; ++I) F[3 * i + I] += v[i] * Nxyz[3 * a + I];
    }
    if (CT != HEX8)
    {
      F[0] += 1.0;
      F[4] += 1.0;
      F[8] += 1.0;
    }
  }

  // strain-like B block of node a: Ba[3*v+i] = B(v, 3a+i)          calc_lib.hpp:708-760
  __device__ __forceinline__ void b_block(const double* F, double n0, double n1, double n2, double* Ba)
  {
#pragma unroll
    for (int i = 0; i < 3; ++i)
    {
      const double f0 = F[3 * i], f1 = F[3 * i + 1], f2 = F[3 * i + 2];
      Ba[0 + i] = f0 * n0;
      Ba[3 + i] = f1 * n1;
      Ba[6 + i] = f2 * n2;
      Ba[9 + i] = f0 * n1 + f1 * n0;
      Ba[12 + i] = f1 * n2 + f2 * n1;
      Ba[15 + i] = f2 * n0 + f0 * n2;
    }
  }

  template <int CT, int FORM, bool WANT_K>
  __global__ void __launch_bounds__(GenCfg<CT, FORM>::threads) k_generic(const EvalArgs A, const unsigned char* __restrict__ ele_bad)
  {
    using Cfg = GenCfg<CT, FORM>;
    constexpr int NEN = Cfg::nen, NGP = Cfg::ngp, ND = 3 * NEN, GPD = Cfg::gpd;
    extern __shared__ double smem[];
    const int el = threadIdx.x / NEN, a = threadIdx.x % NEN;
    const int e = blockIdx.x * Cfg::epb + el;
    const bool active = e < A.n_ele;
    double* Xs = smem + (size_t)el * Cfg::ele_doubles;
    double* us = Xs + 3 * NEN;
    double* gpdat = us + 3 * NEN;
    __shared__ int s_bad[Cfg::epb];
    if (a == 0) s_bad[el] = (active && ele_bad && ele_bad[e]) ? 1 : 0;

    int node = 0;
    if (active)
    {
      node = __ldg(A.conn + (size_t)e * NEN + a);
      for (int i = 0; i < 3; ++i)
      {
        Xs[3 * a + i] = __ldg(A.X + 3 * (size_t)node + i);
        us[3 * a + i] = __ldg(A.u + 3 * (size_t)node + i);
      }
    }
    __syncthreads();

    // ---- phase 1: Gauss-point records ----------------------------------------------------------
    if (active)
    {
      constexpr int NREC = mulf_records(NGP, FORM);
      double N0[form_fbar(FORM) ? ND : 1], iF0[9], detF0 = 1.0;
      if (form_fbar(FORM))
      {
        double F0[9];
        jacobian_mapping<NEN>(ele_table<CT>(A.dNtab, e) + (size_t)(NGP + NEN) * ND, Xs, N0);
        if (FORM == FORM_MULF_FBAR)
        {
          // MulfFBarFormulation::prepare: N_XYZ_0 of the reference Jacobian, F_0 from the centroid's history record
          double hl[18], delta[9];
          mulf_load(A.mulf + ((size_t)e * NREC + NGP) * 18, hl);
          mulf_delta<NEN>(ele_table<CT>(A.dNtab, e) + (size_t)(NGP + NEN) * ND, us, hl, delta);
          mat3_mul(delta, hl + 9, F0);
        }
        else
          deformation_gradient<CT>(Xs, us, N0, F0);
        detF0 = inv3(F0, iF0);
      }
      for (int gp = a; gp < NGP; gp += NEN)
      {
        double* R = gpdat + (size_t)gp * GPD;
        double* Nx = R + Cfg::oN;
        const double detJ = jacobian_mapping<NEN>(ele_table<CT>(A.dNtab, e) + (size_t)gp * ND, Xs, Nx);
        double F[9];
        if (form_mulf(FORM))
        {
          // evaluate_mulf_spatial_material_mapping; the B-operator and K_geo below keep the N_XYZ of the reference mapping
          const double* hist = A.mulf + ((size_t)e * NREC + gp) * 18;
          double hl[18], delta[9];
          mulf_load(hist, hl);
          mulf_delta<NEN>(ele_table<CT>(A.dNtab, e) + (size_t)gp * ND, us, hl, delta);
          mat3_mul(delta, hl + 9, F);
        }
        else
          deformation_gradient<CT>(Xs, us, Nx, F);
        const double intfac = detJ * __ldg(A.gpw + gp);
        double fbar = 1.0, iF[9];
        const double detF = form_fbar(FORM) ? inv3(F, iF) : (FORM == FORM_LINEAR ? 1.0 : det3(F));
        if (form_fbar(FORM)) fbar = pow(detF0 / detF, 1.0 / 3.0);
        if (!(detJ > 0.0) || detF == 0.0 || !isfinite(detF) || !isfinite(fbar)) s_bad[el] = 1;
        // right Cauchy-Green of the (modified) deformation gradient, symmetric-6
        double C[6];
        if (FORM == FORM_LINEAR)
        {
          // gl_strain = B_lin u = sym(grad u) (calc_lib.hpp:683-695, 773-799); the material law sees it as C = 2 E + I
          // (4C_mat_elasthyper_service.cpp:80-88), and the B-operator below is built with F = I
          const double H[9] = {F[0] - 1.0, F[1], F[2], F[3], F[4] - 1.0, F[5], F[6], F[7], F[8] - 1.0};
          C[0] = 2.0 * H[0] + 1.0;
          C[1] = 2.0 * H[4] + 1.0;
          C[2] = 2.0 * H[8] + 1.0;
          C[3] = H[1] + H[3];
          C[4] = H[5] + H[7];
          C[5] = H[2] + H[6];
          for (int i = 0; i < 9; ++i) F[i] = (i % 4 == 0) ? 1.0 : 0.0;
        }
        else
        {
          const double s2 = fbar * fbar;
          C[0] = s2 * (F[0] * F[0] + F[3] * F[3] + F[6] * F[6]);
          C[1] = s2 * (F[1] * F[1] + F[4] * F[4] + F[7] * F[7]);
          C[2] = s2 * (F[2] * F[2] + F[5] * F[5] + F[8] * F[8]);
          C[3] = s2 * (F[0] * F[1] + F[3] * F[4] + F[6] * F[7]);
          C[4] = s2 * (F[1] * F[2] + F[4] * F[5] + F[7] * F[8]);
          C[5] = s2 * (F[0] * F[2] + F[3] * F[5] + F[6] * F[8]);
        }
        double S[6], cmat[36];
        material_evaluate(A.mat, C, S, cmat, A.fiber ? A.fiber + (size_t)e * A.fiber_ele + gp * A.fiber_gp : nullptr);
        for (int i = 0; i < 9; ++i) R[Cfg::oF + i] = F[i];
        for (int i = 0; i < 36; ++i) R[Cfg::oC + i] = cmat[i];
        for (int i = 0; i < 6; ++i) R[Cfg::oS + i] = S[i];
        R[Cfg::oFac + 0] = intfac * fbar;  // elastic stiffness   (calc_fbar.hpp:176-178)
        R[Cfg::oFac + 1] = FORM == FORM_LINEAR ? 0.0 : intfac / fbar;  // geometric stiffness (:179-180); none with linear kinematics
        R[Cfg::oFac + 2] = intfac / fbar;  // internal force      (:164-166)
        if (form_fbar(FORM))
        {
          // H operator (lib_fbar.hpp:71-104) and v = (w f/3) cmat:C_gp - (w/f/3) S (lib_fbar.hpp:133-147)
          for (int b = 0; b < NEN; ++b)
            for (int i = 0; i < 3; ++i)
            {
              double h = 0.0;
              for (int d = 0; d < 3; ++d)
                h += iF0[3 * d + i] * N0[3 * b + d] - iF[3 * d + i] * Nx[3 * b + d];
              R[Cfg::oH + 3 * b + i] = h;
            }
          const double is2 = 1.0 / (fbar * fbar);  // unmodified C, strain-like (doubled shear)
          const double rcg[6] = {C[0] * is2, C[1] * is2, C[2] * is2, 2.0 * C[3] * is2, 2.0 * C[4] * is2, 2.0 * C[5] * is2};
          for (int v = 0; v < 6; ++v)
          {
            double ccg = 0.0;
            for (int w = 0; w < 6; ++w) ccg += cmat[6 * v + w] * rcg[w];
            R[Cfg::oV + v] = intfac * fbar / 3.0 * ccg - intfac / fbar / 3.0 * S[v];
          }
        }
      }
    }
    __syncthreads();
    if (!active) return;
    if (s_bad[el])
    {
      if (a == 0) report_error(A.err, __ldg(A.ele_gid + e));
      return;
    }

    // ---- phase 2 + 3 -----------------------------------------------------------------------------
    const bool owned = node < A.n_row_nodes;
    const int64_t rowbase = (owned && A.node_rowbase) ? __ldg(A.node_rowbase + node) : -1;
    const int rowlen = A.node_rowlen ? __ldg(A.node_rowlen + node) : 0;
    double fa[3] = {0, 0, 0};
    constexpr int BCH = WANT_K ? Cfg::bch : NEN;
    for (int b0 = 0; b0 < NEN; b0 += BCH)
    {
      double acc[WANT_K ? BCH * 9 : 1];
      if (WANT_K)
        for (int i = 0; i < BCH * 9; ++i) acc[i] = 0.0;
      for (int gp = 0; gp < NGP; ++gp)
      {
        const double* R = gpdat + (size_t)gp * GPD;
        const double* F = R + Cfg::oF;
        const double* Nx = R + Cfg::oN;
        double Ba[18];
        b_block(F, Nx[3 * a], Nx[3 * a + 1], Nx[3 * a + 2], Ba);
        if (b0 == 0)
        {
          const double ff = R[Cfg::oFac + 2];
          for (int i = 0; i < 3; ++i)
          {
            double s = 0.0;
            for (int v = 0; v < 6; ++v) s += Ba[3 * v + i] * R[Cfg::oS + v];
            fa[i] += ff * s;
          }
        }
        if (WANT_K)
        {
          // CBa[3*v+i] = fac_mat * sum_w cmat[v][w] Ba[w][i]
          double CBa[18];
          const double fm = R[Cfg::oFac + 0];
          for (int v = 0; v < 6; ++v)
            for (int i = 0; i < 3; ++i)
            {
              double s = 0.0;
              for (int w = 0; w < 6; ++w) s += R[Cfg::oC + 6 * v + w] * Ba[3 * w + i];
              CBa[3 * v + i] = fm * s;
            }
          const double* S = R + Cfg::oS;
          const double fg = R[Cfg::oFac + 1];
          const double n0 = Nx[3 * a], n1 = Nx[3 * a + 1], n2 = Nx[3 * a + 2];
          const double SN0 = fg * (S[0] * n0 + S[3] * n1 + S[5] * n2);
          const double SN1 = fg * (S[3] * n0 + S[1] * n1 + S[4] * n2);
          const double SN2 = fg * (S[5] * n0 + S[4] * n1 + S[2] * n2);
          double q[3] = {0, 0, 0};
          if (form_fbar(FORM))
            for (int i = 0; i < 3; ++i)
              for (int v = 0; v < 6; ++v) q[i] += Ba[3 * v + i] * R[Cfg::oV + v];
#pragma unroll
          for (int bb = 0; bb < BCH; ++bb)
          {
            const int b = b0 + bb;
            if (b < NEN)
            {
              const double m0 = Nx[3 * b], m1 = Nx[3 * b + 1], m2 = Nx[3 * b + 2];
              double Bb[18];
              b_block(F, m0, m1, m2, Bb);
              const double geo = SN0 * m0 + SN1 * m1 + SN2 * m2;
#pragma unroll
              for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < 3; ++j)
                {
                  double s = acc[9 * bb + 3 * i + j];
#pragma unroll
                  for (int v = 0; v < 6; ++v) s += CBa[3 * v + i] * Bb[3 * v + j];
                  if (i == j) s += geo;
                  if (form_fbar(FORM)) s += q[i] * R[Cfg::oH + 3 * b + j];
                  acc[9 * bb + 3 * i + j] = s;
                }
            }
          }
        }
      }
      if (WANT_K)
      {
        if (A.Ke_out)
        {
          double* Ke = A.Ke_out + (size_t)e * ND * ND;
          for (int bb = 0; bb < BCH; ++bb)
            if (b0 + bb < NEN)
              for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) Ke[(3 * a + i) + (size_t)ND * (3 * (b0 + bb) + j)] = acc[9 * bb + 3 * i + j];
        }
        if (A.K && rowbase >= 0)
        {
          const unsigned char* sl = A.slots + (size_t)e * NEN * NEN + a * NEN;
          for (int i = 0; i < 3; ++i)
          {
            if (A.dbc && A.dbc[3 * node + i]) continue;
            double* row = A.K + rowbase + (int64_t)i * rowlen;
            for (int bb = 0; bb < BCH; ++bb)
              if (b0 + bb < NEN)
              {
                double* p = row + 3 * (int)sl[b0 + bb];
                red_add(p + 0, acc[9 * bb + 3 * i + 0]);
                red_add(p + 1, acc[9 * bb + 3 * i + 1]);
                red_add(p + 2, acc[9 * bb + 3 * i + 2]);
              }
          }
        }
      }
    }
    if (A.fe_out)
      for (int i = 0; i < 3; ++i) A.fe_out[(size_t)e * ND + 3 * a + i] = fa[i];
    if (A.f && owned)
      for (int i = 0; i < 3; ++i) red_add(A.f + 3 * (size_t)node + i, fa[i]);
  }
}  // namespace fc
