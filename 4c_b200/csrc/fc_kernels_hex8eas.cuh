// hex8 with enhanced assumed strains (TECH eas_mild: 9 modes, eas_full: 21 modes, shell_eas: 7 modes), SURVEY 8f F4.
//   EASFormulation, history, condensation      src/solid_3D_ele/4C_solid_3D_ele_calc_eas.hpp:44-291
//   M(xi), Mtilde = detJ0/detJ T0^-T M, integrate_eas, update_alpha
//                                              src/solid_3D_ele/4C_solid_3D_ele_calc_eas_helpers.hpp:189-228, 258-420, 511-585
//   T0^-T                                      src/solid_3D_ele/4C_solid_3D_ele_calc_lib.hpp:1177-1244
//   order of one evaluation / recover          src/solid_3D_ele/4C_solid_3D_ele_calc.cpp:110-240, 244-273
// One warp per element.  Every column of M(xi) has one entry, so Mtilde(:, k) = m_k T(:, row_k) with the six monomials
// m = detJ0/detJ (r, s, t, rs, rt, st); with G = T^T cmat T, H = cmat T and T^T S per Gauss point
//   Kaa(k, l) = sum_gp fac m_k m_l G(row_k, row_l),  Kda(d, l) = sum_gp fac m_l (B(:, d)^T H)(row_l),  s(k) = sum_gp fac m_k (T^T S)(row_k).
// Kaa is inverted in shared memory (Gauss-Jordan with partial pivoting; the reference calls LAPACK), then
//   f += -Kda Kaa^-1 s,  K += -Kda Kaa^-1 Kda^T,  and Kaa^-1, Kda, s go to the element's history for the next recover.
#pragma once
#include "fc_kernels_generic.cuh"

namespace fc
{
  // row of the entry of column k of M(xi), and which monomial it is (0 r, 1 s, 2 t, 3 rs, 4 rt, 5 st); three bits per column
  constexpr unsigned long long eas_pack(const int (&v)[21])
  {
    unsigned long long p = 0;
    for (int k = 0; k < 21; ++k) p |= (unsigned long long)v[k] << (3 * k);
    return p;
  }
  constexpr unsigned long long EAS_ROWS = eas_pack({0, 1, 2, 3, 3, 4, 4, 5, 5, 3, 3, 4, 4, 5, 5, 0, 0, 1, 1, 2, 2});
  constexpr unsigned long long EAS_MONO = eas_pack({0, 1, 2, 0, 1, 1, 2, 0, 2, 4, 5, 3, 4, 3, 5, 3, 4, 3, 5, 4, 5});
  // eas_sosh8 (TECH shell_eas, 7 modes, t = thickness direction; 4C_solid_3D_ele_calc_eas_helpers.hpp:325-349):
  // r, s, t on the normal strains, r and s on the in-plane shear, rt and st on the thickness strain
  constexpr unsigned long long EAS_ROWS_SH7 = eas_pack({0, 1, 2, 3, 3, 2, 2, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0});
  constexpr unsigned long long EAS_MONO_SH7 = eas_pack({0, 1, 2, 0, 1, 4, 5, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0});
  template <int NEAS>
  __host__ __device__ constexpr int eas_row(int k) { return (int)(((NEAS == 7 ? EAS_ROWS_SH7 : EAS_ROWS) >> (3 * k)) & 7ull); }
  template <int NEAS>
  __host__ __device__ constexpr int eas_mono(int k) { return (int)(((NEAS == 7 ? EAS_MONO_SH7 : EAS_MONO) >> (3 * k)) & 7ull); }

  template <int NEAS>
  struct EasCfg
  {
    static constexpr int neas = NEAS;
    // element history in the order EASFormulation::pack writes it (matrices column-major)
    static constexpr int oAlphaInc = 0, oAlpha = NEAS, oS = 2 * NEAS, oInvKaa = 3 * NEAS, oKda = 3 * NEAS + NEAS * NEAS,
                         oStep = oKda + 24 * NEAS, hist_doubles = oStep + 1;
    static constexpr int epb = 4, threads = 32 * epb;
    // Gauss-point record: F(9) N_XYZ(24) cmat(36) S(6) fac m(6) G(36) H(36) T^T S(6) pad
    static constexpr int gF = 0, gN = 9, gC = 33, gS = 69, gFac = 75, gM = 76, gG = 82, gH = 118, gTS = 154, gpd = 160;
    // per element: X(24) u(24) T(36) detJ0 pad alpha(NEAS) s(NEAS) | records | W = [Kaa | I] (NEAS x 2 NEAS) | Kda (24 x NEAS)
    static constexpr int sX = 0, sU = 24, sT = 48, sDet = 84, sAlpha = 86, sS = sAlpha + NEAS, sGp = (sS + NEAS + 1) / 2 * 2,
                         sW = sGp + 8 * gpd, sKda = sW + 2 * NEAS * NEAS, ele_doubles = (sKda + 24 * NEAS + 1) / 2 * 2;
    static constexpr size_t smem_bytes = sizeof(double) * ele_doubles * epb;
  };

  // Entry (r, c) of the strain-like Voigt form of E -> A E A^T (rows r = (a, b), columns c = (p, q) in the order xx yy zz xy yz
  // xz; shear rows and columns carry the doubled shear strains), A row-major 3 x 3.  The reference builds this matrix from the
  // centroid Jacobian and inverts it with LAPACK (evaluate_voigt_transformation_matrix, calc_lib.hpp:1177-1244); the inverse of
  // the map with A is the map with A^-1, so T0^-T is this matrix of J0^-1 and no 6 x 6 system is solved.
  __device__ __forceinline__ double voigt_transformation_entry(const double* Am, int r, int c)
  {
    const int a = r < 3 ? r : (r == 4 ? 1 : 0), b = r < 3 ? r : (r == 3 ? 1 : 2);
    const int p = c < 3 ? c : (c == 4 ? 1 : 0), q = c < 3 ? c : (c == 3 ? 1 : 2);
    if (c < 3) return (r < 3 ? 1.0 : 2.0) * Am[3 * a + p] * Am[3 * b + p];
    if (r < 3) return Am[3 * a + p] * Am[3 * a + q];
    return Am[3 * a + p] * Am[3 * b + q] + Am[3 * b + p] * Am[3 * a + q];
  }

  // inverse of the centroid Jacobian iJ0 (row-major; J0[3 d + i] = dX_i / dxi_d) and det J0; dN0 [8][3], Xs [8][3]
  template <int NEN>
  __device__ __forceinline__ double eas_centroid_jacobian_n(const double* __restrict__ dN0, const double* Xs, double* iJ0)
  {
    double J[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int a = 0; a < NEN; ++a)
      for (int d = 0; d < 3; ++d)
        for (int i = 0; i < 3; ++i) J[3 * d + i] += __ldg(dN0 + 3 * a + d) * Xs[3 * a + i];
    return inv3(J, iJ0);
  }
  __device__ __forceinline__ double eas_centroid_jacobian(const double* __restrict__ dN0, const double* Xs, double* iJ0)
  {
    return eas_centroid_jacobian_n<8>(dN0, Xs, iJ0);
  }

  // T0^-T (row-major 6 x 6) and det J0, one thread
  __device__ inline double eas_centroid_transformation(const double* __restrict__ dN0, const double* Xs, double* Tm)
  {
    double iJ0[9];
    const double detJ0 = eas_centroid_jacobian(dN0, Xs, iJ0);
    for (int r = 0; r < 6; ++r)
      for (int c = 0; c < 6; ++c) Tm[6 * r + c] = voigt_transformation_entry(iJ0, r, c);
    return detJ0;
  }

  // m[6] = scale (r, s, t, rs, rt, st) of Gauss point gp
  __device__ __forceinline__ void eas_monomials(int gp, double scale, double* m)
  {
    double xi[3], w;
    gauss_point<HEX8>(gp, xi, &w);
    m[0] = scale * xi[0];
    m[1] = scale * xi[1];
    m[2] = scale * xi[2];
    m[3] = scale * xi[0] * xi[1];
    m[4] = scale * xi[0] * xi[2];
    m[5] = scale * xi[1] * xi[2];
  }

  // enhanced strains Mtilde alpha (strain-like Voigt) added to the right Cauchy-Green tensor C (symmetric-6, true components)
  template <int NEAS>
  __device__ __forceinline__ void eas_enhance_cauchy_green(const double* Tm, const double* m, const double* alpha, double* C)
  {
    double ea[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < NEAS; ++k)
    {
      const double ma = m[eas_mono<NEAS>(k)] * alpha[k];
      for (int v = 0; v < 6; ++v) ea[v] += Tm[6 * v + eas_row<NEAS>(k)] * ma;
    }
    for (int v = 0; v < 3; ++v) C[v] += 2.0 * ea[v];
    for (int v = 3; v < 6; ++v) C[v] += ea[v];
  }

  // The material law as a call of its own: inlined into k_hex8_eas, nvcc 12.9 -O3 produces a wrong tangent (cmat(1,1) = 329 instead
  // of 13.5 for Neo-Hooke at F ~ I; S is right; the same source inlined into k_generic is right) - found with the parity tests.
  static __device__ __noinline__ void eas_material(const MatParams* mp, const double* C, double* S, double* cmat, const double* fiber)
  {
    material_evaluate(*mp, C, S, cmat, fiber);
  }

  // LINEAR: KINEM linear (evaluate_eas_kinematics, calc_eas_helpers.hpp:617-632): B-operator at F = I, strain = sym grad u + Mtilde
  // alpha; the geometric stiffness term stays, as EASFormulation::add_stiffness_matrix adds it for both kinematics (calc_eas.hpp:150-163)
  template <int NEAS, bool WANT_K, bool LINEAR = false>
  __global__ void __launch_bounds__(EasCfg<NEAS>::threads) k_hex8_eas(const EvalArgs A, const unsigned char* __restrict__ ele_bad)
  {
    using Cfg = EasCfg<NEAS>;
    constexpr int N2 = 2 * NEAS;
    extern __shared__ double smem[];
    const int lane = threadIdx.x & 31, el = threadIdx.x >> 5;
    const int e = blockIdx.x * Cfg::epb + el;
    if (e >= A.n_ele) return;  // the warps are independent: only __syncwarp below
    double* sm = smem + (size_t)el * Cfg::ele_doubles;
    double *Xs = sm + Cfg::sX, *us = sm + Cfg::sU, *Tm = sm + Cfg::sT, *alpha = sm + Cfg::sAlpha, *sv = sm + Cfg::sS,
           *gpdat = sm + Cfg::sGp, *W = sm + Cfg::sW, *Kda = sm + Cfg::sKda;
    double* hist = A.eas + (size_t)e * Cfg::hist_doubles;
    const int a = lane / 3, i = lane % 3;  // lanes 0..23: DOF d = lane = 3 a + i
    int node = 0;
    if (lane < 24)
    {
      node = __ldg(A.conn + (size_t)e * 8 + a);
      Xs[lane] = __ldg(A.X + 3 * (size_t)node + i);
      us[lane] = __ldg(A.u + 3 * (size_t)node + i);
    }
    if (lane < NEAS) alpha[lane] = hist[Cfg::oAlpha + lane];
    __syncwarp();
    {
      // T0^-T: every lane inverts the centroid Jacobian and takes one or two of the 36 entries
      double iJ0[9];
      const double detJ0 = eas_centroid_jacobian(A.dNtab + (size_t)16 * 24, Xs, iJ0);
      Tm[lane] = voigt_transformation_entry(iJ0, lane / 6, lane % 6);
      if (lane < 4) Tm[32 + lane] = voigt_transformation_entry(iJ0, 5, 2 + lane);
      if (lane == 0) sm[Cfg::sDet] = detJ0;
    }
    __syncwarp();

    // ---- Gauss-point records (lanes 0..7) ----------------------------------------------------------
    int bad = (ele_bad && ele_bad[e]) ? 1 : 0;
    if (lane < 8)
    {
      const int gp = lane;
      double* R = gpdat + (size_t)gp * Cfg::gpd;
      const double detJ = jacobian_mapping<8>(A.dNtab + (size_t)gp * 24, Xs, R + Cfg::gN);
      double F[9];
      deformation_gradient<HEX8>(Xs, us, R + Cfg::gN, F);
      const double detF = LINEAR ? 1.0 : det3(F);
      double C[6] = {F[0] * F[0] + F[3] * F[3] + F[6] * F[6], F[1] * F[1] + F[4] * F[4] + F[7] * F[7],
          F[2] * F[2] + F[5] * F[5] + F[8] * F[8], F[0] * F[1] + F[3] * F[4] + F[6] * F[7],
          F[1] * F[2] + F[4] * F[5] + F[7] * F[8], F[0] * F[2] + F[3] * F[5] + F[6] * F[8]};
      if (LINEAR)
      {
        // the material law sees C = 2 sym(grad u) + I, the B-operator F = I
        const double H[9] = {F[0] - 1.0, F[1], F[2], F[3], F[4] - 1.0, F[5], F[6], F[7], F[8] - 1.0};
        C[0] = 2.0 * H[0] + 1.0;
        C[1] = 2.0 * H[4] + 1.0;
        C[2] = 2.0 * H[8] + 1.0;
        C[3] = H[1] + H[3];
        C[4] = H[5] + H[7];
        C[5] = H[2] + H[6];
        for (int k = 0; k < 9; ++k) F[k] = (k % 4 == 0) ? 1.0 : 0.0;
      }
      double m[6];
      eas_monomials(gp, sm[Cfg::sDet] / detJ, m);
      eas_enhance_cauchy_green<NEAS>(Tm, m, alpha, C);
      double S[6], cmat[36];
      eas_material(&A.mat, C, S, cmat, A.fiber ? A.fiber + (size_t)e * A.fiber_ele + gp * A.fiber_gp : nullptr);
      if (!(detJ > 0.0) || detF == 0.0 || !isfinite(detF) || !isfinite(S[0] + S[1] + S[2] + S[3] + S[4] + S[5])) bad = 1;
      for (int k = 0; k < 9; ++k) R[Cfg::gF + k] = F[k];
      for (int k = 0; k < 36; ++k) R[Cfg::gC + k] = cmat[k];
      for (int k = 0; k < 6; ++k) R[Cfg::gS + k] = S[k];
      R[Cfg::gFac] = detJ * __ldg(A.gpw + gp);
      for (int k = 0; k < 6; ++k) R[Cfg::gM + k] = m[k];
    }
    __syncwarp();
    // H = cmat T, then G = T^T H and T^T S of the eight records, entries over the lanes
    for (int idx = lane; idx < 8 * 36; idx += 32)
    {
      double* R = gpdat + (size_t)(idx / 36) * Cfg::gpd;
      const int v = (idx % 36) / 6, c = idx % 6;
      double h = 0.0;
      for (int w = 0; w < 6; ++w) h += R[Cfg::gC + 6 * v + w] * Tm[6 * w + c];
      R[Cfg::gH + 6 * v + c] = h;
    }
    __syncwarp();
    for (int idx = lane; idx < 8 * 42; idx += 32)
    {
      double* R = gpdat + (size_t)(idx / 42) * Cfg::gpd;
      const int r = (idx % 42) / 7, c = idx % 7;
      double g = 0.0;
      if (c < 6)
      {
        for (int v = 0; v < 6; ++v) g += Tm[6 * v + r] * R[Cfg::gH + 6 * v + c];
        R[Cfg::gG + 6 * r + c] = g;
      }
      else
      {
        for (int v = 0; v < 6; ++v) g += Tm[6 * v + r] * R[Cfg::gS + v];
        R[Cfg::gTS + r] = g;
      }
    }
    __syncwarp();
    if (__any_sync(0xffffffffu, bad))
    {
      if (lane == 0) report_error(A.err, __ldg(A.ele_gid + e));
      return;
    }

    // ---- EAS integrals ---------------------------------------------------------------------------------
    for (int idx = lane; idx < NEAS * NEAS; idx += 32)
    {
      const int k = idx / NEAS, l = idx % NEAS;
      const int rk = eas_row<NEAS>(k), rl = eas_row<NEAS>(l), mk = eas_mono<NEAS>(k), ml = eas_mono<NEAS>(l);
      double v = 0.0;
      for (int gp = 0; gp < 8; ++gp)
      {
        const double* R = gpdat + (size_t)gp * Cfg::gpd;
        v += R[Cfg::gFac] * R[Cfg::gM + mk] * R[Cfg::gM + ml] * R[Cfg::gG + 6 * rk + rl];
      }
      W[k * N2 + l] = v;
      W[k * N2 + NEAS + l] = k == l ? 1.0 : 0.0;
    }
    if (lane < NEAS)
    {
      const int rk = eas_row<NEAS>(lane), mk = eas_mono<NEAS>(lane);
      double v = 0.0;
      for (int gp = 0; gp < 8; ++gp)
      {
        const double* R = gpdat + (size_t)gp * Cfg::gpd;
        v += R[Cfg::gFac] * R[Cfg::gM + mk] * R[Cfg::gTS + rk];
      }
      sv[lane] = v;
    }
    double fd = 0.0;  // internal force of DOF d
    if (lane < 24)
    {
      double acc[NEAS];
#pragma unroll
      for (int l = 0; l < NEAS; ++l) acc[l] = 0.0;
      for (int gp = 0; gp < 8; ++gp)
      {
        const double* R = gpdat + (size_t)gp * Cfg::gpd;
        const double* F = R + Cfg::gF + 3 * i;
        const double* Nx = R + Cfg::gN + 3 * a;
        const double Bc[6] = {F[0] * Nx[0], F[1] * Nx[1], F[2] * Nx[2], F[0] * Nx[1] + F[1] * Nx[0], F[1] * Nx[2] + F[2] * Nx[1],
            F[2] * Nx[0] + F[0] * Nx[2]};
        const double fac = R[Cfg::gFac];
        double P[6], fm[6], bs = 0.0;
        for (int c = 0; c < 6; ++c)
        {
          double p = 0.0;
          for (int v = 0; v < 6; ++v) p += Bc[v] * R[Cfg::gH + 6 * v + c];
          P[c] = p;
          fm[c] = fac * R[Cfg::gM + c];
          bs += Bc[c] * R[Cfg::gS + c];
        }
        fd += fac * bs;
#pragma unroll
        for (int l = 0; l < NEAS; ++l) acc[l] += fm[eas_mono<NEAS>(l)] * P[eas_row<NEAS>(l)];
      }
#pragma unroll
      for (int l = 0; l < NEAS; ++l) Kda[lane * NEAS + l] = acc[l];
    }
    __syncwarp();

    // ---- Kaa^-1: Gauss-Jordan on W = [Kaa | I], rows over the lanes ----------------------------------------
    for (int k = 0; k < NEAS; ++k)
    {
      double best = (lane >= k && lane < NEAS) ? fabs(W[lane * N2 + k]) : -1.0;
      int piv = lane;
      for (int off = 16; off > 0; off >>= 1)
      {
        const double ob = __shfl_xor_sync(0xffffffffu, best, off);
        const int op = __shfl_xor_sync(0xffffffffu, piv, off);
        if (ob > best || (ob == best && op < piv))
        {
          best = ob;
          piv = op;
        }
      }
      if (piv != k)
        for (int c = lane; c < N2; c += 32)
        {
          const double t = W[k * N2 + c];
          W[k * N2 + c] = W[piv * N2 + c];
          W[piv * N2 + c] = t;
        }
      __syncwarp();
      const double ip = 1.0 / W[k * N2 + k];
      __syncwarp();
      for (int c = lane; c < N2; c += 32) W[k * N2 + c] *= ip;
      __syncwarp();
      if (lane < NEAS && lane != k)
      {
        const double m = W[lane * N2 + k];
        for (int c = 0; c < N2; ++c) W[lane * N2 + c] -= m * W[k * N2 + c];
      }
      __syncwarp();
    }

    // ---- condensation; history for the next recover ------------------------------------------------------------
    double mk[NEAS];
    if (lane < 24)
    {
#pragma unroll
      for (int l = 0; l < NEAS; ++l) mk[l] = 0.0;
      for (int k = 0; k < NEAS; ++k)
      {
        const double kd = Kda[lane * NEAS + k];
#pragma unroll
        for (int l = 0; l < NEAS; ++l) mk[l] -= kd * W[k * N2 + NEAS + l];
      }
#pragma unroll
      for (int l = 0; l < NEAS; ++l) fd += mk[l] * sv[l];
    }
    if (lane < NEAS) hist[Cfg::oS + lane] = sv[lane];
    for (int idx = lane; idx < NEAS * NEAS; idx += 32) hist[Cfg::oInvKaa + idx] = W[(idx % NEAS) * N2 + NEAS + idx / NEAS];
    for (int idx = lane; idx < 24 * NEAS; idx += 32) hist[Cfg::oKda + idx] = Kda[(idx % 24) * NEAS + idx / 24];

    const bool owned = lane < 24 && node < A.n_row_nodes;
    if (lane < 24)
    {
      if (A.fe_out) A.fe_out[(size_t)e * 24 + lane] = fd;
      if (A.f && owned) red_add(A.f + 3 * (size_t)node + i, fd);
    }
    if (!WANT_K) return;
    __syncwarp();

    // ---- K row d: B^T cmat B + K_geo - Kda Kaa^-1 Kda^T ----------------------------------------------------------
    if (lane < 24)
    {
      double acc[24];
#pragma unroll
      for (int c = 0; c < 24; ++c) acc[c] = 0.0;
      for (int gp = 0; gp < 8; ++gp)
      {
        const double* R = gpdat + (size_t)gp * Cfg::gpd;
        const double* Fm = R + Cfg::gF;
        const double* F = Fm + 3 * i;
        const double* Nx = R + Cfg::gN;
        const double n0 = Nx[3 * a], n1 = Nx[3 * a + 1], n2 = Nx[3 * a + 2];
        const double Bc[6] = {F[0] * n0, F[1] * n1, F[2] * n2, F[0] * n1 + F[1] * n0, F[1] * n2 + F[2] * n1, F[2] * n0 + F[0] * n2};
        const double fac = R[Cfg::gFac];
        double CB[6];
        for (int v = 0; v < 6; ++v)
        {
          double s = 0.0;
          for (int w = 0; w < 6; ++w) s += R[Cfg::gC + 6 * v + w] * Bc[w];
          CB[v] = fac * s;
        }
        const double* S = R + Cfg::gS;
        const double SN0 = fac * (S[0] * n0 + S[3] * n1 + S[5] * n2);
        const double SN1 = fac * (S[3] * n0 + S[1] * n1 + S[4] * n2);
        const double SN2 = fac * (S[5] * n0 + S[4] * n1 + S[2] * n2);
#pragma unroll
        for (int b = 0; b < 8; ++b)
        {
          const double m0 = Nx[3 * b], m1 = Nx[3 * b + 1], m2 = Nx[3 * b + 2];
          double Bb[18];
          b_block(Fm, m0, m1, m2, Bb);
          const double geo = SN0 * m0 + SN1 * m1 + SN2 * m2;
#pragma unroll
          for (int j = 0; j < 3; ++j)
          {
            double s = acc[3 * b + j];
#pragma unroll
            for (int v = 0; v < 6; ++v) s += CB[v] * Bb[3 * v + j];
            acc[3 * b + j] = s + (i == j ? geo : 0.0);
          }
        }
      }
#pragma unroll
      for (int c = 0; c < 24; ++c)
      {
        double s = acc[c];
#pragma unroll
        for (int k = 0; k < NEAS; ++k) s += mk[k] * Kda[c * NEAS + k];
        acc[c] = s;
      }
      if (A.Ke_out)
      {
        double* Ke = A.Ke_out + (size_t)e * 576;
        for (int c = 0; c < 24; ++c) Ke[lane + 24 * c] = acc[c];
      }
      const int64_t rowbase = (owned && A.node_rowbase) ? __ldg(A.node_rowbase + node) : -1;
      if (A.K && rowbase >= 0 && !(A.dbc && A.dbc[3 * node + i]))
      {
        const int rowlen = __ldg(A.node_rowlen + node);
        const unsigned char* sl = A.slots + (size_t)e * 64 + a * 8;
        double* row = A.K + rowbase + (int64_t)i * rowlen;
#pragma unroll
        for (int b = 0; b < 8; ++b)
        {
          double* p = row + 3 * (int)sl[b];
          red_add(p + 0, acc[3 * b + 0]);
          red_add(p + 1, acc[3 * b + 1]);
          red_add(p + 2, acc[3 * b + 2]);
        }
      }
    }
  }

  // struct_calc_recover (default step): alpha_inc = -Kaa^-1 (s + Kda^T du), alpha += step_length alpha_inc.  One warp per element.
  template <int NEAS>
  __global__ void k_hex8_eas_recover(const int* __restrict__ conn, int n_ele, const double* __restrict__ du, double step_length,
      double* __restrict__ hist_all)
  {
    using Cfg = EasCfg<NEAS>;
    __shared__ double s_tmp[8][NEAS], s_du[8][24];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int e = blockIdx.x * 8 + w;
    if (e >= n_ele) return;
    double* hist = hist_all + (size_t)e * Cfg::hist_doubles;
    if (lane < 24) s_du[w][lane] = __ldg(du + 3 * (size_t)__ldg(conn + (size_t)e * 8 + lane / 3) + lane % 3);
    __syncwarp();
    if (lane < NEAS)
    {
      double v = hist[Cfg::oS + lane];
      for (int d = 0; d < 24; ++d) v += hist[Cfg::oKda + d + 24 * lane] * s_du[w][d];
      s_tmp[w][lane] = v;
    }
    __syncwarp();
    if (lane < NEAS)
    {
      double v = 0.0;
      for (int l = 0; l < NEAS; ++l) v += hist[Cfg::oInvKaa + lane + NEAS * l] * s_tmp[w][l];
      hist[Cfg::oAlphaInc + lane] = -v;
      hist[Cfg::oAlpha + lane] += step_length * (-v);
    }
    if (lane == 0) hist[Cfg::oStep] = step_length;
  }

  // new elements: everything zero, old_step_length = 1 (EASHistoryData, calc_eas.hpp:29-37)
  static __global__ void k_eas_history_init(long long n_ele, int hist_doubles, double* __restrict__ hist)
  {
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= n_ele * hist_doubles) return;
    hist[tid] = (tid % hist_doubles == hist_doubles - 1) ? 1.0 : 0.0;
  }
}  // namespace fc
