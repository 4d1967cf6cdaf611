// Solid shells: assumed natural strains (TECH shell_ans) on hex8 and wedge6, optionally with the EAS modes of the solid shell
// (TECH shell_eas_ans: eas_sosh8, 7 modes, on hex8; eas_sosw6, 1 mode, on wedge6).  The element's t direction is the thickness.
//   sampling points and their data, local B-operator, local strains, geometric stiffness
//                                            src/solid_3D_ele/4C_solid_3D_ele_calc_shell_ans_lib.hpp:24-47, 61-330, 332-504, 506-560
//   ShellANSFormulation                      src/solid_3D_ele/4C_solid_3D_ele_calc_shell_ans.hpp:52-196
//   ShellEASANSFormulation                   src/solid_3D_ele/4C_solid_3D_ele_calc_shell_eas_ans.hpp:48-345
// One warp per element, like k_hex8_eas (fc_kernels_hex8eas.cuh), whose condensation it repeats.  Strains and B-operator live in
// the natural coordinates (rr, ss, tt, rs, st, rt): the in-plane rows at the Gauss point, the rows tt / st / rt interpolated from
// the sampling points; T^-T of the Gauss point's Jacobian (the Voigt transformation of J^-1, no 6 x 6 solve) takes both to the
// global frame.  The geometric stiffness contracts T^-T^T S with the natural-coordinate products of the shape function derivatives.
#pragma once
#include "fc_kernels_hex8eas.cuh"

namespace fc
{
  template <int CT, int NEAS>
  struct ShellCfg
  {
    static constexpr int nen = Cell<CT>::nen, ngp = Cell<CT>::ngp, nd = 3 * nen;
    static constexpr int nsp = CT == HEX8 ? 8 : 5;  // sampling points
    static constexpr int ne = NEAS > 0 ? NEAS : 1;
    // element history (EAS only), the order of the pack functions: alpha_inc | alpha | s | Kaa^-1 | Kda [nd x n] | old_step_length
    static constexpr int oAlphaInc = 0, oAlpha = NEAS, oS = 2 * NEAS, oInvKaa = 3 * NEAS, oKda = 3 * NEAS + NEAS * NEAS,
                         oStep = oKda + nd * NEAS, hist_doubles = NEAS > 0 ? oStep + 1 : 0;
    static constexpr int epb = 2, threads = 32 * epb;
    // Gauss-point record: dN(3 nen) T^-T(36) B(6 nd) cmat(36) S(6) T^-T^T S(6) fac w(8) | EAS: m(6) G(36) H(36) T0^T S(6)
    static constexpr int gD = 0, gT = gD + 3 * nen, gB = gT + 36, gC = gB + 6 * nd, gS = gC + 36, gSl = gS + 6, gFac = gSl + 6, gW = gFac + 1,
                         gM = gW + 8, gG = gM + 6, gH = gG + 36, gTS = gH + 36, gpd = ((NEAS > 0 ? gTS + 6 : gM) + 1) / 2 * 2;
    // per element: X u | sampling points: dN(3 nen) Jref(9) Jcur(9) bop(3 nd) each | T0^-T(36) detJ0 pad alpha s | records | W | Kda
    static constexpr int spd = 3 * nen + 18 + 3 * nd;
    static constexpr int sX = 0, sU = nd, sSp = 2 * nd, sT0 = sSp + nsp * spd, sDet = sT0 + 36, sAlpha = sDet + 2, sS = sAlpha + ne,
                         sGp = (sS + ne + 1) / 2 * 2, sW = sGp + ngp * gpd, sKda = sW + 2 * ne * ne,
                         ele_doubles = (sKda + nd * ne + 1) / 2 * 2;
    static constexpr size_t smem_bytes = sizeof(double) * ele_doubles * epb;
  };

  // row of the entry of EAS mode k and its monomial (0 r, 1 s, 2 t, 3 rs, 4 rt, 5 st): hex8 the table of eas_sosh8, wedge6 one
  // mode, t on the thickness strain (4C_solid_3D_ele_calc_eas_helpers.hpp:350-367)
  template <int CT, int NEAS>
  __host__ __device__ constexpr int shell_row(int k) { return CT == HEX8 ? eas_row<7>(k) : 2; }
  template <int CT, int NEAS>
  __host__ __device__ constexpr int shell_mono(int k) { return CT == HEX8 ? eas_mono<7>(k) : 2; }

  // parameter coordinates of sampling point k (4C_solid_3D_ele_calc_shell_ans_lib.hpp:24-47)
  template <int CT>
  __device__ __forceinline__ void shell_sampling_point(int k, double* xi)
  {
    if (CT == HEX8)
    {
      // A B C D: edge mids of the mid-surface; E F G H: its corners
      const int r[8] = {0, 1, 0, -1, -1, 1, 1, -1}, s[8] = {-1, 0, 1, 0, -1, -1, 1, 1};
      xi[0] = r[k];
      xi[1] = s[k];
    }
    else
    {
      const double r[5] = {0.5, 0.0, 0.0, 1.0, 0.0}, s[5] = {0.0, 0.5, 0.0, 0.0, 1.0};
      xi[0] = r[k];
      xi[1] = s[k];
    }
    xi[2] = 0.0;
  }

  // interpolation weights of the sampling points at (r, s): w[0..3] the four (hex8) / three (wedge6) points of the tt row,
  // w[4..5] the st row, w[6..7] the rt row; spi[] says which sampling point each weight belongs to
  template <int CT>
  __device__ __forceinline__ void shell_weights(double r, double s, double* w)
  {
    if (CT == HEX8)
    {
      w[0] = 0.25 * (1 - r) * (1 - s);  // E
      w[1] = 0.25 * (1 + r) * (1 - s);  // F
      w[2] = 0.25 * (1 + r) * (1 + s);  // G
      w[3] = 0.25 * (1 - r) * (1 + s);  // H
      w[4] = 0.5 * (1 + r);             // B
      w[5] = 0.5 * (1 - r);             // D
      w[6] = 0.5 * (1 - s);             // A
      w[7] = 0.5 * (1 + s);             // C
    }
    else
    {
      w[0] = 1 - r - s;  // C
      w[1] = r;          // D
      w[2] = s;          // E
      w[3] = 0.0;
      w[4] = r;  // B
      w[5] = 0.0;
      w[6] = s;  // A
      w[7] = 0.0;
    }
  }
  // sampling point of weight slot q (slots without a point repeat a valid index with weight 0)
  template <int CT>
  __host__ __device__ constexpr int shell_slot_point(int q)
  {
    return CT == HEX8 ? (q < 4 ? 4 + q : (q == 4 ? 1 : (q == 5 ? 3 : (q == 6 ? 0 : 2)))) : (q < 3 ? 2 + q : (q == 3 ? 2 : (q < 6 ? 1 : 0)));
  }

  // natural-coordinate Green-Lagrange strains (strain-like) at a Gauss point: in-plane rows from the Jacobians J (reference) and
  // jc (current) of the point, rows tt / st / rt from the metric differences of the sampling points; sp = sampling point records
  template <int CT>
  __device__ __forceinline__ void shell_local_strain(const double* J, const double* jc, const double* w, const double* sp, int spd, int oJr,
      double* El)
  {
    double g00 = 0, G00 = 0, g11 = 0, G11 = 0, g01 = 0, G01 = 0;
    for (int i = 0; i < 3; ++i)
    {
      g00 += jc[i] * jc[i];
      G00 += J[i] * J[i];
      g11 += jc[3 + i] * jc[3 + i];
      G11 += J[3 + i] * J[3 + i];
      g01 += jc[i] * jc[3 + i];
      G01 += J[i] * J[3 + i];
    }
    El[0] = 0.5 * (g00 - G00);
    El[1] = 0.5 * (g11 - G11);
    El[3] = g01 - G01;
    double e[3] = {0.0, 0.0, 0.0};
    for (int q = 0; q < 8; ++q)
    {
      const double* Jr = sp + (size_t)shell_slot_point<CT>(q) * spd + oJr;  // Jref(9) then Jcur(9), row-major [3 d + i]
      const double* Jc = Jr + 9;
      const int d1 = q < 4 ? 2 : (q < 6 ? 1 : 0);
      double g = 0.0, G = 0.0;
      for (int i = 0; i < 3; ++i)
      {
        g += Jc[3 * d1 + i] * Jc[6 + i];
        G += Jr[3 * d1 + i] * Jr[6 + i];
      }
      e[q < 4 ? 0 : (q < 6 ? 1 : 2)] += w[q] * (g - G);
    }
    El[2] = 0.5 * e[0];
    El[4] = e[1];
    El[5] = e[2];
  }

  // like eas_material (fc_kernels_hex8eas.cuh): the material law as a call of its own
  static __device__ __noinline__ void shell_material(const MatParams* mp, const double* C, double* S, double* cmat, const double* fiber)
  {
    material_evaluate(*mp, C, S, cmat, fiber);
  }

  // sampling point data of one element: lane k < nsp fills record k (dN, Jref, Jcur, the rows tt / st / rt of the local B-operator)
  template <int CT>
  __device__ __forceinline__ void shell_sampling_data(int k, const double* Xs, const double* us, double* rec)
  {
    constexpr int NEN = Cell<CT>::nen;
    double xi[3];
    shell_sampling_point<CT>(k, xi);
    double* dN = rec;
    shape_deriv1<CT>(xi, dN);
    double* Jr = rec + 3 * NEN;
    double* Jc = Jr + 9;
    for (int q = 0; q < 9; ++q) Jr[q] = 0.0;
    for (int a = 0; a < NEN; ++a)
      for (int d = 0; d < 3; ++d)
        for (int i = 0; i < 3; ++i) Jr[3 * d + i] += dN[3 * a + d] * Xs[3 * a + i];
    for (int q = 0; q < 9; ++q) Jc[q] = Jr[q];
    for (int a = 0; a < NEN; ++a)
      for (int d = 0; d < 3; ++d)
        for (int i = 0; i < 3; ++i) Jc[3 * d + i] += dN[3 * a + d] * us[3 * a + i];
    double* bop = Jc + 9;  // [3][nd]
    for (int a = 0; a < NEN; ++a)
      for (int i = 0; i < 3; ++i)
      {
        const double d0 = dN[3 * a], d1 = dN[3 * a + 1], d2 = dN[3 * a + 2];
        bop[0 * 3 * NEN + 3 * a + i] = d2 * Jc[6 + i];
        bop[1 * 3 * NEN + 3 * a + i] = d1 * Jc[6 + i] + d2 * Jc[3 + i];
        bop[2 * 3 * NEN + 3 * a + i] = d0 * Jc[6 + i] + d2 * Jc[i];
      }
  }

  template <int CT, int NEAS, bool WANT_K>
  __global__ void __launch_bounds__(ShellCfg<CT, NEAS>::threads) k_shell(const EvalArgs A, const unsigned char* __restrict__ ele_bad)
  {
    using Cfg = ShellCfg<CT, NEAS>;
    constexpr int NEN = Cfg::nen, NGP = Cfg::ngp, ND = Cfg::nd, NSP = Cfg::nsp, NE = Cfg::ne, N2 = 2 * NE;
    extern __shared__ double smem[];
    const int lane = threadIdx.x & 31, el = threadIdx.x >> 5;
    const int e = blockIdx.x * Cfg::epb + el;
    if (e >= A.n_ele) return;  // independent warps: only __syncwarp below
    double* sm = smem + (size_t)el * Cfg::ele_doubles;
    double *Xs = sm + Cfg::sX, *us = sm + Cfg::sU, *sp = sm + Cfg::sSp, *T0 = sm + Cfg::sT0, *alpha = sm + Cfg::sAlpha, *sv = sm + Cfg::sS,
           *gpdat = sm + Cfg::sGp, *W = sm + Cfg::sW, *Kda = sm + Cfg::sKda;
    double* hist = NEAS > 0 ? A.eas + (size_t)e * Cfg::hist_doubles : nullptr;
    const int a = lane / 3, i = lane % 3;  // lanes < nd: DOF d = lane = 3 a + i
    int node = 0;
    if (lane < ND)
    {
      node = __ldg(A.conn + (size_t)e * NEN + a);
      Xs[lane] = __ldg(A.X + 3 * (size_t)node + i);
      us[lane] = __ldg(A.u + 3 * (size_t)node + i);
    }
    if (NEAS > 0 && lane < NEAS) alpha[lane] = hist[Cfg::oAlpha + lane];
    __syncwarp();
    if (lane < NSP) shell_sampling_data<CT>(lane, Xs, us, sp + (size_t)lane * Cfg::spd);
    if (NEAS > 0)
    {
      double iJ0[9];
      const double detJ0 = eas_centroid_jacobian_n<NEN>(A.dNtab + (size_t)(NGP + NEN) * ND, Xs, iJ0);
      T0[lane] = voigt_transformation_entry(iJ0, lane / 6, lane % 6);
      if (lane < 4) T0[32 + lane] = voigt_transformation_entry(iJ0, 5, 2 + lane);
      if (lane == 0) sm[Cfg::sDet] = detJ0;
    }
    __syncwarp();

    // ---- Gauss-point records (lanes < ngp) ------------------------------------------------------------
    int bad = (ele_bad && ele_bad[e]) ? 1 : 0;
    if (lane < NGP)
    {
      const int gp = lane;
      double* R = gpdat + (size_t)gp * Cfg::gpd;
      double xi[3], wq;
      gauss_point<CT>(gp, xi, &wq);
      const double* dNg = A.dNtab + (size_t)gp * ND;
      double J[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, jc[9], iJ[9];
      for (int b = 0; b < NEN; ++b)
      {
        const double d0 = __ldg(dNg + 3 * b), d1 = __ldg(dNg + 3 * b + 1), d2 = __ldg(dNg + 3 * b + 2);
        R[Cfg::gD + 3 * b] = d0;
        R[Cfg::gD + 3 * b + 1] = d1;
        R[Cfg::gD + 3 * b + 2] = d2;
        for (int c = 0; c < 3; ++c)
        {
          J[c] += d0 * Xs[3 * b + c];
          J[3 + c] += d1 * Xs[3 * b + c];
          J[6 + c] += d2 * Xs[3 * b + c];
        }
      }
      for (int q = 0; q < 9; ++q) jc[q] = J[q];
      for (int b = 0; b < NEN; ++b)
        for (int d = 0; d < 3; ++d)
          for (int c = 0; c < 3; ++c) jc[3 * d + c] += R[Cfg::gD + 3 * b + d] * us[3 * b + c];
      const double detJ = inv3(J, iJ);
      double Tm[36];
      for (int r = 0; r < 6; ++r)
        for (int c = 0; c < 6; ++c) Tm[6 * r + c] = voigt_transformation_entry(iJ, r, c);
      for (int q = 0; q < 36; ++q) R[Cfg::gT + q] = Tm[q];
      double w[8];
      shell_weights<CT>(xi[0], xi[1], w);
      for (int q = 0; q < 8; ++q) R[Cfg::gW + q] = w[q];
      // B = T^-T B_loc, column by column
      for (int c = 0; c < ND; ++c)
      {
        const int b = c / 3, k = c % 3;
        const double d0 = R[Cfg::gD + 3 * b], d1 = R[Cfg::gD + 3 * b + 1];
        double Bl[6];
        Bl[0] = d0 * jc[k];
        Bl[1] = d1 * jc[3 + k];
        Bl[3] = d0 * jc[3 + k] + d1 * jc[k];
        double b2 = 0.0, b4 = 0.0, b5 = 0.0;
        for (int q = 0; q < 4; ++q) b2 += w[q] * sp[(size_t)shell_slot_point<CT>(q) * Cfg::spd + 3 * NEN + 18 + c];
        for (int q = 4; q < 6; ++q) b4 += w[q] * sp[(size_t)shell_slot_point<CT>(q) * Cfg::spd + 3 * NEN + 18 + ND + c];
        for (int q = 6; q < 8; ++q) b5 += w[q] * sp[(size_t)shell_slot_point<CT>(q) * Cfg::spd + 3 * NEN + 18 + 2 * ND + c];
        Bl[2] = b2;
        Bl[4] = b4;
        Bl[5] = b5;
        for (int r = 0; r < 6; ++r)
        {
          double v = 0.0;
          for (int q = 0; q < 6; ++q) v += Tm[6 * r + q] * Bl[q];
          R[Cfg::gB + r * ND + c] = v;
        }
      }
      double El[6], E[6];
      shell_local_strain<CT>(J, jc, w, sp, Cfg::spd, 3 * NEN, El);
      for (int r = 0; r < 6; ++r)
      {
        double v = 0.0;
        for (int q = 0; q < 6; ++q) v += Tm[6 * r + q] * El[q];
        E[r] = v;
      }
      double C[6] = {2.0 * E[0] + 1.0, 2.0 * E[1] + 1.0, 2.0 * E[2] + 1.0, E[3], E[4], E[5]};
      if (NEAS > 0)
      {
        double m[6];
        const double sc = sm[Cfg::sDet] / detJ;
        m[0] = sc * xi[0]; m[1] = sc * xi[1]; m[2] = sc * xi[2];
        m[3] = sc * xi[0] * xi[1]; m[4] = sc * xi[0] * xi[2]; m[5] = sc * xi[1] * xi[2];
        for (int q = 0; q < 6; ++q) R[Cfg::gM + q] = m[q];
        double ea[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int k = 0; k < NE; ++k)
        {
          const double ma = m[shell_mono<CT, NEAS>(k)] * alpha[k];
          for (int v = 0; v < 6; ++v) ea[v] += T0[6 * v + shell_row<CT, NEAS>(k)] * ma;
        }
        for (int v = 0; v < 3; ++v) C[v] += 2.0 * ea[v];
        for (int v = 3; v < 6; ++v) C[v] += ea[v];
      }
      double S[6], cmat[36];
      shell_material(&A.mat, C, S, cmat, A.fiber ? A.fiber + (size_t)e * A.fiber_ele + gp * A.fiber_gp : nullptr);
      if (!(detJ > 0.0) || !isfinite(S[0] + S[1] + S[2] + S[3] + S[4] + S[5])) bad = 1;
      for (int q = 0; q < 36; ++q) R[Cfg::gC + q] = cmat[q];
      for (int q = 0; q < 6; ++q) R[Cfg::gS + q] = S[q];
      for (int c = 0; c < 6; ++c)
      {
        double v = 0.0;
        for (int r = 0; r < 6; ++r) v += Tm[6 * r + c] * S[r];
        R[Cfg::gSl + c] = v;
      }
      R[Cfg::gFac] = detJ * __ldg(A.gpw + gp);
    }
    __syncwarp();
    if (__any_sync(0xffffffffu, bad))
    {
      if (lane == 0) report_error(A.err, __ldg(A.ele_gid + e));
      return;
    }

    // ---- EAS integrals: as in k_hex8_eas with G = T0^T cmat T0, H = cmat T0, T0^T S and the stored B ------------------------
    if (NEAS > 0)
    {
      for (int idx = lane; idx < NGP * 36; idx += 32)
      {
        double* R = gpdat + (size_t)(idx / 36) * Cfg::gpd;
        const int v = (idx % 36) / 6, c = idx % 6;
        double h = 0.0;
        for (int w = 0; w < 6; ++w) h += R[Cfg::gC + 6 * v + w] * T0[6 * w + c];
        R[Cfg::gH + 6 * v + c] = h;
      }
      __syncwarp();
      for (int idx = lane; idx < NGP * 42; idx += 32)
      {
        double* R = gpdat + (size_t)(idx / 42) * Cfg::gpd;
        const int r = (idx % 42) / 7, c = idx % 7;
        double g = 0.0;
        if (c < 6)
        {
          for (int v = 0; v < 6; ++v) g += T0[6 * v + r] * R[Cfg::gH + 6 * v + c];
          R[Cfg::gG + 6 * r + c] = g;
        }
        else
        {
          for (int v = 0; v < 6; ++v) g += T0[6 * v + r] * R[Cfg::gS + v];
          R[Cfg::gTS + r] = g;
        }
      }
      __syncwarp();
      for (int idx = lane; idx < NE * NE; idx += 32)
      {
        const int k = idx / NE, l = idx % NE;
        const int rk = shell_row<CT, NEAS>(k), rl = shell_row<CT, NEAS>(l), mk = shell_mono<CT, NEAS>(k), ml = shell_mono<CT, NEAS>(l);
        double v = 0.0;
        for (int gp = 0; gp < NGP; ++gp)
        {
          const double* R = gpdat + (size_t)gp * Cfg::gpd;
          v += R[Cfg::gFac] * R[Cfg::gM + mk] * R[Cfg::gM + ml] * R[Cfg::gG + 6 * rk + rl];
        }
        W[k * N2 + l] = v;
        W[k * N2 + NE + l] = k == l ? 1.0 : 0.0;
      }
      if (lane < NE)
      {
        const int rk = shell_row<CT, NEAS>(lane), mk = shell_mono<CT, NEAS>(lane);
        double v = 0.0;
        for (int gp = 0; gp < NGP; ++gp)
        {
          const double* R = gpdat + (size_t)gp * Cfg::gpd;
          v += R[Cfg::gFac] * R[Cfg::gM + mk] * R[Cfg::gTS + rk];
        }
        sv[lane] = v;
      }
    }
    double fd = 0.0;  // internal force of DOF d
    if (lane < ND)
    {
      double acc[NE];
#pragma unroll
      for (int l = 0; l < NE; ++l) acc[l] = 0.0;
      for (int gp = 0; gp < NGP; ++gp)
      {
        const double* R = gpdat + (size_t)gp * Cfg::gpd;
        double Bc[6], bs = 0.0;
        for (int v = 0; v < 6; ++v)
        {
          Bc[v] = R[Cfg::gB + v * ND + lane];
          bs += Bc[v] * R[Cfg::gS + v];
        }
        const double fac = R[Cfg::gFac];
        fd += fac * bs;
        if (NEAS > 0)
        {
          double P[6], fm[6];
          for (int c = 0; c < 6; ++c)
          {
            double p = 0.0;
            for (int v = 0; v < 6; ++v) p += Bc[v] * R[Cfg::gH + 6 * v + c];
            P[c] = p;
            fm[c] = fac * R[Cfg::gM + c];
          }
#pragma unroll
          for (int l = 0; l < NE; ++l) acc[l] += fm[shell_mono<CT, NEAS>(l)] * P[shell_row<CT, NEAS>(l)];
        }
      }
      if (NEAS > 0)
      {
#pragma unroll
        for (int l = 0; l < NE; ++l) Kda[lane * NE + l] = acc[l];
      }
    }
    __syncwarp();
    double mk[NE];
#pragma unroll
    for (int l = 0; l < NE; ++l) mk[l] = 0.0;
    if (NEAS > 0)
    {
      // Kaa^-1: Gauss-Jordan on W = [Kaa | I], rows over the lanes
      for (int k = 0; k < NE; ++k)
      {
        double best = (lane >= k && lane < NE) ? fabs(W[lane * N2 + k]) : -1.0;
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
        if (lane < NE && lane != k)
        {
          const double m = W[lane * N2 + k];
          for (int c = 0; c < N2; ++c) W[lane * N2 + c] -= m * W[k * N2 + c];
        }
        __syncwarp();
      }
      if (lane < ND)
      {
        for (int k = 0; k < NE; ++k)
        {
          const double kd = Kda[lane * NE + k];
#pragma unroll
          for (int l = 0; l < NE; ++l) mk[l] -= kd * W[k * N2 + NE + l];
        }
#pragma unroll
        for (int l = 0; l < NE; ++l) fd += mk[l] * sv[l];
      }
      if (lane < NE) hist[Cfg::oS + lane] = sv[lane];
      for (int idx = lane; idx < NE * NE; idx += 32) hist[Cfg::oInvKaa + idx] = W[(idx % NE) * N2 + NE + idx / NE];
      for (int idx = lane; idx < ND * NE; idx += 32) hist[Cfg::oKda + idx] = Kda[(idx % ND) * NE + idx / ND];
    }
    const bool owned = lane < ND && node < A.n_row_nodes;
    if (lane < ND)
    {
      if (A.fe_out) A.fe_out[(size_t)e * ND + lane] = fd;
      if (A.f && owned) red_add(A.f + 3 * (size_t)node + i, fd);
    }
    if (!WANT_K) return;
    __syncwarp();

    // ---- K row d: B^T cmat B + ANS geometric stiffness - Kda Kaa^-1 Kda^T ------------------------------------------
    if (lane < ND)
    {
      double acc[ND];
#pragma unroll
      for (int c = 0; c < ND; ++c) acc[c] = 0.0;
      for (int gp = 0; gp < NGP; ++gp)
      {
        const double* R = gpdat + (size_t)gp * Cfg::gpd;
        const double fac = R[Cfg::gFac];
        double CB[6];
        for (int v = 0; v < 6; ++v)
        {
          double s = 0.0;
          for (int w = 0; w < 6; ++w) s += R[Cfg::gC + 6 * v + w] * R[Cfg::gB + w * ND + lane];
          CB[v] = fac * s;
        }
#pragma unroll
        for (int c = 0; c < ND; ++c)
        {
          double s = acc[c];
#pragma unroll
          for (int v = 0; v < 6; ++v) s += CB[v] * R[Cfg::gB + v * ND + c];
          acc[c] = s;
        }
        // S . (T^-T G_ab) = (T^-T^T S) . G_ab with the natural-coordinate products G_ab of node a (this lane) and node b
        const double* Sl = R + Cfg::gSl;
        const double* w = R + Cfg::gW;
        const double* dn = R + Cfg::gD;
        const double ar = dn[3 * a], as = dn[3 * a + 1];
        double at[8], a1[8];  // per weight slot: dN_t of node a at the slot's sampling point, and its in-plane partner derivative
        for (int q = 0; q < 8; ++q)
        {
          const double* sd = sp + (size_t)shell_slot_point<CT>(q) * Cfg::spd + 3 * a;
          at[q] = sd[2];
          a1[q] = q < 4 ? 0.0 : (q < 6 ? sd[1] : sd[0]);
        }
#pragma unroll
        for (int b = 0; b < NEN; ++b)
        {
          const double br = dn[3 * b], bs = dn[3 * b + 1];
          double g = Sl[0] * ar * br + Sl[1] * as * bs + Sl[3] * (ar * bs + as * br);
          double g2 = 0.0, g4 = 0.0, g5 = 0.0;
          for (int q = 0; q < 8; ++q)
          {
            const double* sd = sp + (size_t)shell_slot_point<CT>(q) * Cfg::spd + 3 * b;
            if (q < 4)
              g2 += w[q] * at[q] * sd[2];
            else if (q < 6)
              g4 += w[q] * (a1[q] * sd[2] + at[q] * sd[1]);
            else
              g5 += w[q] * (a1[q] * sd[2] + at[q] * sd[0]);
          }
          g += Sl[2] * g2 + Sl[4] * g4 + Sl[5] * g5;
          acc[3 * b + i] += fac * g;
        }
      }
      if (NEAS > 0)
      {
#pragma unroll
        for (int c = 0; c < ND; ++c)
        {
          double s = acc[c];
#pragma unroll
          for (int k = 0; k < NE; ++k) s += mk[k] * Kda[c * NE + k];
          acc[c] = s;
        }
      }
      if (A.Ke_out)
      {
        double* Ke = A.Ke_out + (size_t)e * ND * ND;
        for (int c = 0; c < ND; ++c) Ke[lane + ND * c] = acc[c];
      }
      const int64_t rowbase = (owned && A.node_rowbase) ? __ldg(A.node_rowbase + node) : -1;
      if (A.K && rowbase >= 0 && !(A.dbc && A.dbc[3 * node + i]))
      {
        const int rowlen = __ldg(A.node_rowlen + node);
        const unsigned char* sl = A.slots + (size_t)e * NEN * NEN + a * NEN;
        double* row = A.K + rowbase + (int64_t)i * rowlen;
#pragma unroll
        for (int b = 0; b < NEN; ++b)
        {
          double* p = row + 3 * (int)sl[b];
          red_add(p + 0, acc[3 * b + 0]);
          red_add(p + 1, acc[3 * b + 1]);
          red_add(p + 2, acc[3 * b + 2]);
        }
      }
    }
  }

  // struct_calc_recover of the EAS-ANS elements (update_alpha); one warp per element
  template <int CT, int NEAS>
  __global__ void k_shell_recover(const int* __restrict__ conn, int n_ele, const double* __restrict__ du, double step_length,
      double* __restrict__ hist_all)
  {
    using Cfg = ShellCfg<CT, NEAS>;
    constexpr int ND = Cfg::nd, NEN = Cfg::nen;
    __shared__ double s_tmp[8][Cfg::ne], s_du[8][ND];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int e = blockIdx.x * 8 + w;
    if (e >= n_ele) return;
    double* hist = hist_all + (size_t)e * Cfg::hist_doubles;
    if (lane < ND) s_du[w][lane] = __ldg(du + 3 * (size_t)__ldg(conn + (size_t)e * NEN + lane / 3) + lane % 3);
    __syncwarp();
    if (lane < NEAS)
    {
      double v = hist[Cfg::oS + lane];
      for (int d = 0; d < ND; ++d) v += hist[Cfg::oKda + d + ND * lane] * s_du[w][d];
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
}  // namespace fc
