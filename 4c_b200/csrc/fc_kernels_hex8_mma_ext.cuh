// Tensor-core element kernel for the other hex8 combinations of the path: F-bar + CoupNeoHooke (BASELINE C4)
// and displacement-based St.Venant-Kirchhoff (BASELINE C1).  Same organisation as fc_kernels_hex8nh_mma.cuh
// (a warp owns 4 elements; Gauss-point quantities in m8n8k4 fragment layout in shared memory; every lane ends
// up with the two complete 3x3 blocks K_{a,b0}, K_{a,b0+1}; staged scatter with lane-contiguous FP64 reductions),
// with THREE 8x8 planes per element (the material gradients) and one 3x3 matrix set per Gauss point: every other
// quantity of the closed forms is such a matrix times the material gradient and is applied to the fragments in
// registers, which keeps the shared-memory footprint at the Neo-Hooke kernel's (4 CTAs per SM).
//
// Closed forms (g_a = F^-T grad N_a, p_a = F grad N_a = sum_b x_b (grad N_b . grad N_a), all per Gauss point,
// w detJ = integration factor; reference: 4C_solid_3D_ele_calc_lib.hpp:851-927, 4C_solid_3D_ele_calc_fbar.hpp:65-190,
// 4C_solid_3D_ele_calc_lib_fbar.hpp:49-154, 4C_mat_stvenantkirchhoff.cpp:115-177, 4C_mat_elast_coupneohooke.cpp:60-77):
//
//  F-bar Neo-Hooke, fbar = (det F0 / det F)^(1/3), I3b = det(fbar^2 C):
//    kappa = w detJ 2c I3b^-beta / fbar^3        lambda = w detJ 2c / fbar
//    K_ab  = sum_gp kappa (2 beta g_a g_b^T + g_b g_a^T) + lambda (grad N_a . grad N_b) I        (symmetric part)
//          + (sum_gp v_a) g0_b^T - sum_gp v_a g_b^T,   v_a = 1/3 (kappa (6 beta + 3) g_a - lambda p_a)   (H-operator term,
//            g0_b = F0^-T grad N0_b at the centroid; not symmetric)
//    f_a   = sum_b s_ab x_b - sum_gp kappa g_a,   s_ab = sum_gp lambda grad N_a . grad N_b
//  St.Venant-Kirchhoff (Lame L, mu; S = L tr(E) I + 2 mu E; b = F F^T):
//    K_ab  = sum_gp w detJ [L p_a p_b^T + mu p_b p_a^T + mu (grad N_a . grad N_b) b] + s_ab I,  s_ab = sum_gp w detJ grad N_a . S grad N_b
//    f_a   = sum_b s_ab x_b
//
// DMMA count per element: F-bar 48 (6 s + 18 W + 18 V + 6 vsum), SVK 60 (6 s + 18 W + 36 for the six b_ij-weighted
// (grad N_a . grad N_b) tiles), vs 108+ for a dense B^T C B on the same tensor cores.
#pragma once
#include "fc_kernels_hex8nh_mma.cuh"

namespace fc
{
  enum
  {
    VAR_NH_FBAR = 0,
    VAR_SVK_DISP = 1,
    // KINEM linear with St.Venant-Kirchhoff (4C_solid_3D_ele_calc_displacement_based_linear_kinematics.hpp:36-155): the SVK form
    // with the B-operator at F = I (p_a = grad N_a, b = I), the stress of the linear strain sym(grad u), no geometric
    // stiffness, and f_a = sum_b s_ab X_b on the reference coordinates (sum_b grad N_b X_b^T = I)
    VAR_SVK_LINEAR = 2
  };

  struct Hex8MmaExtCfg
  {
    static constexpr int warps = 4;
    static constexpr int threads = 32 * warps;
    static constexpr int epw = 4;
    static constexpr int epb = warps * epw;
    // per element: 3 planes of 64 with the MATERIAL gradients n_a = J^-1 dN_a in fragment layout - every other
    // Gauss-point quantity of the closed forms is a 3x3 matrix (per Gauss point) times n_a and is applied to the
    // fragments in registers - then one record of 24 doubles per Gauss point, then x rel [8][3]:
    //   F-bar: [0..8] sqrt(kappa) j^-1 J (-> ghat), [9..17] (-> v / sqrt(kappa)), [18] sqrt(lambda), [19] sqrt(kappa),
    //          [20..22] centroid gradient g0 of node gp
    //   SVK:   [0..8] sqrt(w detJ) F (-> phat), [9..17] w detJ S (-> t), [18..23] mu w detJ b_ij
    // Record stride 25, not 24: in phase 2 the 8 lanes of a quad column read the same record and the 4 columns read records q
    // (or 2q) apart; with 24 doubles = 48 words the four addresses fall on banks 0/16/0/16 (0/0/0/0 for 2q) - ncu counted 152 M
    // excess wavefronts per 10^6 elements, 36 % of the shared-memory traffic of a kernel that sits at 84 % of the l1tex peak.
    // 25 doubles = 50 words: banks 0/18/4/22 (0/4/8/12).
    // 417 doubles per element -> still 4 CTAs per SM
    static constexpr int oR = 192, rec = 25, oX = oR + 8 * rec, ele_doubles = oX + 24 + 1;
    static constexpr int ks_stride = 24;
    static constexpr size_t smem_bytes = sizeof(double) * (size_t)epb * ele_doubles;
  };

  template <int VAR, bool WANT_K, bool HAS_DBC, bool HAS_PERM>
  __global__ void __launch_bounds__(Hex8MmaExtCfg::threads, 4) k_hex8_mma_ext(const EvalArgs A, const unsigned char* __restrict__ ele_bad)
  {
    using Cfg = Hex8MmaExtCfg;
    extern __shared__ __align__(16) double smem_mma_ext[];
    double* const smem = smem_mma_ext;
    __shared__ long long s_rowbase[Cfg::epb][8];
    __shared__ int s_rowlen[Cfg::epb][8];
    __shared__ int s_node[Cfg::epb][8];
    __shared__ unsigned long long s_slots[Cfg::epb][8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int el = lane >> 3, t = lane & 7;
    const int wel = warp * Cfg::epw;
    const int slot_e = A.ele_offset + blockIdx.x * Cfg::epb + wel + el;
    const bool active = slot_e < A.n_ele;
    // launch slot -> element: the caller's order, or the L2 sweep order of large meshes (fc_sweep_order.hpp)
    const int e = active ? (HAS_PERM ? __ldg(A.ele_perm + slot_e) : slot_e) : -1;  // HAS_PERM: see k_hex8_neohooke_mma
    double* E = smem + (size_t)(wel + el) * Cfg::ele_doubles;

    // ---- phase 0: lane (el, t) gathers node t of element el ------------------------------------------
    double Xr[3] = {0, 0, 0}, xr[3] = {0, 0, 0};
    int node = 0;
    if (active)
    {
      node = __ldg(A.conn + (size_t)e * 8 + t);
      for (int i = 0; i < 3; ++i)
      {
        Xr[i] = __ldg(A.X + 3 * (size_t)node + i);
        xr[i] = __ldg(A.u + 3 * (size_t)node + i);
      }
    }
    {
      const bool owned = active && node < A.n_row_nodes;
      long long rb = -1;
      int rl = 0;
      unsigned long long sl = 0;
      if (WANT_K && owned && A.K)
      {
        rb = __ldg(A.node_rowbase + node);
        rl = __ldg(A.node_rowlen + node);
        sl = __ldg(reinterpret_cast<const unsigned long long*>(A.slots + (size_t)e * 64) + t);
        if (HAS_DBC)
          for (int i = 0; i < 3; ++i) rl |= (A.dbc[3 * node + i] ? 1 : 0) << (28 + i);
      }
      s_rowbase[wel + el][t] = rb;
      s_rowlen[wel + el][t] = rl;
      s_slots[wel + el][t] = sl;
      s_node[wel + el][t] = owned ? node : -1;
    }
    for (int i = 0; i < 3; ++i)
    {
      const double X0 = __shfl_sync(0xFFFFFFFFu, Xr[i], lane & 24);
      const double u0 = __shfl_sync(0xFFFFFFFFu, xr[i], lane & 24);
      Xr[i] -= X0;
      xr[i] = Xr[i] + (xr[i] - u0);
    }

    // ---- phase 1: lane (el, t) evaluates Gauss point t of element el ------------------------------------
    const double mu = A.mat.young / (2.0 * (1.0 + A.mat.nue));  // = 2c of CoupNeoHooke
    const double lame = A.mat.young * A.mat.nue / ((1.0 + A.mat.nue) * (1.0 - 2.0 * A.mat.nue));
    const double beta = A.mat.nue / (1.0 - 2.0 * A.mat.nue);
    bool bad = false;
    {
      const double p = 0.57735026918962576451;
      const double xi0 = (hex_node_pos(t, 0) > 0) ? p : -p;
      const double xi1 = (hex_node_pos(t, 1) > 0) ? p : -p;
      const double xi2 = (hex_node_pos(t, 2) > 0) ? p : -p;
      double* R = E + Cfg::oR + Cfg::rec * t;
      double detF0 = 1.0;
      if (VAR == VAR_NH_FBAR)
      {
        // centroid pass first (its own sweep over the nodes: 18 accumulators less in the Gauss-point sweep below)
        double jc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, Jc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int a = 0; a < 8; ++a)
        {
          const int src = (lane & 24) | a;
          const double dc[3] = {0.125 * hex_node_pos(a, 0), 0.125 * hex_node_pos(a, 1), 0.125 * hex_node_pos(a, 2)};
#pragma unroll
          for (int i = 0; i < 3; ++i)
          {
            const double xa = __shfl_sync(0xFFFFFFFFu, xr[i], src), Xa = __shfl_sync(0xFFFFFFFFu, Xr[i], src);
#pragma unroll
            for (int d = 0; d < 3; ++d)
            {
              jc[3 * d + i] += dc[d] * xa;
              Jc[3 * d + i] += dc[d] * Xa;
            }
          }
        }
        double jci[9];
        const double detjc = inv3(jc, jci);
        detF0 = detjc / det3(Jc);
        // centroid gradient of node t: g0_t = jc^-1 dN0_t
        const double c0 = 0.125 * hex_node_pos(t, 0), c1 = 0.125 * hex_node_pos(t, 1), c2 = 0.125 * hex_node_pos(t, 2);
        if (active)
          for (int i = 0; i < 3; ++i) R[20 + i] = jci[3 * i] * c0 + jci[3 * i + 1] * c1 + jci[3 * i + 2] * c2;
      }
      double jm[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, Jm[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
      for (int a = 0; a < 8; ++a)
      {
        const int src = (lane & 24) | a;
        double xa[3], Xa[3];
#pragma unroll
        for (int i = 0; i < 3; ++i)
        {
          xa[i] = __shfl_sync(0xFFFFFFFFu, xr[i], src);
          Xa[i] = __shfl_sync(0xFFFFFFFFu, Xr[i], src);
        }
        const double s0 = hex_node_pos(a, 0), s1 = hex_node_pos(a, 1), s2 = hex_node_pos(a, 2);
        const double r = 1.0 + s0 * xi0, s = 1.0 + s1 * xi1, q = 1.0 + s2 * xi2;
        const double dNa[3] = {0.125 * s0 * s * q, 0.125 * s1 * q * r, 0.125 * s2 * r * s};
#pragma unroll
        for (int d = 0; d < 3; ++d)
#pragma unroll
          for (int i = 0; i < 3; ++i)
          {
            jm[3 * d + i] += dNa[d] * xa[i];
            Jm[3 * d + i] += dNa[d] * Xa[i];
          }
      }
      double ji[9], Ji[9];
      const double detj = inv3(jm, ji);
      const double detJ = inv3(Jm, Ji);
      const double detF = detj / detJ;
      // F[i][I] = sum_d jm[d][i] Ji[I][d]
      double F[9];
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int I = 0; I < 3; ++I) F[3 * i + I] = jm[i] * Ji[3 * I] + jm[3 + i] * Ji[3 * I + 1] + jm[6 + i] * Ji[3 * I + 2];

      // material gradients of all nodes at this Gauss point, fragment layout
      if (active)
      {
#pragma unroll
        for (int a = 0; a < 8; ++a)
        {
          // shape function derivatives recomputed (6 flops) instead of kept in 48 registers across the inversions
          const double s0 = hex_node_pos(a, 0), s1 = hex_node_pos(a, 1), s2 = hex_node_pos(a, 2);
          const double r = 1.0 + s0 * xi0, s = 1.0 + s1 * xi1, q = 1.0 + s2 * xi2;
          const double d0 = 0.125 * s0 * s * q, d1 = 0.125 * s1 * q * r, d2 = 0.125 * s2 * r * s;
          const int ix = frag_idx(t, a);
#pragma unroll
          for (int I = 0; I < 3; ++I) E[I * 64 + ix] = Ji[3 * I] * d0 + Ji[3 * I + 1] * d1 + Ji[3 * I + 2] * d2;
        }
      }
      if (VAR == VAR_NH_FBAR)
      {
        const double fbar = cbrt(detF0 / detF);  // reference: std::pow(x, 1/3.) - equal to an ulp
        const double jb = fbar * fbar * fbar * detF;
        const double I3b = jb * jb;
        const double lambda = detJ * mu / fbar;
        const double kappa = detJ * mu * exp(-beta * log(I3b)) / (fbar * fbar * fbar);
        bad = active && (!(detJ > 0.0) || detF == 0.0 || !isfinite(detF) || !isfinite(fbar) || !(I3b > 0.0) || !isfinite(kappa) ||
                            (ele_bad && ele_bad[e]));
        const double sk = sqrt(kappa), sl = sqrt(lambda);
        const double cg = sk * (2.0 * beta + 1.0), cp = -lambda / (3.0 * sk);  // v_a / sk = cg g_a + cp p_a
        if (active)
        {
          // g_a = (j^-1 J) n_a, p_a = F n_a
#pragma unroll
          for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int I = 0; I < 3; ++I)
            {
              const double gm = ji[3 * i] * Jm[I] + ji[3 * i + 1] * Jm[3 + I] + ji[3 * i + 2] * Jm[6 + I];
              R[3 * i + I] = sk * gm;
              R[9 + 3 * i + I] = cg * gm + cp * F[3 * i + I];
            }
          R[18] = sl;
          R[19] = sk;
        }
      }
      else
      {
        // E = (F^T F - I)/2, S = L tr(E) I + 2 mu E, b = F F^T
        double C[6], S[6], b[6];
        const int vi[6] = {0, 1, 2, 0, 1, 0}, vj[6] = {0, 1, 2, 1, 2, 2};  // (00,11,22,01,12,02)
#pragma unroll
        for (int v = 0; v < 6; ++v)
        {
          C[v] = F[vi[v]] * F[vj[v]] + F[3 + vi[v]] * F[3 + vj[v]] + F[6 + vi[v]] * F[6 + vj[v]];
          b[v] = F[3 * vi[v]] * F[3 * vj[v]] + F[3 * vi[v] + 1] * F[3 * vj[v] + 1] + F[3 * vi[v] + 2] * F[3 * vj[v] + 2];
        }
        const double trE = 0.5 * (C[0] + C[1] + C[2] - 3.0);
#pragma unroll
        for (int v = 0; v < 3; ++v) S[v] = lame * trE + mu * (C[v] - 1.0);
#pragma unroll
        for (int v = 3; v < 6; ++v) S[v] = mu * C[v];
        bad = active && (!(detJ > 0.0) || detF == 0.0 || !isfinite(detF) || (ele_bad && ele_bad[e]));
        if (VAR == VAR_SVK_LINEAR)
        {
          const double tr = (F[0] - 1.0) + (F[4] - 1.0) + (F[8] - 1.0);
          S[0] = lame * tr + 2.0 * mu * (F[0] - 1.0);
          S[1] = lame * tr + 2.0 * mu * (F[4] - 1.0);
          S[2] = lame * tr + 2.0 * mu * (F[8] - 1.0);
          S[3] = mu * (F[1] + F[3]);
          S[4] = mu * (F[5] + F[7]);
          S[5] = mu * (F[2] + F[6]);
#pragma unroll
          for (int v = 0; v < 6; ++v) b[v] = v < 3 ? 1.0 : 0.0;
#pragma unroll
          for (int i = 0; i < 9; ++i) F[i] = (i % 4 == 0) ? 1.0 : 0.0;
          bad = active && (!(detJ > 0.0) || (ele_bad && ele_bad[e]));
        }
        const double sw = sqrt(detJ);
        if (active)
        {
#pragma unroll
          for (int i = 0; i < 9; ++i) R[i] = sw * F[i];
          const int sym[9] = {0, 3, 5, 3, 1, 4, 5, 4, 2};
#pragma unroll
          for (int i = 0; i < 9; ++i) R[9 + i] = detJ * S[sym[i]];
#pragma unroll
          for (int v = 0; v < 6; ++v) R[18 + v] = mu * detJ * b[v];
        }
      }
      if (active)
        for (int i = 0; i < 3; ++i) E[Cfg::oX + 3 * t + i] = VAR == VAR_SVK_LINEAR ? Xr[i] : xr[i];
    }
    const unsigned badmask = __ballot_sync(0xFFFFFFFFu, bad);
    const unsigned actmask = __ballot_sync(0xFFFFFFFFu, active);
    if (active && t == 0 && ((badmask >> (lane & 24)) & 0xFFu)) report_error(A.err, __ldg(A.ele_gid + e));
    __syncwarp();

    // ---- phase 2 + 3: one element at a time, whole warp ---------------------------------------------------
    const int a = lane >> 2, q = lane & 3, b0 = 2 * q;
#pragma unroll 1
    for (int k = 0; k < Cfg::epw; ++k)
    {
      if (!((actmask >> (8 * k)) & 1u) || ((badmask >> (8 * k)) & 0xFFu)) continue;  // warp-uniform
      const int ek = HAS_PERM ? __shfl_sync(0xFFFFFFFFu, e, 8 * k) : A.ele_offset + blockIdx.x * Cfg::epb + wel + k;
      const double* G = smem + (size_t)(wel + k) * Cfg::ele_doubles;
      const int fo0 = frag_idx(q, a), fo1 = frag_idx(q + 4, a);
      const double* R0 = G + Cfg::oR + Cfg::rec * q;
      const double* R1 = G + Cfg::oR + Cfg::rec * (q + 4);
      double fn[3][2];  // material gradient fragments: node a, Gauss points q and q+4
#pragma unroll
      for (int I = 0; I < 3; ++I)
      {
        fn[I][0] = G[I * 64 + fo0];
        fn[I][1] = G[I * 64 + fo1];
      }
      // fragment of (3x3 matrix at offset o of the Gauss-point record) * n_a
      auto apply = [&](int o, double(&out)[3][2])
      {
#pragma unroll
        for (int i = 0; i < 3; ++i)
        {
          out[i][0] = R0[o + 3 * i] * fn[0][0] + R0[o + 3 * i + 1] * fn[1][0] + R0[o + 3 * i + 2] * fn[2][0];
          out[i][1] = R1[o + 3 * i] * fn[0][1] + R1[o + 3 * i + 1] * fn[1][1] + R1[o + 3 * i + 2] * fn[2][1];
        }
      };
      double fg[3][2];  // ghat (F-bar) or phat (SVK)
      apply(0, fg);
      // s_ab: lambda-weighted (F-bar) or S-weighted (SVK) gradient products
      double sa[3][2] = {{0.0, 0.0}, {0.0, 0.0}, {0.0, 0.0}};
      double sk0 = 0.0, sk1 = 0.0;
      if (VAR == VAR_NH_FBAR)
      {
        const double sl0 = R0[18], sl1 = R1[18];
        sk0 = R0[19];
        sk1 = R1[19];
#pragma unroll
        for (int i = 0; i < 3; ++i)
        {
          const double n0 = sl0 * fn[i][0], n1 = sl1 * fn[i][1];
          dmma884(sa[i][0], sa[i][1], n0, n0);
          dmma884(sa[i][0], sa[i][1], n1, n1);
        }
      }
      else
      {
        double ft[3][2];
        apply(9, ft);
#pragma unroll
        for (int i = 0; i < 3; ++i)
        {
          dmma884(sa[i][0], sa[i][1], ft[i][0], fn[i][0]);
          dmma884(sa[i][0], sa[i][1], ft[i][1], fn[i][1]);
        }
      }
      const double s0 = sa[0][0] + sa[1][0] + sa[2][0], s1 = sa[0][1] + sa[1][1] + sa[2][1];
      {
        const double* xk = G + Cfg::oX;
        double fpart[3];
#pragma unroll
        for (int i = 0; i < 3; ++i)
        {
          fpart[i] = s0 * xk[3 * b0 + i] + s1 * xk[3 * b0 + 3 + i];
          if (VAR == VAR_NH_FBAR) fpart[i] -= sk0 * fg[i][0] + sk1 * fg[i][1];
          fpart[i] += __shfl_xor_sync(0xFFFFFFFFu, fpart[i], 1);
          fpart[i] += __shfl_xor_sync(0xFFFFFFFFu, fpart[i], 2);
        }
        const int na = s_node[wel + k][a];
        if (q < 3)
        {
          const double v = q == 0 ? fpart[0] : (q == 1 ? fpart[1] : fpart[2]);
          if (A.fe_out) A.fe_out[(size_t)ek * 24 + 3 * a + q] = v;
          if (A.f && na >= 0) red_add(A.f + 3 * (size_t)na + q, v);
        }
      }
      if (!WANT_K) continue;
      double Kb[2][9];
      {
        double W[3][3][2];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
          for (int j = 0; j < 3; ++j)
          {
            W[i][j][0] = 0.0;
            W[i][j][1] = 0.0;
            dmma884(W[i][j][0], W[i][j][1], fg[i][0], fg[j][0]);
            dmma884(W[i][j][0], W[i][j][1], fg[i][1], fg[j][1]);
          }
        const double c1 = VAR == VAR_NH_FBAR ? 2.0 * beta : lame, c2 = VAR == VAR_NH_FBAR ? 1.0 : mu;
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
          for (int j = 0; j < 3; ++j)
          {
            Kb[0][3 * i + j] = c1 * W[i][j][0] + c2 * W[j][i][0] + (i == j && VAR != VAR_SVK_LINEAR ? s0 : 0.0);
            Kb[1][3 * i + j] = c1 * W[i][j][1] + c2 * W[j][i][1] + (i == j && VAR != VAR_SVK_LINEAR ? s1 : 0.0);
          }
      }
      if (VAR == VAR_NH_FBAR)
      {
        // H-operator term: (sum_gp v_a) g0_b^T - sum_gp v_a g_b^T
        double fv[3][2];
        apply(9, fv);
        const double* g0a = G + Cfg::oR + Cfg::rec * b0 + 20;
        const double* g0b = g0a + Cfg::rec;
#pragma unroll
        for (int i = 0; i < 3; ++i)
        {
          double vs0 = 0.0, vs1 = 0.0;
          dmma884(vs0, vs1, fv[i][0], sk0);
          dmma884(vs0, vs1, fv[i][1], sk1);
#pragma unroll
          for (int j = 0; j < 3; ++j)
          {
            double v0 = 0.0, v1 = 0.0;
            dmma884(v0, v1, fv[i][0], fg[j][0]);
            dmma884(v0, v1, fv[i][1], fg[j][1]);
            Kb[0][3 * i + j] += vs0 * g0a[j] - v0;
            Kb[1][3 * i + j] += vs1 * g0b[j] - v1;
          }
        }
      }
      else
      {
        // mu w detJ b_ij (grad N_a . grad N_b): six weighted tiles, the weight rides on the A fragment
        const int vi[6] = {0, 1, 2, 0, 1, 0}, vj[6] = {0, 1, 2, 1, 2, 2};
#pragma unroll
        for (int v = 0; v < (VAR == VAR_SVK_LINEAR ? 3 : 6); ++v)  // b = I with linear kinematics: the diagonal tiles only
        {
          double t0 = 0.0, t1 = 0.0;
          const double w0 = R0[18 + v], w1 = R1[18 + v];
#pragma unroll
          for (int J = 0; J < 3; ++J)
          {
            dmma884(t0, t1, w0 * fn[J][0], fn[J][0]);
            dmma884(t0, t1, w1 * fn[J][1], fn[J][1]);
          }
          Kb[0][3 * vi[v] + vj[v]] += t0;
          Kb[1][3 * vi[v] + vj[v]] += t1;
          if (v >= 3)
          {
            Kb[0][3 * vj[v] + vi[v]] += t0;
            Kb[1][3 * vj[v] + vi[v]] += t1;
          }
        }
      }
      if (A.Ke_out)
      {
        double* Ke = A.Ke_out + (size_t)ek * 576;
        for (int c = 0; c < 2; ++c)
          for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) Ke[(3 * a + i) + 24 * (3 * (b0 + c) + j)] = Kb[c][3 * i + j];
      }
      if (A.K)
      {
        // staged scatter, identical to fc_kernels_hex8nh_mma.cuh (see there)
        double* Ks = smem + (size_t)(wel + k) * Cfg::ele_doubles + ((wel + k) & 1);
        const int ln = lane < 24 ? lane : 23;
        const int b = ln / 3;
#pragma unroll
        for (int half = 0; half < 2; ++half)
        {
          __syncwarp();
          if ((a >> 2) == half)
          {
#pragma unroll
            for (int i = 0; i < 3; ++i)
            {
              double2* row = reinterpret_cast<double2*>(Ks + (3 * (a & 3) + i) * Cfg::ks_stride + 6 * q);
              row[0] = make_double2(Kb[0][3 * i + 0], Kb[0][3 * i + 1]);
              row[1] = make_double2(Kb[0][3 * i + 2], Kb[1][3 * i + 0]);
              row[2] = make_double2(Kb[1][3 * i + 1], Kb[1][3 * i + 2]);
            }
          }
          __syncwarp();
          double v[12];
          double* ptr[4];
          int rl[4];
#pragma unroll
          for (int r = 0; r < 4; ++r)
          {
            const int aa = 4 * half + r;
            const long long rb = s_rowbase[wel + k][aa];
            const int slot = (int)((s_slots[wel + k][aa] >> (8 * b)) & 0xFF);
            rl[r] = s_rowlen[wel + k][aa];
            ptr[r] = (rb >= 0 && lane < 24) ? A.K + rb + 3 * slot + (ln - 3 * b) : nullptr;
#pragma unroll
            for (int i = 0; i < 3; ++i) v[3 * r + i] = Ks[(3 * r + i) * Cfg::ks_stride + ln];
          }
#pragma unroll
          for (int r = 0; r < 4; ++r)
          {
            if (HAS_DBC)
            {
#pragma unroll
              for (int i = 0; i < 3; ++i)
                red_add_if(ptr[r] != nullptr && !((rl[r] >> (28 + i)) & 1), ptr[r] + (long long)i * (rl[r] & 0x0FFFFFFF), v[3 * r + i]);
            }
            else
              red_add3_if(ptr[r] != nullptr, ptr[r], 8ll * rl[r], v[3 * r], v[3 * r + 1], v[3 * r + 2]);
          }
        }
      }
    }
  }
}  // namespace fc
