// Specialised element kernel: hex8, displacement-based, MAT_ElastHyper + ELAST_CoupNeoHooke.
//
// For Psi = c (I1-3) + c/beta (I3^-beta - 1) the reference's
//   gamma = (2c, 0, -2c I3^-beta),  delta5 = 4 c beta I3^-beta,  delta6 = 4 c I3^-beta
// (src/mat/4C_mat_elasthyper_service.cpp:413-431 with the derivatives of
// src/mat/elast/4C_mat_elast_coupneohooke.cpp:60-77), so with the spatial shape-function gradient
// g_a = F^-T grad N_a = (dN x^T)^-1 dN_a the element integrals of
// src/solid_3D_ele/4C_solid_3D_ele_calc_lib.hpp:851-927 collapse to
//   K_ab = sum_gp kappa (2 beta g_a g_b^T + g_b g_a^T) + sum_gp lambda (g_a . b g_b) I3x3
//   f_a  = sum_gp (lambda b g_a - kappa g_a)
//   kappa = w detJ 2c I3^-beta,  lambda = w detJ 2c,  b = F F^T
// (B^T C B + K_geo: the (delta6/2 + gamma2)(g_a.g_b) I terms cancel exactly).  This is the same
// algebra as the reference evaluated in a different order: ~17 kFLOP per element instead of ~94 k.
//
// Work decomposition: 8 threads per element (4 elements per warp, warp-synchronous).
//   phase 0  thread t gathers node t
//   phase 1  thread g evaluates Gauss point g: ghat_a = sqrt(kappa) g_a, m_a = lambda/sqrt(kappa) b g_a -> smem
//   phase 2  thread t accumulates T_ab = sum ghat_a ghat_b^T, s_ab = sum ghat_a . m_b for the node pairs
//            (t, t+k mod 8), k = 0..4 (36 unordered pairs, symmetry of K_e exploited), and f_t
//   phase 3  K_ab = 2 beta T + T^T + s I and its transpose are added to the CSR rows of a and b
#pragma once
#include "fc_kernels_common.cuh"

namespace fc
{
  struct Hex8NHCfg
  {
    static constexpr int epb = 8;              // elements per CTA
    static constexpr int threads = epb * 8;
    static constexpr int gp_doubles = 50;      // ghat[8][3], m[8][3], sqrt(kappa), pad
    static constexpr int ele_doubles = 8 * gp_doubles + 48 + 1;  // + x[8][3], X[8][3], +1 pad (2-bank skew between elements)
    static constexpr size_t smem_bytes = sizeof(double) * (size_t)epb * ele_doubles;
  };

  template <bool WANT_K>
  __global__ void __launch_bounds__(Hex8NHCfg::threads) k_hex8_neohooke(const EvalArgs A, const unsigned char* __restrict__ ele_bad)
  {
    using Cfg = Hex8NHCfg;
    extern __shared__ double smem[];
    const int el = threadIdx.x >> 3, t = threadIdx.x & 7;
    const int lane = threadIdx.x & 31;
    const int e = blockIdx.x * Cfg::epb + el;
    const bool active = e < A.n_ele;
    double* xs = smem + (size_t)el * Cfg::ele_doubles;  // current coordinates [8][3]
    double* Xs = xs + 24;                                // reference coordinates [8][3]
    double* gpd = Xs + 24;                               // [8][gp_doubles]

    int node = 0;
    if (active)
    {
      node = __ldg(A.conn + (size_t)e * 8 + t);
      for (int i = 0; i < 3; ++i)
      {
        const double X = __ldg(A.X + 3 * (size_t)node + i);
        Xs[3 * t + i] = X;
        xs[3 * t + i] = X + __ldg(A.u + 3 * (size_t)node + i);
      }
    }
    __syncwarp();

    // ---- phase 1: Gauss point t ------------------------------------------------------------------
    bool bad = false;
    if (active)
    {
      const double p = 0.57735026918962576451;
      const double xi0 = (hex_node_pos(t, 0) > 0) ? p : -p;
      const double xi1 = (hex_node_pos(t, 1) > 0) ? p : -p;
      const double xi2 = (hex_node_pos(t, 2) > 0) ? p : -p;
      double jm[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, Jm[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      double dN[24];
#pragma unroll
      for (int a = 0; a < 8; ++a)
      {
        const double s0 = hex_node_pos(a, 0), s1 = hex_node_pos(a, 1), s2 = hex_node_pos(a, 2);
        const double r = 1.0 + s0 * xi0, s = 1.0 + s1 * xi1, q = 1.0 + s2 * xi2;
        dN[3 * a + 0] = 0.125 * s0 * s * q;
        dN[3 * a + 1] = 0.125 * s1 * q * r;
        dN[3 * a + 2] = 0.125 * s2 * r * s;
#pragma unroll
        for (int d = 0; d < 3; ++d)
#pragma unroll
          for (int i = 0; i < 3; ++i)
          {
            jm[3 * d + i] += dN[3 * a + d] * xs[3 * a + i];
            Jm[3 * d + i] += dN[3 * a + d] * Xs[3 * a + i];
          }
      }
      double ji[9], Ji[9];
      const double detj = inv3(jm, ji);
      const double detJ = inv3(Jm, Ji);
      // F[i][I] = sum_d Ji[I][d] jm[d][i];  b = F F^T
      double F[9];
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int I = 0; I < 3; ++I) F[3 * i + I] = Ji[3 * I] * jm[i] + Ji[3 * I + 1] * jm[3 + i] + Ji[3 * I + 2] * jm[6 + i];
      double b[9];
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int k = 0; k < 3; ++k) b[3 * i + k] = F[3 * i] * F[3 * k] + F[3 * i + 1] * F[3 * k + 1] + F[3 * i + 2] * F[3 * k + 2];
      const double detF = detj / detJ;
      const double I3 = detF * detF;
      const double c2 = A.mat.young / (2.0 * (1.0 + A.mat.nue));  // 2c
      const double beta = A.mat.nue / (1.0 - 2.0 * A.mat.nue);
      const double lambda = detJ * c2;  // Gauss weight of hex_8point is 1
      const double kappa = lambda * exp(-beta * log(I3));
      bad = !(detJ > 0.0) || !(I3 > 0.0) || !isfinite(kappa);
      const double sk = sqrt(kappa);
      const double lk = lambda / sk;
      double* R = gpd + t * Cfg::gp_doubles;
#pragma unroll
      for (int a = 0; a < 8; ++a)
      {
        const double d0 = dN[3 * a], d1 = dN[3 * a + 1], d2 = dN[3 * a + 2];
        const double g0 = ji[0] * d0 + ji[1] * d1 + ji[2] * d2;
        const double g1 = ji[3] * d0 + ji[4] * d1 + ji[5] * d2;
        const double g2 = ji[6] * d0 + ji[7] * d1 + ji[8] * d2;
        R[6 * a + 0] = sk * g0;
        R[6 * a + 1] = sk * g1;
        R[6 * a + 2] = sk * g2;
        R[6 * a + 3] = lk * (b[0] * g0 + b[1] * g1 + b[2] * g2);
        R[6 * a + 4] = lk * (b[3] * g0 + b[4] * g1 + b[5] * g2);
        R[6 * a + 5] = lk * (b[6] * g0 + b[7] * g1 + b[8] * g2);
      }
      R[48] = sk;
    }
    if (active && ele_bad && ele_bad[e]) bad = true;
    // element-wide error vote over its 8 lanes
    const unsigned grp = 0xFFu << (lane & 24);
    const bool ele_err = (__ballot_sync(0xFFFFFFFFu, bad) & grp) != 0;
    __syncwarp();
    if (active && ele_err && t == 0) report_error(A.err, __ldg(A.ele_gid + e));
    const bool live = active && !ele_err;  // no early exit: all lanes take part in the shuffles below

    // ---- phase 2 ----------------------------------------------------------------------------------
    double T[WANT_K ? 45 : 1], s[WANT_K ? 5 : 1];
    if (WANT_K)
    {
#pragma unroll
      for (int i = 0; i < 45; ++i) T[i] = 0.0;
#pragma unroll
      for (int i = 0; i < 5; ++i) s[i] = 0.0;
    }
    double fa[3] = {0, 0, 0};
    const int npair = live ? (t < 4 ? 5 : 4) : 0;
#pragma unroll 1
    for (int gp = 0; gp < (live ? 8 : 0); ++gp)
    {
      const double* R = gpd + gp * Cfg::gp_doubles;
      const double g0 = R[6 * t], g1 = R[6 * t + 1], g2 = R[6 * t + 2];
      const double sk = R[48];
      fa[0] += sk * (R[6 * t + 3] - g0);
      fa[1] += sk * (R[6 * t + 4] - g1);
      fa[2] += sk * (R[6 * t + 5] - g2);
      if (WANT_K)
      {
#pragma unroll
        for (int k = 0; k < 5; ++k)
        {
          if (k < npair)
          {
            const double* Rb = R + 6 * ((t + k) & 7);
            const double h0 = Rb[0], h1 = Rb[1], h2 = Rb[2];
            T[9 * k + 0] += g0 * h0; T[9 * k + 1] += g0 * h1; T[9 * k + 2] += g0 * h2;
            T[9 * k + 3] += g1 * h0; T[9 * k + 4] += g1 * h1; T[9 * k + 5] += g1 * h2;
            T[9 * k + 6] += g2 * h0; T[9 * k + 7] += g2 * h1; T[9 * k + 8] += g2 * h2;
            s[k] += g0 * Rb[3] + g1 * Rb[4] + g2 * Rb[5];
          }
        }
      }
    }

    // ---- phase 3 ----------------------------------------------------------------------------------
    const bool owned = live && node < A.n_row_nodes;
    if (A.fe_out && live)
      for (int i = 0; i < 3; ++i) A.fe_out[(size_t)e * 24 + 3 * t + i] = fa[i];
    if (A.f && owned)
      for (int i = 0; i < 3; ++i) red_add(A.f + 3 * (size_t)node + i, fa[i]);
    if (!WANT_K) return;  // uniform

    const double beta2 = 2.0 * A.mat.nue / (1.0 - 2.0 * A.mat.nue);
    int64_t rowbase = -1;
    int rowlen = 0;
    unsigned long long myslots = 0;  // slots[e][t][0..7]
    unsigned dbcmask = 0;
    if (A.K)
    {
      if (owned)
      {
        rowbase = __ldg(A.node_rowbase + node);
        rowlen = __ldg(A.node_rowlen + node);
        if (A.dbc)
          for (int i = 0; i < 3; ++i) dbcmask |= (A.dbc[3 * node + i] ? 1u : 0u) << i;
      }
      if (live) myslots = __ldg(reinterpret_cast<const unsigned long long*>(A.slots + (size_t)e * 64) + t);
    }
#pragma unroll
    for (int k = 0; k < 5; ++k)
    {
      const int bnode = (t + k) & 7;
      const int src = (lane & 24) | bnode;
      // partner row data (all lanes participate in the shuffles)
      const int64_t brow = __shfl_sync(0xFFFFFFFFu, rowbase, src);
      const int blen = __shfl_sync(0xFFFFFFFFu, rowlen, src);
      const unsigned long long bslots = __shfl_sync(0xFFFFFFFFu, myslots, src);
      const unsigned bdbc = __shfl_sync(0xFFFFFFFFu, dbcmask, src);
      if (k < npair)
      {
        double Kab[9];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
          for (int j = 0; j < 3; ++j) Kab[3 * i + j] = beta2 * T[9 * k + 3 * i + j] + T[9 * k + 3 * j + i] + (i == j ? s[k] : 0.0);
        if (A.Ke_out)
        {
          double* Ke = A.Ke_out + (size_t)e * 576;
          for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j)
            {
              Ke[(3 * t + i) + 24 * (3 * bnode + j)] = Kab[3 * i + j];
              Ke[(3 * bnode + j) + 24 * (3 * t + i)] = Kab[3 * i + j];
            }
        }
        if (A.K)
        {
          if (rowbase >= 0)
          {
            const int slot = (int)((myslots >> (8 * bnode)) & 0xFF);
#pragma unroll
            for (int i = 0; i < 3; ++i)
              if (!((dbcmask >> i) & 1u))
              {
                double* pr = A.K + rowbase + (int64_t)i * rowlen + 3 * slot;
                red_add(pr + 0, Kab[3 * i + 0]);
                red_add(pr + 1, Kab[3 * i + 1]);
                red_add(pr + 2, Kab[3 * i + 2]);
              }
          }
          if (k > 0 && brow >= 0)
          {
            const int slot = (int)((bslots >> (8 * t)) & 0xFF);
#pragma unroll
            for (int j = 0; j < 3; ++j)
              if (!((bdbc >> j) & 1u))
              {
                double* pr = A.K + brow + (int64_t)j * blen + 3 * slot;
                red_add(pr + 0, Kab[0 + j]);
                red_add(pr + 1, Kab[3 + j]);
                red_add(pr + 2, Kab[6 + j]);
              }
          }
        }
      }
    }
  }
}  // namespace fc
