// tet10, displacement-based CoupNeoHooke (BASELINE C3b): one warp per element, closed form on the FP64 FMA pipe.
//
// Same closed form as fc_kernels_nh_mma_gen.cuh (what it replaces in the reference: SolidEleCalc::
// evaluate_nonlinear_force_stiffness_mass, src/solid_3D_ele/4C_solid_3D_ele_calc.cpp:110-240, with the tet_4point rule of
// 4C_solid_3D_ele_calc_lib_integration.hpp:51-55 and CoupNeoHooke, src/mat/elast/4C_mat_elast_coupneohooke.cpp:60-77):
//   K[3a+i][3b+j] = 2 beta sum_g gh[g][a][i] gh[g][b][j] + sum_g gh[g][a][j] gh[g][b][i] + delta_ij s_ab
//   s_ab          = sum_g nh[g][a] . nh[g][b]
//   f[3b+j]       = sum_a s_ab x_a[j] - sum_g sqrt(kappa_g) gh[g][b][j]
//   gh = sqrt(kappa) j^-1 dN (j = dN x^T), nh = sqrt(lambda) J^-1 dN (J = dN X^T), kappa = w detJ 2c I3^-beta, lambda = w detJ 2c
//
// Why not the tensor-core tile kernel for this cell: 10 nodes and 4 Gauss points fill 39 % of the 16 x 16 x 4 DMMA tiles, and
// building the fragments, staging the tiles and the four CTA barriers cost 2372 warp instructions per element
// (profiles/r02_tet10_70cube.json: issue 43 %, 3.4 warps per scheduler).  Here lane c = 3 b + j owns COLUMN c of the element
// matrix: it keeps gh[.][b][0..2] and nh[.][b][j] in registers, walks the ten row nodes a, reads their vectors as shared-memory
// broadcasts and sends the three entries (3a+i, c), i = 0..2, straight from registers: a warp-wide reduction covers one matrix
// row, 30 lanes on 10 runs of 3 doubles.  No CTA barrier inside the element loop, no staging of K.
//
// The kernel is bound by shared-memory wavefronts (a warp-wide LDS costs its bytes x 32 lanes whether or not the lanes read the
// same word: 92 % LSU at 652 wavefronts per element in the first version, profiles/r02_tet10_warp_v1.json), so everything
// that does not depend on u comes from a per-element geometry cache instead: S0_ab = sum_g w detJ grad N_a . grad N_b and
// detJ_g (104 doubles per element, built once per mesh), which removes J, J^-1, nh and the s_ab sums from the kernel.
#pragma once
#include "fc_kernels_common.cuh"

namespace fc
{
  struct Tet10NhCfg
  {
    static constexpr int nen = 10, ngp = 4;
    static constexpr int warps = 8, threads = 32 * warps;
    static constexpr int epw = 4, epb = warps * epw;  // elements per warp and per CTA (launch slots are consecutive per warp)
    // geometry cache of an element (k_tet10_geom_cache, once per mesh): S0[a][b] = sum_g w_g detJ_g grad N_a . grad N_b
    // (evaluate_jacobian_mapping on the reference coordinates, 4C_solid_3D_ele_calc_lib.hpp:435-448), then detJ_g
    static constexpr int geom_doubles = 104;
    // per-warp shared memory (doubles)
    static constexpr int oX = 0;                   // [10][3] current coordinates relative to node 0
    static constexpr int oInv = 30;                // [4][9] sqrt(kappa) j^-1
    static constexpr int oSk = oInv + 36;          // [4] sqrt(kappa)
    static constexpr int oS = oSk + 4;             // [104] the element's geometry cache
    static constexpr int oG = oS + 104;            // [10][4][4] gh of node a at Gauss point g (3, one pad): 16-byte aligned
    static constexpr int warp_doubles = oG + 160;  // 334
  };

  // a0, a1 or a2 by two lane-constant predicates, as 32-bit selects (nvcc turns a ?: on doubles into branches and moves)
  __device__ __forceinline__ double select3(bool p0, bool p1, double a0, double a1, double a2)
  {
    const int lo = p0 ? __double2loint(a0) : (p1 ? __double2loint(a1) : __double2loint(a2));
    const int hi = p0 ? __double2hiint(a0) : (p1 ? __double2hiint(a1) : __double2hiint(a2));
    return __hiloint2double(hi, lo);
  }

  // X [n_col_nodes][3], conn [n_ele][10]; one thread per element
  static __global__ void k_tet10_geom_cache(const double* __restrict__ X, const int* __restrict__ conn, int n_ele, const double* __restrict__ dNtab,
      const double* __restrict__ gpw, double* __restrict__ geom)
  {
    constexpr int NEN = 10, NGP = 4;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_ele) return;
    double Xe[NEN][3];
    const int n0 = conn[(size_t)e * NEN];
    for (int a = 0; a < NEN; ++a)
      for (int i = 0; i < 3; ++i) Xe[a][i] = X[3 * (size_t)conn[(size_t)e * NEN + a] + i] - X[3 * (size_t)n0 + i];
    double* out = geom + (size_t)e * Tet10NhCfg::geom_doubles;
    for (int i = 0; i < 100; ++i) out[i] = 0.0;
    for (int g = 0; g < NGP; ++g)
    {
      const double* dN = dNtab + (size_t)g * NEN * 3;
      double M[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, Mi[9];
      for (int a = 0; a < NEN; ++a)
        for (int d = 0; d < 3; ++d)
          for (int i = 0; i < 3; ++i) M[3 * d + i] += dN[3 * a + d] * Xe[a][i];
      const double detJ = inv3(M, Mi);
      out[100 + g] = detJ;
      const double wd = gpw[g] * detJ;
      double N[NEN][3];
      for (int a = 0; a < NEN; ++a)
        for (int i = 0; i < 3; ++i) N[a][i] = Mi[3 * i] * dN[3 * a] + Mi[3 * i + 1] * dN[3 * a + 1] + Mi[3 * i + 2] * dN[3 * a + 2];
      for (int a = 0; a < NEN; ++a)
        for (int b = 0; b < NEN; ++b) out[10 * a + b] += wd * (N[a][0] * N[b][0] + N[a][1] * N[b][1] + N[a][2] * N[b][2]);
    }
  }

  template <bool WANT_K, bool HAS_DBC, bool ELE_OUT>
  __global__ void __launch_bounds__(Tet10NhCfg::threads, 3) k_tet10_nh(const EvalArgs A, const unsigned char* __restrict__ ele_bad)
  {
    using Cfg = Tet10NhCfg;
    constexpr int NEN = Cfg::nen, NGP = Cfg::ngp;
    __shared__ double s_dN[NGP][NEN][3];
    __shared__ __align__(16) double s_w[Cfg::warps][Cfg::warp_doubles];
    __shared__ long long s_rowbase[Cfg::warps][NEN];
    __shared__ int s_rowlen[Cfg::warps][NEN];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < NGP * NEN * 3; i += Cfg::threads) (&s_dN[0][0][0])[i] = __ldg(A.dNtab + i);
    __syncthreads();
    const double c2 = A.mat.young / (2.0 * (1.0 + A.mat.nue));  // 2c
    const double beta = A.mat.nue / (1.0 - 2.0 * A.mat.nue), beta2 = 2.0 * beta;
    double* W = s_w[warp];
    const bool col = lane < 30;              // lanes 30, 31 shadow lane 29 and send nothing
    const int b = col ? lane / 3 : 9, j = col ? lane - 3 * b : 2;
    const int slot0 = A.ele_offset + (blockIdx.x * Cfg::warps + warp) * Cfg::epw;

#pragma unroll 1
    for (int k = 0; k < Cfg::epw; ++k)
    {
      const int slot = slot0 + k;
      if (slot >= A.n_ele) break;  // warp-uniform
      // launch slot -> element: the caller's order, or the L2 sweep order of large meshes (fc_sweep_order.hpp)
      const int e = A.ele_perm ? __ldg(A.ele_perm + slot) : slot;
      // ---- node b: current coordinates relative to node 0, row bookkeeping, slots of column b in the ten rows; geometry cache ---
      const int node = __ldg(A.conn + (size_t)e * NEN + b), n0 = __ldg(A.conn + (size_t)e * NEN);
      const double* ge = A.geom + (size_t)e * Cfg::geom_doubles;
      double gv[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) gv[q] = lane + 32 * q < Cfg::geom_doubles ? __ldg(ge + lane + 32 * q) : 0.0;
      const double xr = (__ldg(A.X + 3 * (size_t)node + j) - __ldg(A.X + 3 * (size_t)n0 + j)) +
                        (__ldg(A.u + 3 * (size_t)node + j) - __ldg(A.u + 3 * (size_t)n0 + j));
      unsigned char sl[NEN];
      if (WANT_K && A.K)
      {
#pragma unroll
        for (int a = 0; a < NEN; ++a) sl[a] = __ldg(A.slots + ((size_t)e * NEN + a) * NEN + b);
      }
      __syncwarp();  // the previous element's readers are done
      W[Cfg::oX + 3 * b + j] = xr;
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (lane + 32 * q < Cfg::geom_doubles) W[Cfg::oS + lane + 32 * q] = gv[q];
      const bool owned = node < A.n_row_nodes;
      if (j == 0)
      {
        long long rb = -1;
        int rl = 0;
        if (WANT_K && A.K && owned)
        {
          rb = __ldg(A.node_rowbase + node);
          rl = __ldg(A.node_rowlen + node);
          if (HAS_DBC)
            for (int i = 0; i < 3; ++i) rl |= (A.dbc[3 * node + i] ? 1 : 0) << (28 + i);
        }
        s_rowbase[warp][b] = rb;
        s_rowlen[warp][b] = rl;
      }
      __syncwarp();

      // ---- current Jacobian j = dN x^T: lane (g, d) sums row d; the lanes with d == 0 invert --------------------------------
      bool bad = ele_bad && ele_bad[e];
      {
        const int g = lane >> 3, d = lane & 7;
        double m0 = 0.0, m1 = 0.0, m2 = 0.0;
        if (d < 3)
        {
#pragma unroll
          for (int a = 0; a < NEN; ++a)
          {
            const double dn = s_dN[g][a][d];
            m0 += dn * W[Cfg::oX + 3 * a];
            m1 += dn * W[Cfg::oX + 3 * a + 1];
            m2 += dn * W[Cfg::oX + 3 * a + 2];
          }
        }
        double M[9], Mi[9];
        M[0] = m0;
        M[1] = m1;
        M[2] = m2;
        M[3] = __shfl_down_sync(0xFFFFFFFFu, m0, 1);
        M[4] = __shfl_down_sync(0xFFFFFFFFu, m1, 1);
        M[5] = __shfl_down_sync(0xFFFFFFFFu, m2, 1);
        M[6] = __shfl_down_sync(0xFFFFFFFFu, m0, 2);
        M[7] = __shfl_down_sync(0xFFFFFFFFu, m1, 2);
        M[8] = __shfl_down_sync(0xFFFFFFFFu, m2, 2);
        if (d == 0)
        {
          const double detj = inv3(M, Mi), detJ = W[Cfg::oS + 100 + g];
          const double detF = detj / detJ;
          const double I3 = detF * detF;
          const double lambda = detJ * __ldg(A.gpw + g) * c2;
          const double kappa = lambda * exp(-beta * log(I3));
          if (!(detJ > 0.0) || !(I3 > 0.0) || !isfinite(kappa)) bad = true;
          const double sk = sqrt(kappa);
          double* R = W + Cfg::oInv + 9 * g;
#pragma unroll
          for (int i = 0; i < 9; ++i) R[i] = sk * Mi[i];
          W[Cfg::oSk + g] = sk;
        }
      }
      bad = __any_sync(0xFFFFFFFFu, bad);
      if (bad)
      {
        if (lane == 0) report_error(A.err, __ldg(A.ele_gid + e));
        continue;
      }
      __syncwarp();

      // ---- lane (b, j): component j of gh of node b at the four Gauss points; then all of node b's gh into registers ----------
      double gb[NGP][3], gbj[NGP], fpart = 0.0;
#pragma unroll
      for (int g = 0; g < NGP; ++g)
      {
        const double* R = W + Cfg::oInv + 9 * g + 3 * j;
        const double gh = R[0] * s_dN[g][b][0] + R[1] * s_dN[g][b][1] + R[2] * s_dN[g][b][2];
        gbj[g] = gh;
        W[Cfg::oG + (4 * b + g) * 4 + j] = gh;
        fpart -= W[Cfg::oSk + g] * gh;
      }
      __syncwarp();
      if (WANT_K)
      {
#pragma unroll
        for (int g = 0; g < NGP; ++g)
        {
          const double2 v = *reinterpret_cast<const double2*>(W + Cfg::oG + (4 * b + g) * 4);
          gb[g][0] = v.x;
          gb[g][1] = v.y;
          gb[g][2] = W[Cfg::oG + (4 * b + g) * 4 + 2];
        }
      }

      // ---- the ten row nodes ------------------------------------------------------------------------------------------------
#pragma unroll
      for (int a = 0; a < NEN; ++a)
      {
        const double s = c2 * W[Cfg::oS + 10 * a + b];
        fpart += s * W[Cfg::oX + 3 * a + j];  // s_ab = s_ba: lane (b, j) sums its own residual entry
        if (!WANT_K) continue;
        const double* Ga = W + Cfg::oG + 16 * a;
        double P[3] = {0.0, 0.0, 0.0}, Q[3] = {0.0, 0.0, 0.0};
#pragma unroll
        for (int g = 0; g < NGP; ++g)
        {
          const double2 ga01 = *reinterpret_cast<const double2*>(Ga + 4 * g);
          const double ga2 = Ga[4 * g + 2];
          const double gaj = select3(j == 0, j == 1, ga01.x, ga01.y, ga2);
          P[0] += ga01.x * gbj[g];
          P[1] += ga01.y * gbj[g];
          P[2] += ga2 * gbj[g];
          Q[0] += gaj * gb[g][0];
          Q[1] += gaj * gb[g][1];
          Q[2] += gaj * gb[g][2];
        }
        double Kc[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) Kc[i] = beta2 * P[i] + Q[i] + (i == j ? s : 0.0);
        if (ELE_OUT)
        {
          if (A.Ke_out && col)
          {
            double* Ke = A.Ke_out + (size_t)e * 900 + (size_t)30 * lane + 3 * a;  // column-major, column c = lane
            Ke[0] = Kc[0];
            Ke[1] = Kc[1];
            Ke[2] = Kc[2];
          }
        }
        if (A.K)
        {
          const long long rb = s_rowbase[warp][a];  // warp-uniform
          if (rb >= 0)
          {
            const int rl = s_rowlen[warp][a];
            double* p = A.K + rb + 3 * (int)sl[a] + j;
            if (HAS_DBC)
            {
#pragma unroll
              for (int i = 0; i < 3; ++i) red_add_if(col && !((rl >> (28 + i)) & 1), p + (long long)i * (rl & 0x0FFFFFFF), Kc[i]);
            }
            else
              red_add3_if(col, p, 8ll * rl, Kc[0], Kc[1], Kc[2]);
          }
        }
      }
      if (col)
      {
        if (ELE_OUT && A.fe_out) A.fe_out[(size_t)e * 30 + lane] = fpart;
        if (A.f && owned) red_add(A.f + 3 * (size_t)node + j, fpart);
      }
    }
  }
}  // namespace fc
