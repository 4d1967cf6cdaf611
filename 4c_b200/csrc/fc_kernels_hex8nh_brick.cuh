// Owner-computed assembly: hex8, displacement-based, CoupNeoHooke, struct_calc_nlnstiff (f and K).
//
// Same element arithmetic as k_hex8_neohooke_mma (closed form, 18 DMMA per element, geometry cache), but the sum over the
// elements of a node happens in SHARED MEMORY and every CSR row is written to HBM exactly once with plain stores:
//   * no zero-fill of K and f (the reference re-creates the matrix in SparseMatrix::zero before every evaluation,
//     src/core/linalg/src/sparse/4C_linalg_sparsematrix.cpp:353-378) and no read-modify-write of the value array:
//     DRAM traffic per evaluation = the K values once + the inputs, against 3x the K values for memset + reductions;
//   * no atomics: the result is bit-identical from run to run, like the reference's serial element loop
//     (src/core/fem/src/discretization/4C_fem_discretization_evaluate.cpp:83-102).
//
// A CTA owns one unit of the brick plan (fc_brick_plan.hpp): NL = 7 x 5 node lines, swept along the lattice axis in which
// the node numbering runs fastest.  Per step t it evaluates the <= 48 elements of layer t (between node planes t-1 and t):
//   A  all 384 threads: thread (k, g) gathers node g of element k and evaluates Gauss point g -> DMMA fragments in smem
//   C  four colour rounds; in a round the 12 warps take the 12 elements of one colour (elements of a colour share no
//      node): 18 DMMA, epilogue, then the 3x3 blocks of the rows this unit owns are added into the accumulator window
//      (2 planes x 35 nodes x 3 DOF rows x 81 entries, CSR row layout) with plain shared-memory read-modify-writes
//   D  node plane t-1 is complete: its rows leave the window as contiguous 1944-byte runs (coalesced stores, or
//      load-add-store when the caller did not ask for FC_EVAL_ZERO_FIRST: the += semantics of SparseMatrix::assemble)
// Elements on the faces of a unit are evaluated by every unit that owns one of their nodes (48 evaluations for 35
// nodes: 1.37x, plus one extra layer per chunk of planes) - FP64 work traded for HBM traffic and the L2 reduction path.
#pragma once
#include "fc_kernels_hex8nh_mma.cuh"

namespace fc
{
  struct BrickPlanDev
  {
    const int* unit_planes;
    const long long* unit_node_off;
    const long long* unit_ele_off;
    const int* node_tab;
    const int* ele_tab;
    const unsigned long long* rowidx_tab;
    int unit_begin;  // first unit of this launch
  };

  struct Hex8BrickCfg
  {
    static constexpr int NL = 35, NE = 48, warps = 12, threads = 32 * warps;
    static constexpr int row_doubles = 243;  // 3 DOF rows x 27 coupled nodes x 3
    static constexpr int oF = row_doubles;   // the 3 force entries of the node sit behind its rows
    static constexpr int acc_stride = 247;
    static constexpr int oSk = 192, oX = 200, frag_doubles = 225;  // per element: 3 fragment planes, sqrt(kappa)[8], x rel [8][3]
    static constexpr int geom_doubles = 72;
    static constexpr size_t smem_bytes = sizeof(double) * (2 * NL * acc_stride + (size_t)NE * frag_doubles) + sizeof(unsigned long long) * (NE * 8 + NE) +
                                         sizeof(long long) * 2 * NL + sizeof(int) * (3 * 2 * NL + 2 * NE);
  };

  template <bool STORE, bool HAS_DBC>
  __global__ void __launch_bounds__(Hex8BrickCfg::threads, 1) k_hex8_nh_brick(const EvalArgs A, const BrickPlanDev P, const unsigned char* __restrict__ ele_bad)
  {
    using Cfg = Hex8BrickCfg;
    constexpr int NL = Cfg::NL, NE = Cfg::NE, AS = Cfg::acc_stride, FD = Cfg::frag_doubles;
    extern __shared__ __align__(16) double smem_brick[];
    double* const acc = smem_brick;                   // [2][NL][AS]
    double* const frag = acc + 2 * NL * AS;           // [NE][FD]
    unsigned long long* const s_slots = reinterpret_cast<unsigned long long*>(frag + NE * FD);  // [NE][8] slot row of local node a
    unsigned long long* const s_rows = s_slots + NE * 8;                                      // [NE] accumulator row of every local node
    long long* const win_rb = reinterpret_cast<long long*>(s_rows + NE);                      // [2 NL] CSR offset of DOF row 3n
    int* const win_rl = reinterpret_cast<int*>(win_rb + 2 * NL);                              // [2 NL] entries per DOF row
    int* const win_node = win_rl + 2 * NL;                                                    // [2 NL] node or -1
    int* const win_dbc = win_node + 2 * NL;                                                   // [2 NL] Dirichlet bits of the 3 DOF rows
    int* const s_ele = win_dbc + 2 * NL;                                                      // [NE] element or -1
    int* const s_skip = s_ele + NE;                                                           // [NE] 1 = empty slot / failed element

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int unit = P.unit_begin + blockIdx.x;
    const int NP = __ldg(P.unit_planes + unit);
    const int* const nodes = P.node_tab + __ldg(P.unit_node_off + unit);
    const long long eoff = __ldg(P.unit_ele_off + unit);
    const int* const eles = P.ele_tab + eoff;
    const unsigned long long* const rows = P.rowidx_tab + eoff;

    const double c2 = A.mat.young / (2.0 * (1.0 + A.mat.nue));  // 2c
    const double beta = A.mat.nue / (1.0 - 2.0 * A.mat.nue);
    const double beta2 = 2.0 * beta;

    for (int i = tid; i < 2 * NL * AS; i += Cfg::threads) acc[i] = 0.0;

    for (int t = 0; t <= NP; ++t)
    {
      // ---- node plane t enters the window (its accumulator rows are zero) ----------------------------------------
      if (t < NP && tid < NL)
      {
        const int n = __ldg(nodes + (size_t)t * NL + tid), w = (t & 1) * NL + tid;
        win_node[w] = n;
        win_rl[w] = 0;
        if (n >= 0)
        {
          win_rb[w] = __ldg(A.node_rowbase + n);
          win_rl[w] = __ldg(A.node_rowlen + n);
          int fl = 0;
          if (HAS_DBC)
            for (int i = 0; i < 3; ++i) fl |= (A.dbc[3 * n + i] ? 1 : 0) << i;
          win_dbc[w] = fl;
        }
      }
      // ---- phase A: thread (k, g) = node g, then Gauss point g, of element slot k --------------------------------
      {
        const int k = tid >> 3, g = tid & 7;
        const int e = __ldg(eles + (size_t)t * NE + k);
        const bool active = e >= 0;
        double* const E = frag + (size_t)k * FD;
        double xr[3] = {0, 0, 0};
        int node = 0;
        if (active)
        {
          node = __ldg(A.conn + (size_t)e * 8 + g);
          for (int i = 0; i < 3; ++i) xr[i] = __ldg(A.X + 3 * (size_t)node + i) + __ldg(A.u + 3 * (size_t)node + i);
          s_slots[k * 8 + g] = __ldg(reinterpret_cast<const unsigned long long*>(A.slots + (size_t)e * 64) + g);
        }
        const unsigned owned_mask = __ballot_sync(0xFFFFFFFFu, active && node < A.n_row_nodes);
        // current coordinates relative to node 0 of the element (sum_a dN_a = 0: j is unchanged)
        for (int i = 0; i < 3; ++i) xr[i] -= __shfl_sync(0xFFFFFFFFu, xr[i], lane & 24);
        double2* const Ec = reinterpret_cast<double2*>(E + (k & 1));
        Ec[2 * g + 0] = make_double2(xr[0], xr[1]);
        Ec[2 * g + 1] = make_double2(xr[2], 0.0);
        __syncwarp();
        const double p = 0.57735026918962576451;
        const double xi0 = ((g ^ (g >> 1)) & 1) ? p : -p;
        const double xi1 = (g & 2) ? p : -p;
        const double xi2 = (g & 4) ? p : -p;
        const double detJ = active ? __ldg(A.geom + (size_t)e * Cfg::geom_doubles + 64 + g) : 1.0;
        double jm[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        double dN[24];
#pragma unroll
        for (int a = 0; a < 8; ++a)
        {
          const double2 q0 = Ec[2 * a], q1 = Ec[2 * a + 1];
          const double xa[3] = {q0.x, q0.y, q1.x};
          const double s0 = hex_node_pos(a, 0), s1 = hex_node_pos(a, 1), s2 = hex_node_pos(a, 2);
          const double r = 1.0 + s0 * xi0, s = 1.0 + s1 * xi1, q = 1.0 + s2 * xi2;
          dN[3 * a + 0] = 0.125 * s0 * s * q;
          dN[3 * a + 1] = 0.125 * s1 * q * r;
          dN[3 * a + 2] = 0.125 * s2 * r * s;
#pragma unroll
          for (int d = 0; d < 3; ++d)
#pragma unroll
            for (int i = 0; i < 3; ++i) jm[3 * d + i] += dN[3 * a + d] * xa[i];
        }
        double ji[9];
        const double detj = inv3(jm, ji);
        const double detF = detj / detJ;
        const double I3 = detF * detF;
        const double lambda = detJ * c2;  // Gauss weight of hex_8point is 1
        const double kappa = lambda * exp(-beta * log(I3));
        const bool bad = active && (!(detJ > 0.0) || !(I3 > 0.0) || !isfinite(kappa) || (ele_bad && ele_bad[e]));
        const double sk = sqrt(kappa);
        __syncwarp();  // every lane has read the coordinates: the fragment area may be overwritten
        if (active)
        {
#pragma unroll
          for (int a = 0; a < 8; ++a)
          {
            const double d0 = dN[3 * a], d1 = dN[3 * a + 1], d2 = dN[3 * a + 2];
            const int ix = frag_idx(g, a);
            E[0 * 64 + ix] = sk * (ji[0] * d0 + ji[1] * d1 + ji[2] * d2);
            E[1 * 64 + ix] = sk * (ji[3] * d0 + ji[4] * d1 + ji[5] * d2);
            E[2 * 64 + ix] = sk * (ji[6] * d0 + ji[7] * d1 + ji[8] * d2);
          }
          E[Cfg::oSk + g] = sk;
          for (int i = 0; i < 3; ++i) E[Cfg::oX + 3 * g + i] = xr[i];
        }
        const bool bad_el = (__ballot_sync(0xFFFFFFFFu, bad) >> (lane & 24)) & 0xFFu;
        if (g == 0)
        {
          const unsigned long long rw = active ? __ldg(rows + (size_t)t * NE + k) : ~0ull;
          s_ele[k] = e;
          s_rows[k] = rw;
          s_skip[k] = (!active || bad_el) ? 1 : 0;
          // a failing element is reported once: by the unit that owns its first owned node
          if (active && bad_el)
          {
            const unsigned om = (owned_mask >> (lane & 24)) & 0xFFu;
            if (om && ((rw >> (8 * (__ffs(om) - 1))) & 0xFF) != 0xFF) report_error(A.err, __ldg(A.ele_gid + e));
          }
        }
      }
      __syncthreads();

      // ---- phase C: four colour rounds, one element per warp ------------------------------------------------------
      const int a = lane >> 2, q = lane & 3, b0 = 2 * q;
#pragma unroll 1
      for (int r = 0; r < 4; ++r)
      {
        const int k = 12 * r + warp;
        if (!s_skip[k])
        {
          const int ek = s_ele[k];
          const double2 sg = __ldg(reinterpret_cast<const double2*>(A.geom + (size_t)ek * Cfg::geom_doubles) + lane);
          const double* G = frag + (size_t)k * FD;
          const int fo0 = frag_idx(q, a), fo1 = frag_idx(q + 4, a);
          double fg[3][2];
#pragma unroll
          for (int c = 0; c < 3; ++c)
          {
            fg[c][0] = G[c * 64 + fo0];
            fg[c][1] = G[c * 64 + fo1];
          }
          const int wr = (int)((s_rows[k] >> (8 * a)) & 0xFF);  // accumulator row of local node a, 0xFF = not ours
          double* const base = acc + (size_t)(wr == 0xFF ? 0 : wr) * AS;
          const double s0 = c2 * sg.x, s1 = c2 * sg.y;
          {
            const double* xk = G + Cfg::oX;
            const double sk0 = G[Cfg::oSk + q], sk1 = G[Cfg::oSk + 4 + q];
            double fpart[3];
#pragma unroll
            for (int i = 0; i < 3; ++i)
            {
              fpart[i] = s0 * xk[3 * b0 + i] + s1 * xk[3 * b0 + 3 + i] - (sk0 * fg[i][0] + sk1 * fg[i][1]);
              fpart[i] += __shfl_xor_sync(0xFFFFFFFFu, fpart[i], 1);
              fpart[i] += __shfl_xor_sync(0xFFFFFFFFu, fpart[i], 2);
            }
            if (q < 3 && wr != 0xFF) base[Cfg::oF + q] += q == 0 ? fpart[0] : (q == 1 ? fpart[1] : fpart[2]);
          }
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
          if (wr != 0xFF)
          {
            const unsigned long long sl = s_slots[k * 8 + a];
            const int RL = win_rl[wr];
            double* const p0 = base + 3 * (int)((sl >> (8 * b0)) & 0xFF);
            double* const p1 = base + 3 * (int)((sl >> (8 * b0 + 8)) & 0xFF);
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
              for (int j = 0; j < 3; ++j)
              {
                p0[i * RL + j] += beta2 * W[i][j][0] + W[j][i][0] + (i == j ? s0 : 0.0);
                p1[i * RL + j] += beta2 * W[i][j][1] + W[j][i][1] + (i == j ? s1 : 0.0);
              }
          }
        }
        __syncthreads();
      }

      // ---- phase D: node plane t-1 is complete ---------------------------------------------------------------------
      if (t >= 1)
      {
        const int par = (t - 1) & 1;
        for (int l = warp; l < NL; l += Cfg::warps)
        {
          const int w = par * NL + l;
          const int n = win_node[w];
          if (n < 0) continue;
          double* const src = acc + (size_t)w * AS;
          const int RL = win_rl[w], len = 3 * RL;
          double* const dst = A.K + win_rb[w];
          const int fl = HAS_DBC ? win_dbc[w] : 0;
          for (int i = lane; i < len; i += 32)
          {
            double v = src[i];
            src[i] = 0.0;
            if (HAS_DBC && fl)
            {
              // Dirichlet rows are skipped by the matrix scatter (4C_linalg_sparsematrix.cpp:480): zero after
              // SparseMatrix::zero, untouched otherwise
              const int row = i / RL;
              if ((fl >> row) & 1)
              {
                if (!STORE) continue;
                v = 0.0;
              }
            }
            if (STORE)
              dst[i] = v;
            else
              dst[i] += v;
          }
          if (lane < 3)
          {
            const double v = src[Cfg::oF + lane];
            src[Cfg::oF + lane] = 0.0;
            if (A.f)
            {
              if (STORE)
                A.f[3 * (size_t)n + lane] = v;
              else
                A.f[3 * (size_t)n + lane] += v;
            }
          }
        }
      }
      __syncthreads();
    }
  }
}  // namespace fc
