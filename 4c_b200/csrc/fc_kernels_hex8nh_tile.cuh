// Tile kernel: hex8, displacement-based, CoupNeoHooke, stiffness + force, assembled.
//
// Same closed form and FP64 tensor-core contraction as fc_kernels_hex8nh_mma.cuh, different scatter.  The
// per-element kernel is bound by the L2 reduction path: 576 FP64 reductions per element arrive as ~220 32-byte
// sectors (profiles/r01_mma_kernel_ncu_full_c.md), 3.4 x the sectors of the CSR values they end up in, because every
// entry receives 2.35 element contributions on average and because the 6-double runs of one element row straddle
// sectors.  Here a CTA owns a TILE of <= 16 elements (4 x 2 x 2 on a structured cube, fc_tile_plan.hpp), sums the
// element matrices of the tile in shared memory and adds every (row node, column node) block of the tile to the CSR
// values ONCE, in CSR order: 637 blocks instead of 1024 per tile and runs of 9..81 contiguous doubles.
//
// No shared-memory atomics: the 9 entries (i, j) of a block are split between three warps by transpose orbit,
//   warp w: (w, w), (p, r), (r, p)   with {p, r} = {0, 1, 2} \ {w}
// so every warp needs exactly three 8x8x8 products (W_ww, W_pr, W_rp: 6 DMMA, 18 per element as before), each warp
// walks ALL elements of the tile one after the other and the three warps never touch the same address.  The fourth
// warp computes the internal force meanwhile.
//
//   phase 0/1  lane (slot, t): gather node t / evaluate Gauss point t of the element in that slot -> fragments (smem)
//   phase 2    warps 0..2: for every element of the tile: 6 fragment loads, 6 DMMA, epilogue, 6 smem adds per lane
//              warp 3: internal force of every element (reductions into f)
//   phase 3    all threads: the tile's blocks, one reduction per double, lanes contiguous in CSR order
#pragma once
#include "fc_kernels_hex8nh_mma.cuh"

namespace fc
{
  struct TilePlanDev
  {
    const int* tile_ele;                // [n_tiles][16]
    const unsigned* tile_blk;           // [n_tiles][16][32]
    const long long* tile_ptr;          // [n_tiles + 1]
    const unsigned long long* rec;      // per block: CSR offset | rowlen << 44 | Dirichlet flags << 60
    int tile_begin;                     // first tile of this launch
    int n_tiles;                        // end of the tile range of this launch
  };

  struct Hex8TileCfg
  {
    static constexpr int warps = 4;
    static constexpr int threads = 32 * warps;
    static constexpr int epb = 16;  // element slots per tile
    // per element: 3 component planes of 64 (fragment layout), sqrt(kappa)[8], x rel [8][3]; odd stride: the 4
    // element groups of a warp store their fragments to disjoint banks
    static constexpr int oSk = 192, oX = 200, ele_doubles = 227;
    static constexpr int geom_doubles = 72;
    static constexpr int frag_doubles = epb * ele_doubles;  // 3632, even: the block planes behind it are 16-byte aligned
    static size_t smem_bytes(int max_blocks) { return sizeof(double) * ((size_t)frag_doubles + 9 * (size_t)max_blocks + 2); }
  };

  template <bool HAS_DBC>
  __global__ void __launch_bounds__(Hex8TileCfg::threads, 3) k_hex8_nh_tile(const EvalArgs A, const TilePlanDev P, const unsigned char* __restrict__ ele_bad)
  {
    using Cfg = Hex8TileCfg;
    extern __shared__ __align__(16) double smem_tile[];
    double* const smem = smem_tile;
    double* const planes = smem + Cfg::frag_doubles;
    __shared__ int s_node[Cfg::epb][8];            // owned row node or -1 (force scatter)
    __shared__ int s_ele[Cfg::epb];                // element index, -1 = slot empty or element failed
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int el = lane >> 3, t = lane & 7;  // phase 0/1 role
    const int slot = warp * 4 + el;
    const int tile = P.tile_begin + blockIdx.x;
    const int e = __ldg(P.tile_ele + (size_t)tile * Cfg::epb + slot);
    const bool active = e >= 0;
    const long long rec0 = __ldg(P.tile_ptr + tile);
    const int nb = (int)(__ldg(P.tile_ptr + tile + 1) - rec0);
    const int PS = 3 * nb;  // doubles per DOF-row plane
    double* E = smem + (size_t)slot * Cfg::ele_doubles;
    // block records of the tile: requested now, parked in registers, needed in phase 3
    unsigned long long rr[8];
#pragma unroll
    for (int m = 0; m < 8; ++m) rr[m] = (tid + Cfg::threads * m) < nb ? __ldg(P.rec + rec0 + tid + Cfg::threads * m) : 0ull;
    // geometry (s part, accumulator layout) of the first elements phase 2 will work on: requested before phase 0
    double2 sgr[8];
#pragma unroll
    for (int m = 0; m < 8; ++m)
    {
      const int em = __ldg(P.tile_ele + (size_t)tile * Cfg::epb + m);
      sgr[m] = (em >= 0 && (m < 4 || warp == 3)) ? __ldg(reinterpret_cast<const double2*>(A.geom + (size_t)em * Cfg::geom_doubles) + lane) : make_double2(0.0, 0.0);
    }

    // ---- phase 0 ----------------------------------------------------------------------------------
    double xr[3] = {0, 0, 0};
    {
      int node = 0;
      double X0[3] = {0, 0, 0};
      if (active)
      {
        node = __ldg(A.conn + (size_t)e * 8 + t);
        for (int i = 0; i < 3; ++i)
        {
          X0[i] = __ldg(A.X + 3 * (size_t)node + i);
          xr[i] = __ldg(A.u + 3 * (size_t)node + i);
        }
      }
      s_node[slot][t] = (active && node < A.n_row_nodes) ? node : -1;
      // coordinates relative to node 0 of the element (sum_a dN_a = 0, so j is unchanged)
      for (int i = 0; i < 3; ++i)
      {
        const double Xn = __shfl_sync(0xFFFFFFFFu, X0[i], lane & 24);
        const double un = __shfl_sync(0xFFFFFFFFu, xr[i], lane & 24);
        xr[i] = (X0[i] - Xn) + (xr[i] - un);
      }
    }
    // all 8 nodes of the element, exchanged through the (not yet written) fragment area of the element
    double2* const Ec = reinterpret_cast<double2*>(E + (slot & 1));
    Ec[2 * t + 0] = make_double2(xr[0], xr[1]);
    Ec[2 * t + 1] = make_double2(xr[2], 0.0);
    // zero the block planes of the tile
    for (int r = tid; 2 * r < 3 * PS; r += Cfg::threads) reinterpret_cast<double2*>(planes)[r] = make_double2(0.0, 0.0);
    __syncwarp();
    // ---- phase 1: Gauss point t of the element in this slot ------------------------------------------
    bool bad = false;
    {
      const double p = 0.57735026918962576451;
      const double xi0 = ((t ^ (t >> 1)) & 1) ? p : -p;
      const double xi1 = (t & 2) ? p : -p;
      const double xi2 = (t & 4) ? p : -p;
      const double detJ = active ? __ldg(A.geom + (size_t)e * Cfg::geom_doubles + 64 + t) : 1.0;
      double jm[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      double dN[24];
#pragma unroll
      for (int a = 0; a < 8; ++a)
      {
        const double2 c0 = Ec[2 * a], c1 = Ec[2 * a + 1];
        const double xa[3] = {c0.x, c0.y, c1.x};
        const double s0 = hex_node_pos(a, 0), s1 = hex_node_pos(a, 1), s2 = hex_node_pos(a, 2);
        const double r = 1.0 + s0 * xi0, s = 1.0 + s1 * xi1, q = 1.0 + s2 * xi2;
        dN[3 * a + 0] = 0.125 * s0 * s * q;
        dN[3 * a + 1] = 0.125 * s1 * q * r;
        dN[3 * a + 2] = 0.125 * s2 * r * s;
#pragma unroll
        for (int d = 0; d < 3; ++d)
#pragma unroll
          for (int i = 0; i < 3; ++i)
            jm[3 * d + i] += dN[3 * a + d] * xa[i];
      }
      double ji[9];
      const double detj = inv3(jm, ji);
      const double detF = detj / detJ;
      const double I3 = detF * detF;
      const double c2 = A.mat.young / (2.0 * (1.0 + A.mat.nue));  // 2c
      const double beta = A.mat.nue / (1.0 - 2.0 * A.mat.nue);
      const double lambda = detJ * c2;  // Gauss weight of hex_8point is 1
      const double kappa = lambda * exp(-beta * log(I3));
      bad = active && (!(detJ > 0.0) || !(I3 > 0.0) || !isfinite(kappa) || (ele_bad && ele_bad[e]));
      const double sk = sqrt(kappa);
      __syncwarp();  // every lane has read the coordinates: the fragment area may be overwritten
      if (active)
      {
#pragma unroll
        for (int a = 0; a < 8; ++a)
        {
          const double d0 = dN[3 * a], d1 = dN[3 * a + 1], d2 = dN[3 * a + 2];
          const int ix = frag_idx(t, a);
          E[0 * 64 + ix] = sk * (ji[0] * d0 + ji[1] * d1 + ji[2] * d2);
          E[1 * 64 + ix] = sk * (ji[3] * d0 + ji[4] * d1 + ji[5] * d2);
          E[2 * 64 + ix] = sk * (ji[6] * d0 + ji[7] * d1 + ji[8] * d2);
        }
        E[Cfg::oSk + t] = sk;
        for (int i = 0; i < 3; ++i) E[Cfg::oX + 3 * t + i] = xr[i];
      }
    }
    {
      const unsigned badmask = __ballot_sync(0xFFFFFFFFu, bad);
      const bool ebad = (badmask >> (lane & 24)) & 0xFFu;
      if (t == 0)
      {
        if (active && ebad) report_error(A.err, __ldg(A.ele_gid + e));
        s_ele[slot] = (active && !ebad) ? e : -1;
      }
    }
    __syncthreads();

    // ---- phase 2 ----------------------------------------------------------------------------------
    const int a = lane >> 2, q = lane & 3, b0 = 2 * q;
    const double c2 = A.mat.young / (2.0 * (1.0 + A.mat.nue));  // 2c
    const int fo0 = frag_idx(q, a), fo1 = frag_idx(q + 4, a);
    auto load_sg = [&](int k) -> double2 {
      const int ek = k < Cfg::epb ? s_ele[k] : -1;
      return ek >= 0 ? __ldg(reinterpret_cast<const double2*>(A.geom + (size_t)ek * Cfg::geom_doubles) + lane) : make_double2(0.0, 0.0);
    };
    if (warp < 3)
    {
      const double beta2 = 2.0 * A.mat.nue / (1.0 - 2.0 * A.mat.nue);
      // transpose orbit of this warp: diagonal (w, w) and the pair (p, r), (r, p)
      const int w = warp, p = warp == 0 ? 1 : 0, r = warp == 2 ? 1 : 2;
      double* const plane_w = planes + w * PS + w;  // entry (row a DOF w, column b DOF w) of block i sits at plane_w[3 i]
      double* const plane_p = planes + p * PS + r;  // (row DOF p, column DOF r)
      double* const plane_r = planes + r * PS + p;  // (row DOF r, column DOF p)
      const int ow = 64 * w, op = 64 * p, orr = 64 * r;
      // tile blocks of accumulator lane (a, q): (a, 2q) | (a, 2q+1) << 16, read one element ahead like the geometry
      const unsigned* const blk = P.tile_blk + (size_t)tile * Cfg::epb * 32 + lane;
      // geometry and block indices four elements ahead (global latency), fragments one element ahead
      unsigned bpr[4];
#pragma unroll
      for (int m = 0; m < 4; ++m) bpr[m] = __ldg(blk + 32 * m);
      double gw0 = smem[ow + fo0], gw1 = smem[ow + fo1], gp0 = smem[op + fo0], gp1 = smem[op + fo1], gr0 = smem[orr + fo0], gr1 = smem[orr + fo1];
#pragma unroll 1
      for (int k0 = 0; k0 < Cfg::epb; k0 += 4)
      {
#pragma unroll
        for (int m = 0; m < 4; ++m)
        {
          const int k = k0 + m;
          const double2 sg = sgr[m];
          const unsigned bp = bpr[m];
          const double a_w0 = gw0, a_w1 = gw1, a_p0 = gp0, a_p1 = gp1, a_r0 = gr0, a_r1 = gr1;
          if (k0 + 4 < Cfg::epb)
          {
            sgr[m] = load_sg(k + 4);
            bpr[m] = __ldg(blk + 32 * (k + 4));
          }
          if (k + 1 < Cfg::epb)
          {
            const double* G = smem + (size_t)(k + 1) * Cfg::ele_doubles;
            gw0 = G[ow + fo0], gw1 = G[ow + fo1];
            gp0 = G[op + fo0], gp1 = G[op + fo1];
            gr0 = G[orr + fo0], gr1 = G[orr + fo1];
          }
          if (s_ele[k] < 0) continue;  // warp-uniform
          double D0 = 0.0, D1 = 0.0, Q0 = 0.0, Q1 = 0.0, R0 = 0.0, R1 = 0.0;
          dmma884(D0, D1, a_w0, a_w0);
          dmma884(Q0, Q1, a_p0, a_r0);
          dmma884(R0, R1, a_r0, a_p0);
          dmma884(D0, D1, a_w1, a_w1);
          dmma884(Q0, Q1, a_p1, a_r1);
          dmma884(R0, R1, a_r1, a_p1);
          if ((bp & 0xFFFFu) != 0xFFFFu)  // row node a is an owned row
          {
            const int i0 = 3 * (int)(bp & 0xFFFFu), i1 = 3 * (int)(bp >> 16);
            // the six entries of a lane are distinct addresses: all loads first, then the adds and stores
            const double t0 = plane_w[i0], t1 = plane_w[i1], t2 = plane_p[i0], t3 = plane_p[i1], t4 = plane_r[i0], t5 = plane_r[i1];
            const double v0 = t0 + ((beta2 + 1.0) * D0 + c2 * sg.x), v1 = t1 + ((beta2 + 1.0) * D1 + c2 * sg.y);
            const double v2 = t2 + (beta2 * Q0 + R0), v3 = t3 + (beta2 * Q1 + R1);
            const double v4 = t4 + (beta2 * R0 + Q0), v5 = t5 + (beta2 * R1 + Q1);
            plane_w[i0] = v0;
            plane_w[i1] = v1;
            plane_p[i0] = v2;
            plane_p[i1] = v3;
            plane_r[i0] = v4;
            plane_r[i1] = v5;
          }
          __syncwarp();  // the next element of this warp may add to the same entries from other lanes
        }
      }
    }
    else
    {
      // internal force: f_a = sum_b s_ab x_b - sum_gp sqrt(kappa) ghat[gp][a]; lanes of the same row node a reduce
      // over their column pairs / Gauss points
#pragma unroll 1
      for (int k0 = 0; k0 < Cfg::epb; k0 += 8)
      {
#pragma unroll
        for (int m = 0; m < 8; ++m)
        {
          const int k = k0 + m;
          const double2 sg = sgr[m];
          if (k0 + 8 < Cfg::epb) sgr[m] = load_sg(k + 8);
          if (s_ele[k] < 0) continue;
          const double* G = smem + (size_t)k * Cfg::ele_doubles;
          const double* xk = G + Cfg::oX;
          const double sk0 = G[Cfg::oSk + q], sk1 = G[Cfg::oSk + 4 + q];
          const double s0 = c2 * sg.x, s1 = c2 * sg.y;
          double fpart[3];
#pragma unroll
          for (int i = 0; i < 3; ++i)
          {
            fpart[i] = s0 * xk[3 * b0 + i] + s1 * xk[3 * b0 + 3 + i] - (sk0 * G[i * 64 + fo0] + sk1 * G[i * 64 + fo1]);
            fpart[i] += __shfl_xor_sync(0xFFFFFFFFu, fpart[i], 1);
            fpart[i] += __shfl_xor_sync(0xFFFFFFFFu, fpart[i], 2);
          }
          const int na = s_node[k][a];
          if (q < 3 && A.f && na >= 0) red_add(A.f + 3 * (size_t)na + q, q == 0 ? fpart[0] : (q == 1 ? fpart[1] : fpart[2]));
        }
      }
    }
    __syncthreads();

    // ---- phase 3: every double of the tile once, lanes contiguous in CSR order ---------------------------
    // the fragment area is dead: it takes the block records
    unsigned long long* const srec = reinterpret_cast<unsigned long long*>(smem);
#pragma unroll
    for (int m = 0; m < 8; ++m)
      if (tid + Cfg::threads * m < nb) srec[tid + Cfg::threads * m] = rr[m];
    __syncthreads();
    if (A.K && !(A.debug & 1))
    {
#pragma unroll
      for (int i = 0; i < 3; ++i)
      {
        const double* pl = planes + i * PS;
#pragma unroll 4
        for (int x = tid; x < PS; x += Cfg::threads)
        {
          const int blk = x / 3, j = x - 3 * blk;
          const unsigned long long rc = srec[blk];
          if (HAS_DBC && ((rc >> (60 + i)) & 1ull)) continue;
          const long long off = (long long)(rc & 0xFFFFFFFFFFFull) + (long long)i * (long long)((rc >> 44) & 0xFFFFull) + j;
          red_add(A.K + off, pl[x]);
        }
      }
    }
  }

  // Dirichlet flags of the block records (bits 60..62 = DOF rows 3a..3a+2 are skipped by the matrix scatter)
  static __global__ void k_tile_set_dbc(unsigned long long* rec, const int* __restrict__ rownode, long long n, const unsigned char* __restrict__ dbc)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned long long r = rec[i] & ~(7ull << 60);
    if (dbc)
    {
      const int a = rownode[i];
      for (int d = 0; d < 3; ++d) r |= (unsigned long long)(dbc[3 * (size_t)a + d] ? 1 : 0) << (60 + d);
    }
    rec[i] = r;
  }
}  // namespace fc
