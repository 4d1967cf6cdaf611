// Specialised element kernel, tensor-core variant: hex8, displacement-based, CoupNeoHooke.
//
// Same closed form as fc_kernels_hex8nh.cuh,
//   K_ab[i][j] = 2 beta W_ij[a][b] + W_ji[a][b] + delta_ij s[a][b]
//   W_ij[a][b] = sum_gp ghat[gp][a][i] ghat[gp][b][j]      ghat = sqrt(kappa) g
//   s[a][b]    = sum_gp sum_i nhat[gp][a][i] nhat[gp][b][i] nhat = sqrt(lambda) grad N
//   f_a        = sum_b s[a][b] x_b - sum_gp sqrt(kappa) ghat[gp][a]
// but the contractions over Gauss points are dense 8x8x8 products per component pair (i,j): nine
// for W, three accumulated into s.  They run on the FP64 tensor cores (mma.sync m8n8k4 f64, DMMA):
// one warp multiplies one element at a time, A and B fragments are the same registers, and the
// accumulator layout leaves every lane holding the two complete 3x3 blocks K_{a,b0}, K_{a,b0+1}
// (a = lane/4, b0 = 2 (lane%4)), so the epilogue and the CSR scatter need no data exchange.
//
// Work decomposition: a warp owns 4 elements.
//   phase 0  lane (el, t) gathers node t of element el (coordinates relative to node 0)
//   phase 1  lane (el, g) evaluates Gauss point g of element el -> ghat, nhat in fragment layout (smem)
//   phase 2  for each of the 4 elements: 6 fragment loads, 18 DMMA (none for the force-only action)
//   phase 3  epilogue + fire-and-forget FP64 reductions into the CSR rows / residual
//
// Geometry cache: s[a][b] / (2c) = sum_gp w detJ grad N_a . grad N_b and detJ at the Gauss points depend on the
// reference coordinates only.  The reference recomputes them in every Newton iteration
// (4C_solid_3D_ele_calc_lib.hpp:435-448); here k_hex8_geom_cache stores them once per mesh (72 doubles per element,
// in the accumulator layout of phase 2) and the evaluation reads them back: 6 DMMA, the material Jacobian, its
// inverse, three fragment planes and 42 % of the shared memory per element less.
#pragma once
#include "fc_kernels_common.cuh"

namespace fc
{
  struct Hex8MmaCfg
  {
    static constexpr int warps = 4;
    static constexpr int threads = 32 * warps;
    static constexpr int epw = 4;                       // elements per warp
    static constexpr int epb = warps * epw;             // elements per CTA
    // per element: 3 component planes of 64 (fragment layout), sqrt(kappa)[8], x rel [8][3]; the K_e staging of the
    // scatter (12 rows x 24) overlays the element's area and sets its size
    // (odd stride 291: the 4 element groups of a warp store their fragments to disjoint banks, and with the parity
    // shift below the 16-byte coordinate loads of the 4 groups fall into 4 different bank groups)
    static constexpr int oSk = 192, oX = 200, ele_doubles = 12 * 24 + 3;
    static constexpr int geom_doubles = 72;  // geometry cache per element: G[8][8], detJ[8]
    static constexpr int ks_stride = 24;  // K_e staging for the scatter: 12 rows x 24 inside the element's own area
    static constexpr size_t smem_bytes = sizeof(double) * (size_t)epb * ele_doubles;
  };

  __device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b)
  {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
  }

  // position of (gp, node) inside a 64-entry component plane: one k-step (4 Gauss points) of one
  // half of the nodes is a contiguous, bank-conflict-free 128-byte line for the fragment loads
  // (the node field is XOR-swizzled with the k-step so that the 8 Gauss-point lanes of phase 1 store conflict-free)
  __device__ __forceinline__ int frag_idx(int gp, int node) { return 32 * (gp >> 2) + 16 * (node >> 2) + 4 * (gp & 3) + ((node & 3) ^ (2 * (gp >> 2))); }

  // HAS_PERM: launch slot -> element through EvalArgs::ele_perm (L2 sweep order of large meshes, interior-first order of a
  // partitioned mesh).  Without it the element is the slot and its index is recomputed instead of kept: measured on one box at C2,
  // 1.430 ms against 1.466 ms for the variant that carries the indirection (2 more spilled registers at the 96-register limit).
  // Occupancy: 4 CTAs per SM for the stiffness kernel (123 registers, nothing spilled; with 5 CTAs the 96-register limit spilled 8
  // registers: same box at C2 1.404 -> 1.366 ms), 5 for the force-only kernel (0.58 vs 0.62 ms).
  template <bool WANT_K, bool HAS_DBC, bool HAS_PERM>
  __global__ void __launch_bounds__(Hex8MmaCfg::threads, WANT_K ? 4 : 5) k_hex8_neohooke_mma(const EvalArgs A, const unsigned char* __restrict__ ele_bad)
  {
    using Cfg = Hex8MmaCfg;
    extern __shared__ __align__(16) double smem_mma[];
    double* const smem = smem_mma;
    // scatter record of node t: x = CSR offset of DOF row 3n (bits 0..43) | entries per DOF row (bits 44..59) |
    // Dirichlet flags of the 3 DOF rows (bits 60..62), all ones = not an owned row; y = slot row (position of
    // every b in the CSR row of this node)
    __shared__ ulonglong2 s_info[Cfg::epb][8];
    __shared__ int s_node[Cfg::epb][8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int el = lane >> 3, t = lane & 7;   // phase 0/1 role
    const int wel = warp * Cfg::epw;          // first element slot of this warp inside the CTA
    const int slot_e = A.ele_offset + blockIdx.x * Cfg::epb + wel + el;
    const bool active = slot_e < A.n_ele;
    // launch slot -> element: the caller's order, or the L2 sweep order of large meshes (fc_sweep_order.hpp)
    const int e = active ? (HAS_PERM ? __ldg(A.ele_perm + slot_e) : slot_e) : -1;
    double* E = smem + (size_t)(wel + el) * Cfg::ele_doubles;

    // ---- phase 0 ----------------------------------------------------------------------------------
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
    unsigned fastmask;  // bit 8k: all 8 nodes of element k are owned rows without Dirichlet flags
    {
      const bool owned = active && node < A.n_row_nodes;
      ulonglong2 inf = make_ulonglong2(~0ull, 0ull);
      bool plain = false;
      if (WANT_K && owned && A.K)
      {
        const unsigned long long rb = (unsigned long long)__ldg(A.node_rowbase + node);
        const unsigned long long rl = (unsigned long long)__ldg(A.node_rowlen + node);
        inf.y = __ldg(reinterpret_cast<const unsigned long long*>(A.slots + (size_t)e * 64) + t);
        unsigned long long fl = 0;
        if (HAS_DBC)
          for (int i = 0; i < 3; ++i) fl |= (unsigned long long)(A.dbc[3 * node + i] ? 1 : 0) << (60 + i);
        inf.x = rb | (rl << 44) | fl;
        plain = fl == 0;
      }
      s_info[wel + el][t] = inf;
      s_node[wel + el][t] = owned ? node : -1;
      const unsigned pm = __ballot_sync(0xFFFFFFFFu, plain);
      fastmask = pm & (pm >> 1);
      fastmask &= fastmask >> 2;
      fastmask &= fastmask >> 4;
    }
    // coordinates relative to node 0 of the element (sum_a dN_a = 0, so J and F are unchanged)
    for (int i = 0; i < 3; ++i)
    {
      const double X0 = __shfl_sync(0xFFFFFFFFu, Xr[i], lane & 24);
      const double u0 = __shfl_sync(0xFFFFFFFFu, xr[i], lane & 24);
      Xr[i] -= X0;
      xr[i] = Xr[i] + (xr[i] - u0);
    }
    // phase 1 needs all 8 nodes of the element: exchanged through the (not yet written) fragment area of the
    // element, 32 bytes per node (current coordinates only), read back as broadcast 16-byte loads (a register shuffle costs the LSU 1.5
    // wavefronts per 32-bit half: 40 wavefronts per element instead of 9)
    double2* const Ec = reinterpret_cast<double2*>(E + ((wel + el) & 1));
    Ec[2 * t + 0] = make_double2(xr[0], xr[1]);
    Ec[2 * t + 1] = make_double2(xr[2], 0.0);
    __syncwarp();
    // ---- phase 1: Gauss point t of element el ----------------------------------------------------
    bool bad = false;
    {
      const double p = 0.57735026918962576451;
      // Gauss point t sits in the octant of node t: signs (x, y, z) of hex8 node t = (bit0 ^ bit1, bit1, bit2)
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
    const unsigned badmask = __ballot_sync(0xFFFFFFFFu, bad);
    const unsigned actmask = __ballot_sync(0xFFFFFFFFu, active);
    if (active && t == 0 && ((badmask >> (lane & 24)) & 0xFFu)) report_error(A.err, __ldg(A.ele_gid + e));
    __syncwarp();

    // ---- phase 2 + 3: one element at a time, whole warp ------------------------------------------
    const int a = lane >> 2, q = lane & 3, b0 = 2 * q;
    const double beta2 = 2.0 * A.mat.nue / (1.0 - 2.0 * A.mat.nue);
    const double c2 = A.mat.young / (2.0 * (1.0 + A.mat.nue));  // 2c
    // cached geometric part of s[a][2q], s[a][2q+1], one element ahead of its use
    auto load_sg = [&](int k) -> double2 {
      int ek;
      if (HAS_PERM)
        ek = __shfl_sync(0xFFFFFFFFu, e, (8 * k) & 31);
      else
      {
        ek = A.ele_offset + blockIdx.x * Cfg::epb + wel + k;
        if (ek >= A.n_ele) ek = -1;
      }
      return (k < Cfg::epw && ek >= 0) ? __ldg(reinterpret_cast<const double2*>(A.geom + (size_t)ek * Cfg::geom_doubles) + lane) : make_double2(0.0, 0.0);
    };
    double2 sg_next = load_sg(0);
#pragma unroll 1
    for (int k = 0; k < Cfg::epw; ++k)
    {
      const double2 sg = sg_next;
      sg_next = load_sg(k + 1);
      if (!((actmask >> (8 * k)) & 1u) || ((badmask >> (8 * k)) & 0xFFu)) continue;  // warp-uniform
      const int ek = HAS_PERM ? __shfl_sync(0xFFFFFFFFu, e, 8 * k) : A.ele_offset + blockIdx.x * Cfg::epb + wel + k;
      const double* G = smem + (size_t)(wel + k) * Cfg::ele_doubles;
      const int fo0 = frag_idx(q, a), fo1 = frag_idx(q + 4, a);
      double fg[3][2];
#pragma unroll
      for (int c = 0; c < 3; ++c)
      {
        fg[c][0] = G[c * 64 + fo0];
        fg[c][1] = G[c * 64 + fo1];
      }
      const double s0 = c2 * sg.x, s1 = c2 * sg.y;
      // internal force: lanes of the same row node a reduce over their column pairs / Gauss points
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
        const int na = s_node[wel + k][a];
        if (q < 3)
        {
          const double v = q == 0 ? fpart[0] : (q == 1 ? fpart[1] : fpart[2]);
          if (A.fe_out) A.fe_out[(size_t)ek * 24 + 3 * a + q] = v;
          if (A.f && na >= 0) red_add(A.f + 3 * (size_t)na + q, v);
        }
      }
      if (!WANT_K) continue;
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
      double Kb[2][9];
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
        {
          Kb[0][3 * i + j] = beta2 * W[i][j][0] + W[j][i][0] + (i == j ? s0 : 0.0);
          Kb[1][3 * i + j] = beta2 * W[i][j][1] + W[j][i][1] + (i == j ? s1 : 0.0);
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
        // stage K_e row-major in shared memory, then let 3 adjacent lanes add the 3 contiguous doubles
        // of one CSR node block in ONE instruction: the memory system processes reductions per 32-byte
        // request, so lane-contiguous triples run 2.3x faster than one scattered double per lane
        // (profiles/r01_red_patterns.md)
        // The staging buffer is the element's own fragment area (already consumed into registers),
        // filled in two halves of 4 row nodes: no extra shared memory, 4 CTAs per SM.
        // (+1 double for odd elements: ele_doubles is odd, the 16-byte staging stores need an even offset)
        double* Ks = smem + (size_t)(wel + k) * Cfg::ele_doubles + ((wel + k) & 1);
        const bool fast = (fastmask >> (8 * k)) & 1u;  // warp-uniform
        const int b = lane / 3, jj = lane - 3 * b;     // scatter role of lanes 0..23: column node b, component jj
#pragma unroll
        for (int half = 0; half < 2; ++half)
        {
          __syncwarp();
          if ((a >> 2) == half)
          {
            // 6 contiguous doubles per row as three 16-byte stores: conflict-free for row stride 24
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
          // ptxas turns every predicated reduction into its own branch region (BSSY/BRA/REDG/BSYNC), so the lanes
          // that scatter sit inside ONE divergent region per half and the reductions of an element whose 8 nodes
          // are all owned rows without Dirichlet flags (the common case) carry no predicate at all
          if (lane < 24)
          {
            double v[12];
            ulonglong2 inf[4];
#pragma unroll
            for (int r = 0; r < 4; ++r)
            {
              inf[r] = s_info[wel + k][4 * half + r];
#pragma unroll
              for (int i = 0; i < 3; ++i) v[3 * r + i] = Ks[(3 * r + i) * Cfg::ks_stride + lane];
            }
            if (fast)
            {
#pragma unroll
              for (int r = 0; r < 4; ++r)
              {
                const long long off = (long long)(inf[r].x & 0xFFFFFFFFFFFull) + 3 * (int)((inf[r].y >> (8 * b)) & 0xFF) + jj;
                red_add3(A.K + off, (long long)((inf[r].x >> 44) & 0xFFFF) * 8, v[3 * r], v[3 * r + 1], v[3 * r + 2]);
              }
            }
            else
            {
#pragma unroll
              for (int r = 0; r < 4; ++r)
              {
                if ((long long)inf[r].x < 0) continue;  // not an owned row (warp-uniform)
                const long long off = (long long)(inf[r].x & 0xFFFFFFFFFFFull) + 3 * (int)((inf[r].y >> (8 * b)) & 0xFF) + jj;
                const long long rl = (long long)((inf[r].x >> 44) & 0xFFFF);
#pragma unroll
                for (int i = 0; i < 3; ++i)
                  if (!((inf[r].x >> (60 + i)) & 1)) red_add(A.K + off + i * rl, v[3 * r + i]);
              }
            }
          }
        }
      }
    }
  }
}  // namespace fc
