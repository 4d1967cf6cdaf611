// Tensor-core closed-form kernel for displacement-based CoupNeoHooke on any cell type of the path
// (tet10: BASELINE C3b, hex27: C3a).  Same closed form as fc_kernels_hex8nh_mma.cuh,
//   K_ab[i][j] = 2 beta W_ij[a][b] + W_ji[a][b] + delta_ij s[a][b]
//   W_ij[a][b] = sum_gp ghat[gp][a][i] ghat[gp][b][j],  ghat = sqrt(kappa) j^-1 dN   (j = dN x^T, current Jacobian)
//   s[a][b]    = sum_gp nhat[gp][a] . nhat[gp][b],       nhat = sqrt(lambda) J^-1 dN  (J = dN X^T)
//   f_a        = sum_b s[a][b] x_b - sum_gp sqrt(kappa) ghat[gp][a]
//   kappa = w detJ 2c I3^-beta, lambda = w detJ 2c, I3 = (det j / det J)^2
// (F = I + u N_XYZ^T of the non-hex8 cells, 4C_solid_3D_ele_calc_lib.hpp:585-602, equals x N_XYZ^T for an
// isoparametric element, so g = F^-T grad N = j^-1 dN for every cell type), with the node dimension padded to
// MT tiles of 8 and the Gauss-point dimension to KS k-steps of 4 of mma.sync.m8n8k4.f64:
//   tet10: 2x2 tiles, 1 k-step, 48 DMMA per element      hex27: 4x4 tiles, 7 k-steps, 1344 DMMA per element
// instead of the 66 kFLOP / 2.7 MFLOP B^T C B of the reference formulation on FP64 FMA.
//
// A CTA (4 warps) works on EPB elements together:
//   phase 0   thread (el, a): gather node a (coordinates relative to node 0), CSR row bookkeeping
//   phase 1a  thread (el, gp, which): J or j at the Gauss point, inverse + determinant
//   phase 1b  thread (el, gp): kappa, lambda, error flags; inverses scaled by sqrt(kappa), sqrt(lambda)
//   phase 1c  warp per (el, k-step, node tile): the 32 entries of every component plane in fragment order
//   phase 2   warp per (el, row tile mt): for every column tile nt: 12 KS DMMA, epilogue, K block staged in
//             shared memory and added to the CSR rows with lane-contiguous FP64 reductions; then the residual rows
#pragma once
#include "fc_kernels_hex8nh_mma.cuh"
#include "fc_nurbs.cuh"

namespace fc
{
  template <int CT>
  struct NhGenCfg
  {
    static constexpr int nen = Cell<CT>::nen, ngp = Cell<CT>::ngp;
    static constexpr int MT = (nen + 7) / 8, KS = (ngp + 3) / 4, NP = 8 * MT, GP4 = 4 * KS;
    static constexpr bool is27 = CT == HEX27 || CT == NURBS27;  // nurbs27: the same kernel on the element's own shape tables (fc_nurbs.cuh)
    static constexpr int epb = CT == TET10 ? 8 : (is27 ? 1 : 4);
    static constexpr int warps = 4, threads = 32 * warps;
    // CTAs per SM the register allocation leaves room for.  hex27: 5 (96 registers, no spills) measured 4.01 ms at 48^3 against
    // 4.38 ms with 4 (113 registers) and 4.58 ms with 6 (80 registers, spills)
    static constexpr int min_blocks = is27 ? 5 : 1;
    static constexpr int plane = KS * MT * 32;
    // shape-function derivatives at the Gauss points behind the [ngp + nen + 1][nen][3] table (fill_tables, fc_host_util.hpp)
    static constexpr int oTabNodeMajor = (ngp + nen + 1) * nen * 3;   // [nen][ngp][3]
    static constexpr int oTabFragment = oTabNodeMajor + nen * ngp * 3;  // [3][KS * MT][32]
    // hex27 reads s_ab / 2c = sum_gp w detJ grad N_a . grad N_b and detJ from a geometry cache (k_gen_geom_cache, once per
    // mesh, like the hex8 kernel): the three nhat planes, J^-1 and 3 of the 12 DMMA per tile step go away, the element needs
    // 22 KB of shared memory less.  tet10 recomputes them (measured with the cache: stiffness 9.88 -> 10.10 ms, residual-only
    // 3.20 -> 2.86 ms; its cache adds 28 % to the bytes the kernel moves).
    static constexpr bool geom = is27;
    static constexpr int nplanes = geom ? 3 : 6;
    static constexpr int geom_doubles = 2 * 32 * MT * MT + ((ngp + 7) & ~7);  // [mt][nt][lane][2] accumulator layout, then detJ[gp]
    static constexpr int oInv = nplanes * plane;     // [GP4][20]: sk*j^-1 (9), sl*J^-1 (9) | scratch: inv, det
    static constexpr int oX = oInv + GP4 * 20;       // [NP][3] current coordinates rel. node 0, zero padded
    static constexpr int oXm = oX + 3 * NP;          // [nen][3] material coordinates rel. node 0
    static constexpr int ele_doubles = (oXm + 3 * nen + 1) & ~1;
    static constexpr int stage = 288;                // per warp: 12 rows x 24
    static constexpr size_t smem_bytes = sizeof(double) * ((size_t)epb * ele_doubles + warps * stage);
  };

  static_assert(NhGenCfg<NURBS27>::oTabNodeMajor == NurbsTab::oNodeMajor && NhGenCfg<NURBS27>::oTabFragment == NurbsTab::oFragment &&
                    NhGenCfg<NURBS27>::MT == NurbsTab::MT && NhGenCfg<NURBS27>::KS == NurbsTab::KS,
      "per-element nurbs27 table and the shared table of the Lagrange cells: same layout");

  // geometry cache of k_nh_mma_gen (cells with NhGenCfg::geom): per element G[a][b] = sum_gp w detJ grad N_a . grad N_b in the
  // accumulator layout of the (mt, nt) tiles - lane (a8, q) holds (8 mt + a8, 8 nt + 2 q + {0, 1}) - followed by detJ at the Gauss
  // points.  evaluate_jacobian_mapping (4C_solid_3D_ele_calc_lib.hpp:435-448) on the reference coordinates: constant for a mesh.
  template <int CT>
  __global__ void __launch_bounds__(128) k_gen_geom_cache(const double* __restrict__ X, const int* __restrict__ conn, int n_ele,
      const double* __restrict__ dNtab, const double* __restrict__ gpw, double* __restrict__ geom)
  {
    using Cfg = NhGenCfg<CT>;
    constexpr int NEN = Cfg::nen, NGP = Cfg::ngp, MT = Cfg::MT;
    __shared__ double sX[NEN][3];
    __shared__ double sN[NGP][NEN][3];  // sqrt(w detJ) J^-1 dN_a
    const int e = blockIdx.x, tid = threadIdx.x;
    if (e >= n_ele) return;
    double* out = geom + (size_t)e * Cfg::geom_doubles;
    for (int i = tid; i < 3 * NEN; i += blockDim.x)
    {
      const int a = i / 3, d = i % 3;
      sX[a][d] = X[3 * (size_t)conn[(size_t)e * NEN + a] + d] - X[3 * (size_t)conn[(size_t)e * NEN] + d];
    }
    __syncthreads();
    for (int gp = tid; gp < NGP; gp += blockDim.x)
    {
      const double* dN = ele_table<CT>(dNtab, e) + (size_t)gp * NEN * 3;
      double M[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, Mi[9];
      for (int a = 0; a < NEN; ++a)
        for (int d = 0; d < 3; ++d)
          for (int i = 0; i < 3; ++i) M[3 * d + i] += dN[3 * a + d] * sX[a][i];
      const double detJ = inv3(M, Mi);
      const double wd = gpw[gp] * detJ, sd = sqrt(wd > 0.0 ? wd : 0.0);
      for (int a = 0; a < NEN; ++a)
        for (int i = 0; i < 3; ++i) sN[gp][a][i] = sd * (Mi[3 * i] * dN[3 * a] + Mi[3 * i + 1] * dN[3 * a + 1] + Mi[3 * i + 2] * dN[3 * a + 2]);
      out[2 * 32 * MT * MT + gp] = detJ;
    }
    __syncthreads();
    for (int idx = tid; idx < 2 * 32 * MT * MT; idx += blockDim.x)
    {
      const int c = idx & 1, ln = (idx >> 1) & 31, tile = idx >> 6;
      const int a = 8 * (tile / MT) + (ln >> 2), b = 8 * (tile % MT) + 2 * (ln & 3) + c;
      double g = 0.0;
      if (a < NEN && b < NEN)
        for (int gp = 0; gp < NGP; ++gp)
          for (int d = 0; d < 3; ++d) g += sN[gp][a][d] * sN[gp][b][d];
      out[idx] = g;
    }
  }

  template <int CT, bool WANT_K, bool HAS_DBC>
  __global__ void __launch_bounds__(NhGenCfg<CT>::threads, NhGenCfg<CT>::min_blocks) k_nh_mma_gen(const EvalArgs A, const unsigned char* __restrict__ ele_bad)
  {
    using Cfg = NhGenCfg<CT>;
    constexpr int NEN = Cfg::nen, NGP = Cfg::ngp, MT = Cfg::MT, KS = Cfg::KS, NP = Cfg::NP, EPB = Cfg::epb;
    extern __shared__ __align__(16) double smem_nhg[];
    __shared__ long long s_rowbase[EPB][NP];
    __shared__ int s_rowlen[EPB][NP];
    __shared__ int s_node[EPB][NP];
    __shared__ int s_bad[EPB];
    // slot of column node b in the row of node a.  hex27: the whole element map is loaded once in phase 0 (measured 5.77 -> 5.03 ms
    // at 48^3: inside the tile loop the byte loads' latency was exposed 4 times per element); tet10: per row tile inside the loop
    // (phase-0 loading measured slower there, 9.9 -> 10.4 ms)
    constexpr bool SLOTS_PHASE0 = Cfg::is27;
    __shared__ unsigned char s_slots[SLOTS_PHASE0 ? 1 : Cfg::warps * 8][NP];
    // phase-0 variant: the element's nen x nen bytes as aligned 32-bit words (the byte-wise loop cost 9 000 cycles per element)
    constexpr int SLW = (NEN * NEN + 6) / 4 + 1;
    __shared__ unsigned int s_slw[SLOTS_PHASE0 ? EPB : 1][SLOTS_PHASE0 ? SLW : 1];
    __shared__ int s_slshift[EPB];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ebase = A.ele_offset + blockIdx.x * EPB;  // launch slots [ele_offset, n_ele)
    const double c2 = A.mat.young / (2.0 * (1.0 + A.mat.nue));  // 2c
    const double beta = A.mat.nue / (1.0 - 2.0 * A.mat.nue);

    // ---- phase 0 -----------------------------------------------------------------------------------------
    for (int it = tid; it < EPB * NP; it += Cfg::threads)
    {
      const int el = it / NP, a = it % NP;
      // launch slot -> element: the caller's order, or the L2 sweep order of large meshes (fc_sweep_order.hpp)
      const int e = ebase + el < A.n_ele ? (A.ele_perm ? __ldg(A.ele_perm + ebase + el) : ebase + el) : -1;
      double* E = smem_nhg + (size_t)el * Cfg::ele_doubles;
      long long rb = -1;
      int rl = 0, nd = -1;
      double xc[3] = {0, 0, 0};
      if (e >= 0 && a < NEN)
      {
        const int node = __ldg(A.conn + (size_t)e * NEN + a), n0 = __ldg(A.conn + (size_t)e * NEN);
        for (int i = 0; i < 3; ++i)
        {
          const double Xr = __ldg(A.X + 3 * (size_t)node + i) - __ldg(A.X + 3 * (size_t)n0 + i);
          E[Cfg::oXm + 3 * a + i] = Xr;
          xc[i] = Xr + (__ldg(A.u + 3 * (size_t)node + i) - __ldg(A.u + 3 * (size_t)n0 + i));
        }
        if (node < A.n_row_nodes)
        {
          nd = node;
          if (WANT_K && A.K)
          {
            rb = __ldg(A.node_rowbase + node);
            rl = __ldg(A.node_rowlen + node);
            if (HAS_DBC)
              for (int i = 0; i < 3; ++i) rl |= (A.dbc[3 * node + i] ? 1 : 0) << (28 + i);
          }
        }
      }
      for (int i = 0; i < 3; ++i) E[Cfg::oX + 3 * a + i] = xc[i];
      s_rowbase[el][a] = rb;
      s_rowlen[el][a] = rl;
      s_node[el][a] = nd;
      if (a == 0) s_bad[el] = (e >= 0 && ele_bad && ele_bad[e]) ? 1 : 0;
    }
    if (SLOTS_PHASE0 && WANT_K && A.K)
    {
      for (int it = tid; it < EPB * SLW; it += Cfg::threads)
      {
        const int el = it / SLW, w = it % SLW;
        const int slot_e = ebase + el;
        if (slot_e >= A.n_ele) continue;
        const int e = A.ele_perm ? __ldg(A.ele_perm + slot_e) : slot_e;
        const unsigned char* base = A.slots + (size_t)e * NEN * NEN;
        const int shift = (int)(reinterpret_cast<uintptr_t>(base) & 3u);
        if (w == 0) s_slshift[el] = shift;
        // words of the element's byte range; the last one may reach up to 3 bytes past it, inside the same aligned word
        if (4 * w < shift + NEN * NEN) s_slw[el][w] = __ldg(reinterpret_cast<const unsigned int*>(base - shift) + w);
      }
    }
    __syncthreads();

    // ---- phase 1a: rows of the Jacobians (J = dN X^T unless cached, j = dN x^T), one (Gauss point, matrix, row) per thread; the
    // node-major table makes a warp's loads of one node contiguous ---------------------------------------------------------
    constexpr int NW = Cfg::geom ? 1 : 2;
    for (int it = tid; it < EPB * NGP * 3 * NW; it += Cfg::threads)
    {
      const int d = it % 3, which = Cfg::geom ? 1 : (it / 3) & 1, gp = (it / (3 * NW)) % NGP, el = it / (3 * NW * NGP);
      if (ebase + el >= A.n_ele) continue;
      double* E = smem_nhg + (size_t)el * Cfg::ele_doubles;
      const double* crd = E + (which ? Cfg::oX : Cfg::oXm);
      const double* dN = ele_table<CT>(A.dNtab, A.ele_perm ? __ldg(A.ele_perm + ebase + el) : ebase + el) + Cfg::oTabNodeMajor + gp * 3 + d;
      double m0 = 0.0, m1 = 0.0, m2 = 0.0;
#pragma unroll
      for (int a = 0; a < NEN; ++a)
      {
        const double dn = __ldg(dN + a * NGP * 3);
        m0 += dn * crd[3 * a];
        m1 += dn * crd[3 * a + 1];
        m2 += dn * crd[3 * a + 2];
      }
      double* R = E + Cfg::oInv + gp * 20 + (which ? 0 : 9) + 3 * d;
      R[0] = m0;
      R[1] = m1;
      R[2] = m2;
    }
    if (Cfg::geom)  // detJ from the geometry cache, on the threads the loop above leaves idle first
      for (int it = Cfg::threads - 1 - tid; it < EPB * NGP; it += Cfg::threads)
      {
        const int gp = it % NGP, el = it / NGP;
        if (ebase + el >= A.n_ele) continue;
        const int eg = A.ele_perm ? __ldg(A.ele_perm + ebase + el) : ebase + el;
        smem_nhg[(size_t)el * Cfg::ele_doubles + Cfg::oInv + gp * 20 + 18] = __ldg(A.geom + (size_t)eg * Cfg::geom_doubles + 2 * 32 * MT * MT + gp);
      }
    __syncthreads();

    // ---- phase 1b: inverses, Gauss-point scalars ---------------------------------------------------------------
    for (int it = tid; it < EPB * Cfg::GP4; it += Cfg::threads)
    {
      const int gp = it % Cfg::GP4, el = it / Cfg::GP4;
      const int e = ebase + el;
      double* R = smem_nhg + (size_t)el * Cfg::ele_doubles + Cfg::oInv + gp * 20;
      if (e >= A.n_ele || gp >= NGP)
      {
        for (int i = 0; i < 20; ++i) R[i] = 0.0;
        continue;
      }
      double M[9], Mi[9];
      for (int i = 0; i < 9; ++i) M[i] = R[i];
      const double detj = inv3(M, Mi);
      for (int i = 0; i < 9; ++i) R[i] = Mi[i];
      double detJ = R[18];
      if (!Cfg::geom)
      {
        for (int i = 0; i < 9; ++i) M[i] = R[9 + i];
        detJ = inv3(M, Mi);
        for (int i = 0; i < 9; ++i) R[9 + i] = Mi[i];
      }
      const double detF = detj / detJ;
      const double I3 = detF * detF;
      const double lambda = detJ * __ldg(A.gpw + gp) * c2;
      const double kappa = lambda * exp(-beta * log(I3));
      if (!(detJ > 0.0) || !(I3 > 0.0) || !isfinite(kappa)) s_bad[el] = 1;
      const double sk = sqrt(kappa), sl = sqrt(lambda);
      for (int i = 0; i < 9; ++i)
      {
        R[i] *= sk;
        if (!Cfg::geom) R[9 + i] *= sl;
      }
      R[18] = sk;
    }
    __syncthreads();

    // ---- phase 1c: component planes in fragment order --------------------------------------------------------
    const int a8 = lane >> 2, q = lane & 3;
#pragma unroll
    for (int blk = warp; blk < EPB * KS * MT; blk += Cfg::warps)
    {
      const int mt = blk % MT, ks = (blk / MT) % KS, el = blk / (MT * KS);
      double* E = smem_nhg + (size_t)el * Cfg::ele_doubles;
      const int gp = 4 * ks + q;
      double out[6] = {0, 0, 0, 0, 0, 0};
      if (ebase + el < A.n_ele)
      {
        // fragment-ordered table: one coalesced load per direction, zero where the tile is padding (R is zero for padded Gauss points)
        const double* dN = ele_table<CT>(A.dNtab, A.ele_perm ? __ldg(A.ele_perm + ebase + el) : ebase + el) + Cfg::oTabFragment + (ks * MT + mt) * 32 + lane;
        const double d0 = __ldg(dN), d1 = __ldg(dN + Cfg::plane), d2 = __ldg(dN + 2 * Cfg::plane);
        const double* R = E + Cfg::oInv + gp * 20;
#pragma unroll
        for (int i = 0; i < 3; ++i)
        {
          out[i] = R[3 * i] * d0 + R[3 * i + 1] * d1 + R[3 * i + 2] * d2;
          if (!Cfg::geom) out[3 + i] = R[9 + 3 * i] * d0 + R[9 + 3 * i + 1] * d1 + R[9 + 3 * i + 2] * d2;
        }
      }
      const int off = (ks * MT + mt) * 32 + lane;
#pragma unroll
      for (int c = 0; c < Cfg::nplanes; ++c) E[c * Cfg::plane + off] = out[c];
    }
    __syncthreads();

    // ---- phase 2: warp per (element, row tile) ----------------------------------------------------------------
    const double beta2 = 2.0 * beta;
    const int b0 = 2 * q;
    double* Ks = smem_nhg + (size_t)EPB * Cfg::ele_doubles + (size_t)warp * Cfg::stage;
#pragma unroll 1
    for (int item = warp; item < EPB * MT; item += Cfg::warps)
    {
      const int mt = item % MT, el = item / MT;
      if (ebase + el >= A.n_ele) continue;
      const int e = A.ele_perm ? __ldg(A.ele_perm + ebase + el) : ebase + el;
      if (s_bad[el])
      {
        if (mt == 0 && lane == 0) report_error(A.err, __ldg(A.ele_gid + e));
        continue;
      }
      const double* E = smem_nhg + (size_t)el * Cfg::ele_doubles;
      const int arow = 8 * mt + a8;  // this lane's row node
      if (!SLOTS_PHASE0 && WANT_K && A.K)
      {
        // slot rows of the 8 row nodes of this tile
        __syncwarp();
        for (int i = lane; i < 8 * NP; i += 32)
        {
          const int r = i / NP, b = i % NP;
          const int aa = 8 * mt + r;
          s_slots[warp * 8 + r][b] = (aa < NEN && b < NEN) ? __ldg(A.slots + ((size_t)e * NEN + aa) * NEN + b) : (unsigned char)0xFF;
        }
        __syncwarp();
      }
      double fpart[3] = {0.0, 0.0, 0.0};
      const double2* const gtile = Cfg::geom ? reinterpret_cast<const double2*>(A.geom + (size_t)e * Cfg::geom_doubles) + (size_t)mt * MT * 32 + lane : nullptr;
      double2 sg_next = Cfg::geom ? __ldg(gtile) : make_double2(0.0, 0.0);
#pragma unroll 1
      for (int nt = 0; nt < MT; ++nt)
      {
        const double2 sg = sg_next;
        if (Cfg::geom && nt + 1 < MT) sg_next = __ldg(gtile + 32 * (nt + 1));
        double W[3][3][2], sa[3][2];
#pragma unroll
        for (int i = 0; i < 3; ++i)
        {
          sa[i][0] = sa[i][1] = 0.0;
#pragma unroll
          for (int j = 0; j < 3; ++j) W[i][j][0] = W[i][j][1] = 0.0;
        }
#pragma unroll
        for (int ks = 0; ks < KS; ++ks)
        {
          double fa[6], fb[6];
          const int oa = (ks * MT + mt) * 32 + lane, ob = (ks * MT + nt) * 32 + lane;
#pragma unroll
          for (int c = 0; c < Cfg::nplanes; ++c)
          {
            fa[c] = E[c * Cfg::plane + oa];
            fb[c] = E[c * Cfg::plane + ob];
          }
#pragma unroll
          for (int i = 0; i < 3; ++i)
          {
            if (!Cfg::geom) dmma884(sa[i][0], sa[i][1], fa[3 + i], fb[3 + i]);
            if (WANT_K)
            {
#pragma unroll
              for (int j = 0; j < 3; ++j) dmma884(W[i][j][0], W[i][j][1], fa[i], fb[j]);
            }
          }
        }
        const double s0 = Cfg::geom ? c2 * sg.x : sa[0][0] + sa[1][0] + sa[2][0], s1 = Cfg::geom ? c2 * sg.y : sa[0][1] + sa[1][1] + sa[2][1];
        {
          const double* xk = E + Cfg::oX + 3 * (8 * nt + b0);
#pragma unroll
          for (int i = 0; i < 3; ++i) fpart[i] += s0 * xk[i] + s1 * xk[3 + i];
        }
        if (!WANT_K) continue;
        double Kb[2][9];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
          for (int j = 0; j < 3; ++j)
          {
            Kb[0][3 * i + j] = beta2 * W[i][j][0] + W[j][i][0] + (i == j ? s0 : 0.0);
            Kb[1][3 * i + j] = beta2 * W[i][j][1] + W[j][i][1] + (i == j ? s1 : 0.0);
          }
        if (A.Ke_out && arow < NEN)
        {
          double* Ke = A.Ke_out + (size_t)e * (9 * NEN * NEN);
          for (int c = 0; c < 2; ++c)
          {
            const int bcol = 8 * nt + b0 + c;
            if (bcol < NEN)
              for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) Ke[(3 * arow + i) + (size_t)(3 * NEN) * (3 * bcol + j)] = Kb[c][3 * i + j];
          }
        }
        if (A.K)
        {
          const int ln = lane < 24 ? lane : 23;
          const int b = ln / 3;
          const int bcol = 8 * nt + b;
#pragma unroll
          for (int half = 0; half < 2; ++half)
          {
            if (8 * mt + 4 * half >= NEN) break;  // warp-uniform: no row nodes in this half
            __syncwarp();
            if ((a8 >> 2) == half)
            {
#pragma unroll
              for (int i = 0; i < 3; ++i)
              {
                double2* row = reinterpret_cast<double2*>(Ks + (3 * (a8 & 3) + i) * 24 + 6 * q);
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
              const int rr = 4 * half + r, aa = 8 * mt + rr;
              const bool ok = aa < NEN && bcol < NEN && lane < 24;
              const long long rb = ok ? s_rowbase[el][aa] : -1;
              const int slot = !ok ? 0
                               : (SLOTS_PHASE0 ? (int)(reinterpret_cast<const unsigned char*>(s_slw[el]) + s_slshift[el])[aa * NEN + bcol]
                                               : (int)s_slots[warp * 8 + rr][bcol]);
              rl[r] = ok ? s_rowlen[el][aa] : 0;
              ptr[r] = rb >= 0 ? A.K + rb + 3 * slot + (ln - 3 * b) : nullptr;
#pragma unroll
              for (int i = 0; i < 3; ++i) v[3 * r + i] = Ks[(3 * r + i) * 24 + ln];
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
      // residual rows of this tile: f_a = sum_b s_ab x_b - sum_gp sqrt(kappa) ghat_a
#pragma unroll
      for (int ks = 0; ks < KS; ++ks)
      {
        const double sk = E[Cfg::oInv + (4 * ks + q) * 20 + 18];
        const int oa = (ks * MT + mt) * 32 + lane;
#pragma unroll
        for (int i = 0; i < 3; ++i) fpart[i] -= sk * E[i * Cfg::plane + oa];
      }
#pragma unroll
      for (int i = 0; i < 3; ++i)
      {
        fpart[i] += __shfl_xor_sync(0xFFFFFFFFu, fpart[i], 1);
        fpart[i] += __shfl_xor_sync(0xFFFFFFFFu, fpart[i], 2);
      }
      if (q < 3 && arow < NEN)
      {
        const double v = q == 0 ? fpart[0] : (q == 1 ? fpart[1] : fpart[2]);
        const int na = s_node[el][arow];
        if (A.fe_out) A.fe_out[(size_t)e * (3 * NEN) + 3 * arow + q] = v;
        if (A.f && na >= 0) red_add(A.f + 3 * (size_t)na + q, v);
      }
    }
  }
}  // namespace fc
