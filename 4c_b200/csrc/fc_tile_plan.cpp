// Host construction of the hex8 tile plan (fc_tile_plan.hpp).
#include "fc_tile_plan.hpp"

#include <algorithm>
#include <cmath>
#include <thread>

namespace
{
  using fc_host::TilePlanHost;
  constexpr int TE = TilePlanHost::tile_eles;

  // elements sorted by (bin of 4 x 2 x 2 lattice cells, position inside the bin, element index); the lattice is the
  // bounding box divided into cells of the mean element extent per axis, bins ascend z-major like the node numbering of a
  // structured cube, so consecutive tiles sweep the CSR rows front to back
  void order_elements(int n_col_nodes, const double* X, int ne, const int* conn, std::vector<std::pair<unsigned long long, int>>& keyed)
  {
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    for (int n = 0; n < n_col_nodes; ++n)
      for (int d = 0; d < 3; ++d)
      {
        lo[d] = std::min(lo[d], X[3 * (size_t)n + d]);
        hi[d] = std::max(hi[d], X[3 * (size_t)n + d]);
      }
    // lattice cell = mean element extent per axis
    double hsum[3] = {0, 0, 0};
    for (int e = 0; e < ne; ++e)
      for (int d = 0; d < 3; ++d)
      {
        double mn = 1e300, mx = -1e300;
        for (int a = 0; a < 8; ++a)
        {
          const double x = X[3 * (size_t)conn[(size_t)e * 8 + a] + d];
          mn = std::min(mn, x);
          mx = std::max(mx, x);
        }
        hsum[d] += mx - mn;
      }
    long long nd[3];
    double hd[3];
    for (int d = 0; d < 3; ++d)
    {
      const double ext = hi[d] - lo[d], h = hsum[d] / std::max(1, ne);
      nd[d] = (ext > 0.0 && h > 0.0) ? std::max(1LL, std::llround(ext / h)) : 1;
      nd[d] = std::min(nd[d], 1LL << 20);
      hd[d] = ext > 0.0 ? ext / (double)nd[d] : 1.0;
    }
    const unsigned long long NBx = (unsigned long long)(nd[0] + 3) / 4, NBy = (unsigned long long)(nd[1] + 1) / 2;
    keyed.resize((size_t)ne);
    for (int e = 0; e < ne; ++e)
    {
      long long q[3];
      for (int d = 0; d < 3; ++d)
      {
        double c = 0.0;
        for (int a = 0; a < 8; ++a) c += X[3 * (size_t)conn[(size_t)e * 8 + a] + d];
        q[d] = (long long)std::floor((0.125 * c - lo[d]) / hd[d]);
        q[d] = std::max(0LL, std::min(nd[d] - 1, q[d]));
      }
      const unsigned long long bin = ((unsigned long long)(q[2] >> 1) * NBy + (unsigned long long)(q[1] >> 1)) * NBx + (unsigned long long)(q[0] >> 2);
      const unsigned long long sub = (unsigned long long)(((q[2] & 1) * 2 + (q[1] & 1)) * 4 + (q[0] & 3));
      keyed[e] = {bin << 4 | sub, e};
    }
    std::sort(keyed.begin(), keyed.end());
  }

  struct TileWork
  {
    int rows[8 * TE];             // owned nodes of the tile, ascending
    int ncol[8 * TE];             // column nodes of each row
    int cols[8 * TE][8 * TE];     // ascending
  };

  // blocks of one tile; returns the number of blocks or -1 (duplicate node in an element / pair missing from the graph)
  int process_tile(const int* eles, const int* conn, int nrow, const int64_t* nodeptr, const int* nodeadj, TileWork& W,
      unsigned* blk_out /*[16][32]*/, std::vector<unsigned long long>& rec, std::vector<int>& rownode)
  {
    int nr = 0;
    for (int s = 0; s < TE; ++s)
    {
      if (eles[s] < 0) continue;
      const int* en = conn + (size_t)eles[s] * 8;
      for (int a = 0; a < 8; ++a)
      {
        for (int b = a + 1; b < 8; ++b)
          if (en[a] == en[b]) return -1;
        if (en[a] < nrow) W.rows[nr++] = en[a];
      }
    }
    std::sort(W.rows, W.rows + nr);
    nr = (int)(std::unique(W.rows, W.rows + nr) - W.rows);
    for (int r = 0; r < nr; ++r) W.ncol[r] = 0;
    for (int s = 0; s < TE; ++s)
    {
      if (eles[s] < 0) continue;
      const int* en = conn + (size_t)eles[s] * 8;
      for (int a = 0; a < 8; ++a)
      {
        if (en[a] >= nrow) continue;
        const int r = (int)(std::lower_bound(W.rows, W.rows + nr, en[a]) - W.rows);
        for (int b = 0; b < 8; ++b) W.cols[r][W.ncol[r]++] = en[b];
      }
    }
    int start[8 * TE + 1];
    start[0] = 0;
    for (int r = 0; r < nr; ++r)
    {
      std::sort(W.cols[r], W.cols[r] + W.ncol[r]);
      W.ncol[r] = (int)(std::unique(W.cols[r], W.cols[r] + W.ncol[r]) - W.cols[r]);
      start[r + 1] = start[r] + W.ncol[r];
      const int a = W.rows[r];
      const int* adj = nodeadj + nodeptr[a];
      const int L = (int)(nodeptr[a + 1] - nodeptr[a]);
      for (int c = 0; c < W.ncol[r]; ++c)
      {
        const int* p = std::lower_bound(adj, adj + L, W.cols[r][c]);
        if (p == adj + L || *p != W.cols[r][c]) return -1;
        rec.push_back((unsigned long long)(9 * nodeptr[a] + 3 * (int64_t)(p - adj)) | (unsigned long long)(3 * L) << 44);
        rownode.push_back(a);
      }
    }
    for (int s = 0; s < TE; ++s)
      for (int a = 0; a < 8; ++a)
        for (int q = 0; q < 4; ++q)
        {
          unsigned v = 0xFFFFFFFFu;
          if (eles[s] >= 0 && conn[(size_t)eles[s] * 8 + a] < nrow)
          {
            const int* en = conn + (size_t)eles[s] * 8;
            const int r = (int)(std::lower_bound(W.rows, W.rows + nr, en[a]) - W.rows);
            const unsigned i0 = (unsigned)(start[r] + (std::lower_bound(W.cols[r], W.cols[r] + W.ncol[r], en[2 * q]) - W.cols[r]));
            const unsigned i1 = (unsigned)(start[r] + (std::lower_bound(W.cols[r], W.cols[r] + W.ncol[r], en[2 * q + 1]) - W.cols[r]));
            v = i0 | i1 << 16;
          }
          blk_out[s * 32 + a * 4 + q] = v;
        }
    return start[nr];
  }
}  // namespace

namespace fc_host
{
  bool l2_sweep_order(int n_col_nodes, const double* X, int n_ele, int nen, const int* conn, long long layer_elems, std::vector<int>& perm)
  {
    perm.clear();
    if (n_ele <= 0 || layer_elems <= 0 || n_ele <= 2 * layer_elems) return false;
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300}, hsum[3] = {0, 0, 0};
    for (int n = 0; n < n_col_nodes; ++n)
      for (int d = 0; d < 3; ++d)
      {
        lo[d] = std::min(lo[d], X[3 * (size_t)n + d]);
        hi[d] = std::max(hi[d], X[3 * (size_t)n + d]);
      }
    for (int e = 0; e < n_ele; ++e)
      for (int d = 0; d < 3; ++d)
      {
        double mn = 1e300, mx = -1e300;
        for (int a = 0; a < nen; ++a)
        {
          const double x = X[3 * (size_t)conn[(size_t)e * nen + a] + d];
          mn = std::min(mn, x);
          mx = std::max(mx, x);
        }
        hsum[d] += mx - mn;
      }
    long long nd[3];
    double hd[3];
    for (int d = 0; d < 3; ++d)
    {
      const double ext = hi[d] - lo[d], h = hsum[d] / n_ele;
      nd[d] = (ext > 0.0 && h > 0.0) ? std::max(1LL, std::llround(ext / h)) : 1;
      nd[d] = std::min(nd[d], 1LL << 20);
      hd[d] = ext > 0.0 ? ext / (double)nd[d] : 1.0;
    }
    // elements per lattice cell (6 for a Kuhn-split tet mesh)
    const long long per_cell = std::max(1LL, (long long)std::llround((double)n_ele / ((double)nd[0] * (double)nd[1] * (double)nd[2])));
    if (nd[0] * nd[1] * per_cell <= layer_elems) return false;  // a z layer fits: keep the caller's order
    const long long W = std::max(4LL, layer_elems / (nd[0] * per_cell));
    if (W >= nd[1]) return false;
    std::vector<std::pair<unsigned long long, int>> keyed((size_t)n_ele);
    for (int e = 0; e < n_ele; ++e)
    {
      long long q[3];
      for (int d = 0; d < 3; ++d)
      {
        double c = 0.0;
        for (int a = 0; a < nen; ++a) c += X[3 * (size_t)conn[(size_t)e * nen + a] + d];
        q[d] = (long long)std::floor((c / nen - lo[d]) / hd[d]);
        q[d] = std::max(0LL, std::min(nd[d] - 1, q[d]));
      }
      const unsigned long long key = (((unsigned long long)(q[1] / W) * (unsigned long long)nd[2] + (unsigned long long)q[2]) * (unsigned long long)nd[1] +
                                         (unsigned long long)q[1]) * (unsigned long long)nd[0] + (unsigned long long)q[0];
      keyed[e] = {key, e};
    }
    std::sort(keyed.begin(), keyed.end());
    perm.resize((size_t)n_ele);
    bool identity = true;
    for (int k = 0; k < n_ele; ++k)
    {
      perm[k] = keyed[k].second;
      identity = identity && perm[k] == k;
    }
    if (identity) perm.clear();
    return !identity;
  }

  bool build_tile_plan_host(int n_col_nodes, const double* X, int n_ele, const int* conn, int n_row_nodes,
      const int64_t* nodeptr, const int* nodeadj, TilePlanHost& plan)
  {
    plan = TilePlanHost();
    if (n_ele <= 0) return false;
    std::vector<std::pair<unsigned long long, int>> keyed;
    order_elements(n_col_nodes, X, n_ele, conn, keyed);
    // a tile closes when it is full, or when the bin changes and it is at least half full
    plan.tile_ele.reserve((size_t)n_ele + (size_t)n_ele / 4 + TE);
    {
      int cnt = 0;
      unsigned long long bin = ~0ull;
      auto close = [&]() {
        for (; cnt % TE != 0; ++cnt) plan.tile_ele.push_back(-1);
        cnt = 0;
      };
      for (int k = 0; k < n_ele; ++k)
      {
        const unsigned long long b = keyed[k].first >> 4;
        if (cnt == TE || (cnt >= TE / 2 && b != bin)) close();
        bin = b;
        plan.tile_ele.push_back(keyed[k].second);
        ++cnt;
      }
      close();
    }
    std::vector<std::pair<unsigned long long, int>>().swap(keyed);
    const int nt = (int)(plan.tile_ele.size() / TE);
    plan.n_tiles = nt;
    plan.tile_blk.assign((size_t)nt * TE * 32, 0xFFFFFFFFu);
    plan.tile_ptr.assign((size_t)nt + 1, 0);
    unsigned nthr = std::max(1u, std::min(std::thread::hardware_concurrency(), 32u));
    if (nt < 256) nthr = 1;
    std::vector<std::vector<unsigned long long>> part_rec(nthr);
    std::vector<std::vector<int>> part_row(nthr);
    std::vector<int> failed(nthr, 0);
    auto work = [&](unsigned t) {
      const int lo = (int)((int64_t)nt * t / nthr), hi = (int)((int64_t)nt * (t + 1) / nthr);
      std::vector<TileWork> W(1);
      part_rec[t].reserve((size_t)(hi - lo) * 640);
      part_row[t].reserve((size_t)(hi - lo) * 640);
      for (int k = lo; k < hi; ++k)
      {
        const int nb = process_tile(plan.tile_ele.data() + (size_t)k * TE, conn, n_row_nodes, nodeptr, nodeadj, W[0],
            plan.tile_blk.data() + (size_t)k * TE * 32, part_rec[t], part_row[t]);
        if (nb < 0)
        {
          failed[t] = 1;
          return;
        }
        plan.tile_ptr[k + 1] = nb;
      }
    };
    if (nthr == 1)
      work(0);
    else
    {
      std::vector<std::thread> th;
      for (unsigned t = 0; t < nthr; ++t) th.emplace_back(work, t);
      for (auto& t : th) t.join();
    }
    for (unsigned t = 0; t < nthr; ++t)
      if (failed[t])
      {
        plan = TilePlanHost();
        return false;
      }
    for (int k = 0; k < nt; ++k)
    {
      plan.max_blocks = std::max(plan.max_blocks, (int)plan.tile_ptr[k + 1]);
      plan.tile_ptr[k + 1] += plan.tile_ptr[k];
    }
    plan.rec.reserve((size_t)plan.tile_ptr[nt]);
    plan.rownode.reserve((size_t)plan.tile_ptr[nt]);
    for (unsigned t = 0; t < nthr; ++t)
    {
      plan.rec.insert(plan.rec.end(), part_rec[t].begin(), part_rec[t].end());
      plan.rownode.insert(plan.rownode.end(), part_row[t].begin(), part_row[t].end());
      std::vector<unsigned long long>().swap(part_rec[t]);
      std::vector<int>().swap(part_row[t]);
    }
    return true;
  }
}  // namespace fc_host

// ---- C ABI (include/fourc_b200.h): the plan as plain arrays, for inspection and the CPU test suite -------------
struct fc_tile_plan
{
  fc_host::TilePlanHost p;
};

extern "C"
{
  int fc_tile_plan_create(int n_col_nodes, const double* X, int n_ele, const int* conn, int n_row_nodes, const int64_t* nodeptr,
      const int* nodeadj, fc_tile_plan** out)
  {
    if (!out || !X || !conn || !nodeptr || !nodeadj || n_col_nodes <= 0 || n_ele < 0 || n_row_nodes < 0) return -1;
    *out = nullptr;
    fc_tile_plan* h = new fc_tile_plan();
    if (!fc_host::build_tile_plan_host(n_col_nodes, X, n_ele, conn, n_row_nodes, nodeptr, nodeadj, h->p))
    {
      delete h;
      return -5;  // FC_ERR_GRAPH: no plan for this mesh (the library then scatters element by element)
    }
    *out = h;
    return 0;
  }

  int fc_tile_plan_sizes(const fc_tile_plan* h, int* n_tiles, int64_t* n_blocks, int* max_blocks)
  {
    if (!h) return -1;
    if (n_tiles) *n_tiles = h->p.n_tiles;
    if (n_blocks) *n_blocks = (int64_t)h->p.rec.size();
    if (max_blocks) *max_blocks = h->p.max_blocks;
    return 0;
  }

  int fc_tile_plan_get(const fc_tile_plan* h, int* tile_ele, uint32_t* tile_blk, int64_t* tile_ptr, uint64_t* rec, int* rownode)
  {
    if (!h) return -1;
    const fc_host::TilePlanHost& p = h->p;
    if (tile_ele) std::copy(p.tile_ele.begin(), p.tile_ele.end(), tile_ele);
    if (tile_blk) std::copy(p.tile_blk.begin(), p.tile_blk.end(), tile_blk);
    if (tile_ptr) std::copy(p.tile_ptr.begin(), p.tile_ptr.end(), tile_ptr);
    if (rec) std::copy(p.rec.begin(), p.rec.end(), rec);
    if (rownode) std::copy(p.rownode.begin(), p.rownode.end(), rownode);
    return 0;
  }

  void fc_tile_plan_destroy(fc_tile_plan* h) { delete h; }

  int fc_l2_sweep_order(int n_col_nodes, const double* X, int n_ele, int nen, const int* conn, int64_t layer_elems, int* perm)
  {
    if (!X || !conn || !perm || n_ele < 0) return -1;
    std::vector<int> p;
    const bool changed = fc_host::l2_sweep_order(n_col_nodes, X, n_ele, nen, conn, layer_elems, p);
    for (int k = 0; k < n_ele; ++k) perm[k] = changed ? p[k] : k;
    return changed ? 1 : 0;
  }
}
