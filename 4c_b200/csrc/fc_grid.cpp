// Structured cuboid mesh + the reference's "PARTITION structured" decomposition, evaluated for ONE
// rank in closed form (pure host integer arithmetic, no device, no communication).
//
// What is reproduced (paths relative to the reference tree):
//   * box split: prime factors of nproc, largest first, each applied to the axis with the largest
//     interval/subdivision ratio; ranges round(i n / sub); rank -> box (r % sx, (r / sx) % sy, r / (sx sy))
//     src/core/io/src/4C_io_gridgenerator.cpp:80-143
//   * element id e = (iz ny + iy) nx + ix, node ids on the (2nx+1)(2ny+1)(2nz+1) lattice, node
//     coordinates i / (2 n) (hi - lo) + lo                    4C_io_gridgenerator.cpp:259-281,328-393
//   * unique node ownership: ranks publish the nodes of their boxes in rank order and every other
//     rank drops them, i.e. the LOWEST rank whose box touches a node owns it
//     src/core/rebalance/src/4C_rebalance_graph_based.cpp:159-214
//     Closed form: along each axis the lowest element touching lattice position p is
//     e(p) = p == 0 ? 0 : (p - 1) / 2, and rank numbers are monotone in every box coordinate, so
//     owner(p) = rank of box (box_x(e(px)), box_y(e(py)), box_z(e(pz))).
//   * element owner = rank owning most of its nodes; tie -> fewer row nodes; still tied -> last such
//     node owner in element node order; every other rank owning one of its nodes ghosts it
//     src/core/fem/src/discretization/4C_fem_discretization_partition.cpp:510-587
//   * DOF numbering 3 (gid - min gid) + j                       src/core/fem/src/dofset/4C_fem_dofset.cpp:344-352
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../include/fourc_b200.h"
#include "fc_fem.cuh"

struct fc_grid
{
  int ct = 0, nen = 8, nranks = 1, rank = 0, node_offset = 0;
  int n[3] = {1, 1, 1};
  double lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
  int sub[3] = {1, 1, 1};
  std::vector<int> range[3];
  int n_row_nodes = 0, n_row_eles = 0;
  std::vector<int> node_gid, node_owner, ele_gid, ele_owner, conn;
  std::vector<double> X;
  std::vector<int> send_count, recv_count, send_nodes;
};

namespace
{
  void subdivisions(int nproc, const int* interval, int* sub)
  {
    std::vector<int> factors;
    int np = nproc;
    for (int f = 2; f <= np;)
    {
      if (np % f == 0)
      {
        factors.push_back(f);
        np /= f;
      }
      else
        ++f;
    }
    sub[0] = sub[1] = sub[2] = 1;
    for (auto it = factors.rbegin(); it != factors.rend(); ++it)
    {
      const double r0 = (double)interval[0] / sub[0], r1 = (double)interval[1] / sub[1], r2 = (double)interval[2] / sub[2];
      if (r0 >= r1 && r0 >= r2)
        sub[0] *= *it;
      else if (r1 >= r0 && r1 >= r2)
        sub[1] *= *it;
      else
        sub[2] *= *it;
    }
  }

  struct Decomp
  {
    const fc_grid* g;
    int box_of(int d, int e) const
    {
      // largest b with range[b] <= e (empty boxes are skipped by taking the last such b)
      const std::vector<int>& r = g->range[d];
      int b = (int)(std::upper_bound(r.begin(), r.end(), e) - r.begin()) - 1;
      return std::min(b, g->sub[d] - 1);
    }
    int rank_of_box(int bx, int by, int bz) const { return bx + g->sub[0] * (by + g->sub[1] * bz); }
    int lowest_ele(int p) const { return p == 0 ? 0 : (p - 1) / 2; }
    int node_owner(int px, int py, int pz) const
    {
      return rank_of_box(box_of(0, lowest_ele(px)), box_of(1, lowest_ele(py)), box_of(2, lowest_ele(pz)));
    }
    // number of lattice positions along axis d owned by box b
    int owned_positions(int d, int b) const
    {
      const int lo = g->range[d][b], hi = g->range[d][b + 1];
      if (hi <= lo) return 0;
      const int per = g->ct == fc::HEX8 ? 1 : 2;
      return per * (hi - lo) + (lo == 0 ? 1 : 0);
    }
    int64_t nodes_per_proc(int r) const
    {
      const int bx = r % g->sub[0], by = (r / g->sub[0]) % g->sub[1], bz = r / (g->sub[0] * g->sub[1]);
      return (int64_t)owned_positions(0, bx) * owned_positions(1, by) * owned_positions(2, bz);
    }
  };
}  // namespace

extern "C"
{
  int fc_grid_create(fc_grid** out, int celltype, int nx, int ny, int nz, const double* lo, const double* hi,
      int nranks, int rank, int node_offset)
  {
    if (!out || nx < 1 || ny < 1 || nz < 1 || nranks < 1 || rank < 0 || rank >= nranks) return FC_ERR_INVALID;
    if (celltype != FC_HEX8 && celltype != FC_HEX27) return FC_ERR_UNSUPPORTED;  // the reference generator has no tets
    const int64_t NX = 2 * (int64_t)nx + 1, NY = 2 * (int64_t)ny + 1, NZ = 2 * (int64_t)nz + 1;
    if (node_offset + NX * NY * NZ > (int64_t)INT32_MAX / 3) return FC_ERR_UNSUPPORTED;  // int32 DOF GIDs
    fc_grid* g = new fc_grid;
    g->ct = celltype;
    g->nen = fc::cell_nen(celltype);
    g->nranks = nranks;
    g->rank = rank;
    g->node_offset = node_offset;
    g->n[0] = nx;
    g->n[1] = ny;
    g->n[2] = nz;
    for (int d = 0; d < 3; ++d)
    {
      g->lo[d] = lo ? lo[d] : 0.0;
      g->hi[d] = hi ? hi[d] : 1.0;
    }
    subdivisions(nranks, g->n, g->sub);
    for (int d = 0; d < 3; ++d)
    {
      g->range[d].resize(g->sub[d] + 1);
      for (int i = 0; i <= g->sub[d]; ++i)
        g->range[d][i] = std::max(0, std::min(g->n[d], (int)std::round(i * (double)g->n[d] / g->sub[d])));
    }
    Decomp D{g};
    const int nen = g->nen, step = celltype == FC_HEX8 ? 2 : 1;
    const int bq[3] = {rank % g->sub[0], (rank / g->sub[0]) % g->sub[1], rank / (g->sub[0] * g->sub[1])};
    int elo[3], ehi[3];  // candidate element range: own box grown by one layer
    for (int d = 0; d < 3; ++d)
    {
      elo[d] = std::max(0, g->range[d][bq[d]] - 1);
      ehi[d] = std::min(g->n[d], g->range[d][bq[d] + 1] + 1);
    }
    std::vector<int64_t> npp(nranks);
    int64_t nglobal = 0;
    for (int r = 0; r < nranks; ++r)
    {
      npp[r] = D.nodes_per_proc(r);
      nglobal += npp[r];
    }
    // local node offsets of the element nodes on the lattice
    std::vector<std::array<int, 3>> noff(nen);
    for (int a = 0; a < nen; ++a)
      for (int d = 0; d < 3; ++d) noff[a][d] = fc::hex_node_pos(a, d) + 1;

    struct Rec
    {
      int p_init, iz, iy, ix, owner;
    };
    std::vector<Rec> own, ghost;
    std::vector<int> nodeowner(nen), numper(nranks, 0);
    const bool box_empty = g->range[0][bq[0]] >= g->range[0][bq[0] + 1] || g->range[1][bq[1]] >= g->range[1][bq[1] + 1] ||
                           g->range[2][bq[2]] >= g->range[2][bq[2] + 1];
    if (!box_empty)
      for (int iz = elo[2]; iz < ehi[2]; ++iz)
        for (int iy = elo[1]; iy < ehi[1]; ++iy)
          for (int ix = elo[0]; ix < ehi[0]; ++ix)
          {
            bool mine = false;
            for (int a = 0; a < nen; ++a)
            {
              nodeowner[a] = D.node_owner(2 * ix + noff[a][0], 2 * iy + noff[a][1], 2 * iz + noff[a][2]);
              mine |= nodeowner[a] == rank;
            }
            if (!mine) continue;
            for (int a = 0; a < nen; ++a) numper[nodeowner[a]] = 0;
            for (int a = 0; a < nen; ++a) numper[nodeowner[a]]++;
            int owner = -1, maxnode = 0;
            int64_t minrow = nglobal;
            for (int a = 0; a < nen; ++a)
            {
              const int cp = nodeowner[a], cnt = numper[cp];
              if (cnt > maxnode || (cnt == maxnode && npp[cp] <= minrow))
              {
                owner = cp;
                maxnode = cnt;
                minrow = npp[cp];
              }
            }
            const int p_init = D.rank_of_box(D.box_of(0, ix), D.box_of(1, iy), D.box_of(2, iz));
            (owner == rank ? own : ghost).push_back(Rec{p_init, iz, iy, ix, owner});
          }
    auto by_visit = [](const Rec& a, const Rec& b)
    {
      if (a.p_init != b.p_init) return a.p_init < b.p_init;
      if (a.iz != b.iz) return a.iz < b.iz;
      if (a.iy != b.iy) return a.iy < b.iy;
      return a.ix < b.ix;
    };
    std::sort(own.begin(), own.end(), by_visit);
    std::sort(ghost.begin(), ghost.end(), by_visit);
    g->n_row_eles = (int)own.size();
    std::vector<Rec> eles(own);
    eles.insert(eles.end(), ghost.begin(), ghost.end());

    // local lattice window for node LIDs
    int plo[3], pn[3];
    for (int d = 0; d < 3; ++d)
    {
      plo[d] = 2 * elo[d];
      pn[d] = (2 * ehi[d] - plo[d]) / step + 1;
    }
    std::vector<int> lid((size_t)pn[0] * pn[1] * pn[2], -1);
    auto widx = [&](int px, int py, int pz) { return ((size_t)((pz - plo[2]) / step) * pn[1] + (py - plo[1]) / step) * pn[0] + (px - plo[0]) / step; };
    auto gid_of = [&](int px, int py, int pz) { return node_offset + (int)(((int64_t)pz * NY + py) * NX + px); };
    // row nodes, ascending GID
    auto owned_lo = [&](int d) { const int l = g->range[d][bq[d]]; return l == 0 ? 0 : (celltype == FC_HEX8 ? 2 * l + 2 : 2 * l + 1); };
    auto owned_hi = [&](int d) { return 2 * g->range[d][bq[d] + 1]; };
    if (!box_empty)
      for (int pz = owned_lo(2); pz <= owned_hi(2); pz += step)
        for (int py = owned_lo(1); py <= owned_hi(1); py += step)
          for (int px = owned_lo(0); px <= owned_hi(0); px += step)
          {
            lid[widx(px, py, pz)] = (int)g->node_gid.size();
            g->node_gid.push_back(gid_of(px, py, pz));
            g->node_owner.push_back(rank);
          }
    g->n_row_nodes = (int)g->node_gid.size();
    // ghost nodes: nodes of column elements owned by others, sorted by (owner, gid)
    struct GN
    {
      int owner, gid, px, py, pz;
    };
    std::vector<GN> gn;
    for (const Rec& r : eles)
      for (int a = 0; a < nen; ++a)
      {
        const int px = 2 * r.ix + noff[a][0], py = 2 * r.iy + noff[a][1], pz = 2 * r.iz + noff[a][2];
        int& l = lid[widx(px, py, pz)];
        if (l == -1)
        {
          l = -2;  // seen, LID assigned after sorting
          gn.push_back(GN{D.node_owner(px, py, pz), gid_of(px, py, pz), px, py, pz});
        }
      }
    std::sort(gn.begin(), gn.end(), [](const GN& a, const GN& b) { return a.owner != b.owner ? a.owner < b.owner : a.gid < b.gid; });
    g->recv_count.assign(nranks, 0);
    for (const GN& n : gn)
    {
      lid[widx(n.px, n.py, n.pz)] = (int)g->node_gid.size();
      g->node_gid.push_back(n.gid);
      g->node_owner.push_back(n.owner);
      g->recv_count[n.owner]++;
    }
    // coordinates (gridgenerator.cpp:268-276)
    g->X.resize(3 * g->node_gid.size());
    for (size_t i = 0; i < g->node_gid.size(); ++i)
    {
      const int64_t pos = g->node_gid[i] - node_offset;
      const int64_t p[3] = {pos % NX, (pos / NX) % NY, pos / (NX * NY)};
      for (int d = 0; d < 3; ++d) g->X[3 * i + d] = static_cast<double>(p[d]) / (2 * g->n[d]) * (g->hi[d] - g->lo[d]) + g->lo[d];
    }
    // elements + halo send plan
    g->ele_gid.resize(eles.size());
    g->ele_owner.resize(eles.size());
    g->conn.resize(eles.size() * nen);
    std::vector<std::pair<int, int>> sends;  // (peer, row node lid)
    for (size_t e = 0; e < eles.size(); ++e)
    {
      const Rec& r = eles[e];
      g->ele_gid[e] = (r.iz * g->n[1] + r.iy) * g->n[0] + r.ix;
      g->ele_owner[e] = r.owner;
      bool mixed = false;
      for (int a = 0; a < nen; ++a)
      {
        const int l = lid[widx(2 * r.ix + noff[a][0], 2 * r.iy + noff[a][1], 2 * r.iz + noff[a][2])];
        g->conn[e * nen + a] = l;
        mixed |= g->node_owner[l] != rank;
      }
      if (mixed)
        for (int a = 0; a < nen; ++a)
        {
          const int la = g->conn[e * nen + a];
          if (la >= g->n_row_nodes) continue;
          for (int b = 0; b < nen; ++b)
          {
            const int p = g->node_owner[g->conn[e * nen + b]];
            if (p != rank) sends.push_back({p, la});
          }
        }
    }
    std::sort(sends.begin(), sends.end());
    sends.erase(std::unique(sends.begin(), sends.end()), sends.end());
    g->send_count.assign(nranks, 0);
    for (auto& s : sends)
    {
      g->send_count[s.first]++;
      g->send_nodes.push_back(s.second);  // row LIDs ascend with GID
    }
    *out = g;
    return FC_OK;
  }

  void fc_grid_destroy(fc_grid* g) { delete g; }

  int fc_grid_sizes(const fc_grid* g, int64_t* s)
  {
    if (!g || !s) return FC_ERR_INVALID;
    s[0] = g->n_row_nodes;
    s[1] = (int64_t)g->node_gid.size();
    s[2] = g->n_row_eles;
    s[3] = (int64_t)g->ele_gid.size();
    s[4] = g->node_offset;
    s[5] = g->nen;
    return FC_OK;
  }

  int fc_grid_subdivisions(const fc_grid* g, int* sub3)
  {
    if (!g || !sub3) return FC_ERR_INVALID;
    for (int d = 0; d < 3; ++d) sub3[d] = g->sub[d];
    return FC_OK;
  }

  int fc_grid_get_nodes(const fc_grid* g, int* node_gid, int* node_owner, double* X)
  {
    if (!g) return FC_ERR_INVALID;
    if (node_gid) std::memcpy(node_gid, g->node_gid.data(), g->node_gid.size() * sizeof(int));
    if (node_owner) std::memcpy(node_owner, g->node_owner.data(), g->node_owner.size() * sizeof(int));
    if (X) std::memcpy(X, g->X.data(), g->X.size() * sizeof(double));
    return FC_OK;
  }

  int fc_grid_get_elements(const fc_grid* g, int* ele_gid, int* ele_owner, int* conn)
  {
    if (!g) return FC_ERR_INVALID;
    if (ele_gid) std::memcpy(ele_gid, g->ele_gid.data(), g->ele_gid.size() * sizeof(int));
    if (ele_owner) std::memcpy(ele_owner, g->ele_owner.data(), g->ele_owner.size() * sizeof(int));
    if (conn) std::memcpy(conn, g->conn.data(), g->conn.size() * sizeof(int));
    return FC_OK;
  }

  int fc_grid_halo_counts(const fc_grid* g, int* send_node_count, int* recv_node_count)
  {
    if (!g) return FC_ERR_INVALID;
    if (send_node_count) std::memcpy(send_node_count, g->send_count.data(), g->nranks * sizeof(int));
    if (recv_node_count) std::memcpy(recv_node_count, g->recv_count.data(), g->nranks * sizeof(int));
    return FC_OK;
  }

  int fc_grid_halo_send_nodes(const fc_grid* g, int* send_node_lids)
  {
    if (!g || !send_node_lids) return FC_ERR_INVALID;
    std::memcpy(send_node_lids, g->send_nodes.data(), g->send_nodes.size() * sizeof(int));
    return FC_OK;
  }
}
