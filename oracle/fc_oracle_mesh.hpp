// ORACLE — TEST INFRASTRUCTURE ONLY. Never linked into or called by the product library.
//
// Restatement of the mesh / decomposition / DOF numbering side of the path
// (paths relative to /root/reference):
//   structured cuboid generator + box partition  src/core/io/src/4C_io_gridgenerator.cpp:41-323, 328-393
//   unique node ownership (lowest touching rank) src/core/rebalance/src/4C_rebalance_graph_based.cpp:159-214
//   element owner / ghosting rule                src/core/fem/src/discretization/4C_fem_discretization_partition.cpp:510-587
//   DOF numbering                                src/core/fem/src/dofset/4C_fem_dofset.cpp:344-352
// The parallel procedure is replayed serially for all ranks at once.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <set>
#include <unordered_map>
#include <vector>

#include "fc_oracle_fem.hpp"

namespace fc_oracle
{
  struct Mesh
  {
    int ct = HEX8;
    int nen = 8;
    std::vector<int> node_gid;  // ascending
    std::vector<double> node_X; // 3 per node, same order as node_gid
    std::vector<int> ele_gid;   // creation order (ascending for generated meshes)
    std::vector<int> ele_nodes; // nen node GIDs per element, reference node order
    std::unordered_map<int, int> node_index;  // gid -> position in node_gid
    std::unordered_map<int, int> ele_index;
    int min_node_gid = 0;
    // nurbs27: the element's knot spans [n_ele][3][6] (Knotvector::get_ele_knots) and the control point weights, one per node
    std::vector<double> nurbs_knots, nurbs_weights;

    void finalize()
    {
      node_index.clear();
      node_index.reserve(node_gid.size() * 2);
      for (size_t i = 0; i < node_gid.size(); ++i) node_index[node_gid[i]] = (int)i;
      ele_index.clear();
      ele_index.reserve(ele_gid.size() * 2);
      for (size_t i = 0; i < ele_gid.size(); ++i) ele_index[ele_gid[i]] = (int)i;
      min_node_gid = node_gid.empty() ? 0 : *std::min_element(node_gid.begin(), node_gid.end());
    }
    // dofset.cpp:344-352 with count = 0 and maxnodenumdf = 3
    int dof_gid(int node_gid_, int j) const { return (node_gid_ - min_node_gid) * 3 + j; }
  };

  // makes element e (index in the mesh's element order) the current NURBS element of this thread for its lifetime
  struct NurbsScope
  {
    double w[27];
    NurbsScope(const Mesh& m, size_t e)
    {
      if (m.ct != NURBS27) return;
      for (int a = 0; a < 27; ++a) w[a] = m.nurbs_weights[(size_t)m.node_index.at(m.ele_nodes[e * 27 + a])];
      nurbs_context().knots = &m.nurbs_knots[18 * e];
      nurbs_context().weights = w;
    }
    ~NurbsScope() { nurbs_context() = NurbsContext{}; }
  };

  struct Partition
  {
    int nranks = 1;
    std::vector<int> node_owner;                 // per mesh node index
    std::vector<int> ele_owner;                  // per mesh element index
    std::vector<std::vector<int>> row_nodes;     // per rank: owned node GIDs ascending
    std::vector<std::vector<int>> col_eles;      // per rank: element GIDs, row elements first then ghosts
    std::vector<int> n_row_eles;                 // per rank
    int subdiv[3] = {1, 1, 1};
  };

  // gridgenerator.cpp:95-118: prime factors of nproc (ascending), applied largest first to the
  // axis with the largest remaining interval/subdivision ratio (x wins ties, then y).
  inline void box_subdivisions(int nproc, const int* interval, unsigned* sub)
  {
    std::vector<int> factors;
    int np = nproc;
    for (int fac = 2; fac < np + 1;)
    {
      if (np % fac == 0)
      {
        factors.push_back(fac);
        np /= fac;
      }
      else
        fac++;
    }
    sub[0] = sub[1] = sub[2] = 1;
    const double d[3] = {(double)interval[0], (double)interval[1], (double)interval[2]};
    for (auto fac = factors.rbegin(); fac != factors.rend(); ++fac)
    {
      const double r[3] = {d[0] / sub[0], d[1] / sub[1], d[2] / sub[2]};
      if (r[0] >= r[1] && r[0] >= r[2])
        sub[0] *= *fac;
      else if (r[1] >= r[0] && r[1] >= r[2])
        sub[1] *= *fac;
      else if (r[2] >= r[0] && r[2] >= r[1])
        sub[2] *= *fac;
    }
  }

  // gridgenerator.cpp:120-137
  inline std::vector<unsigned> box_ranges(int n, unsigned sub)
  {
    std::vector<unsigned> r(sub + 1);
    for (size_t i = 0; i < sub + 1ul; ++i)
      r[i] = std::max(0, std::min(n, static_cast<int>(std::round(i * (double)n / sub))));
    return r;
  }

  // create_hex_element node ids (gridgenerator.cpp:328-393): node with parameter coords (r,s,t)
  // sits at lattice position (ex+r+1, ey+s+1, ez+t+1) on the (2nx+1)(2ny+1)(2nz+1) lattice.
  inline void hex_element_nodes(int ct, int eleid, int node_offset, const int* interval, int* nodeids)
  {
    const size_t ex = 2 * (eleid % interval[0]);
    const size_t ey = 2 * ((eleid / interval[0]) % interval[1]);
    const size_t ez = 2 * (eleid / (interval[0] * interval[1]));
    const size_t nx = 2 * interval[0] + 1;
    const size_t ny = 2 * interval[1] + 1;
    const int nen = num_nodes(ct);
    for (int a = 0; a < nen; ++a)
    {
      double xi[3];
      node_param_coords(ct, a, xi);
      const size_t dx = (size_t)((int)xi[0] + 1), dy = (size_t)((int)xi[1] + 1),
                   dz = (size_t)((int)xi[2] + 1);
      nodeids[a] = node_offset + (int)(((ez + dz) * ny + ey + dy) * nx + ex + dx);
    }
  }

  // Replays unique node ownership + element ownership/ghosting for an arbitrary initial
  // distribution `init_eles[rank]` (element GIDs owned before redistribution).
  inline void build_partition(const Mesh& mesh, const std::vector<std::vector<int>>& init_eles, Partition& part)
  {
    const int nranks = (int)init_eles.size();
    part.nranks = nranks;
    const int nn = (int)mesh.node_gid.size(), ne = (int)mesh.ele_gid.size();
    // rebalance_graph_based.cpp:176-201: rank p keeps the nodes of its elements that no lower
    // rank touches.
    part.node_owner.assign(nn, -1);
    for (int p = 0; p < nranks; ++p)
      for (int eg : init_eles[p])
      {
        const int e = mesh.ele_index.at(eg);
        for (int a = 0; a < mesh.nen; ++a)
        {
          int& o = part.node_owner[mesh.node_index.at(mesh.ele_nodes[(size_t)e * mesh.nen + a])];
          if (o < 0) o = p;  // ranks visited in ascending order -> lowest wins
        }
      }
    part.row_nodes.assign(nranks, {});
    for (int i = 0; i < nn; ++i)
      if (part.node_owner[i] >= 0) part.row_nodes[part.node_owner[i]].push_back(mesh.node_gid[i]);
    for (auto& v : part.row_nodes) std::sort(v.begin(), v.end());
    std::vector<int> nodesperproc(nranks);
    for (int p = 0; p < nranks; ++p) nodesperproc[p] = (int)part.row_nodes[p].size();
    int n_global_rownodes = 0;
    for (int p = 0; p < nranks; ++p) n_global_rownodes += nodesperproc[p];

    // partition.cpp:510-587, elements visited proc by proc in their initial order
    part.ele_owner.assign(ne, -1);
    part.col_eles.assign(nranks, {});
    std::vector<std::vector<int>> ghosts(nranks);
    std::vector<int> numperproc(nranks), nodeowner(mesh.nen);
    for (int p = 0; p < nranks; ++p)
      for (int eg : init_eles[p])
      {
        const int e = mesh.ele_index.at(eg);
        std::fill(numperproc.begin(), numperproc.end(), 0);
        for (int j = 0; j < mesh.nen; ++j)
        {
          nodeowner[j] = part.node_owner[mesh.node_index.at(mesh.ele_nodes[(size_t)e * mesh.nen + j])];
          numperproc[nodeowner[j]]++;
        }
        int owner = -1, maxnode = 0, minrownodes = n_global_rownodes;
        for (int j = 0; j < mesh.nen; ++j)
        {
          const int cp = nodeowner[j];
          const int own = numperproc[cp];
          if (own > maxnode || (own == maxnode && nodesperproc[cp] <= minrownodes))
          {
            owner = cp;
            maxnode = own;
            minrownodes = nodesperproc[cp];
          }
        }
        part.ele_owner[e] = owner;
        for (int q = 0; q < nranks; ++q)
        {
          if (numperproc[q] == 0) continue;  // owns none of the nodes -> neither owns nor ghosts
          if (q == owner)
            part.col_eles[q].push_back(eg);
          else
            ghosts[q].push_back(eg);
        }
      }
    part.n_row_eles.assign(nranks, 0);
    for (int q = 0; q < nranks; ++q)
    {
      part.n_row_eles[q] = (int)part.col_eles[q].size();
      part.col_eles[q].insert(part.col_eles[q].end(), ghosts[q].begin(), ghosts[q].end());
    }
  }

  // create_rectangular_cuboid_discretization with PARTITION structured  (gridgenerator.cpp:41-323)
  inline void create_cuboid(int ct, const int* interval, const double* lo, const double* hi,
      int nranks, int node_offset, Mesh& mesh, Partition& part)
  {
    mesh.ct = ct;
    mesh.nen = num_nodes(ct);
    const int numele = interval[0] * interval[1] * interval[2];
    unsigned sub[3];
    box_subdivisions(nranks, interval, sub);
    for (int d = 0; d < 3; ++d) part.subdiv[d] = (int)sub[d];
    const auto xr = box_ranges(interval[0], sub[0]);
    const auto yr = box_ranges(interval[1], sub[1]);
    const auto zr = box_ranges(interval[2], sub[2]);
    std::vector<std::vector<int>> init(nranks);
    for (int rank = 0; rank < nranks; ++rank)
    {
      const unsigned sec[3] = {rank % sub[0], (rank / sub[0]) % sub[1], rank / (sub[0] * sub[1])};
      for (size_t iz = zr[sec[2]]; iz < zr[sec[2] + 1]; ++iz)
        for (size_t it = yr[sec[1]]; it < yr[sec[1] + 1]; ++it)
          for (size_t ix = xr[sec[0]]; ix < xr[sec[0] + 1]; ++ix)
            init[rank].push_back((int)((iz * interval[1] + it) * interval[0] + ix));
    }
    mesh.ele_gid.resize(numele);
    mesh.ele_nodes.resize((size_t)numele * mesh.nen);
    for (int e = 0; e < numele; ++e)
    {
      mesh.ele_gid[e] = e;
      hex_element_nodes(ct, e, node_offset, interval, &mesh.ele_nodes[(size_t)e * mesh.nen]);
    }
    // nodes that exist = nodes touched by elements
    std::vector<int> gids(mesh.ele_nodes);
    std::sort(gids.begin(), gids.end());
    gids.erase(std::unique(gids.begin(), gids.end()), gids.end());
    mesh.node_gid = gids;
    mesh.node_X.resize(gids.size() * 3);
    const size_t nx = 2 * interval[0] + 1, ny = 2 * interval[1] + 1;
    for (size_t n = 0; n < gids.size(); ++n)
    {
      const int posid = gids[n] - node_offset;
      const size_t i = posid % nx, j = (posid / nx) % ny, k = posid / (nx * ny);
      // gridgenerator.cpp:268-276
      mesh.node_X[3 * n + 0] = static_cast<double>(i) / (2 * interval[0]) * (hi[0] - lo[0]) + lo[0];
      mesh.node_X[3 * n + 1] = static_cast<double>(j) / (2 * interval[1]) * (hi[1] - lo[1]) + lo[1];
      mesh.node_X[3 * n + 2] = static_cast<double>(k) / (2 * interval[2]) * (hi[2] - lo[2]) + lo[2];
    }
    mesh.finalize();
    build_partition(mesh, init, part);
  }

  // single-rank partition for an explicitly given mesh
  inline void trivial_partition(const Mesh& mesh, Partition& part)
  {
    std::vector<std::vector<int>> init(1);
    init[0] = mesh.ele_gid;
    build_partition(mesh, init, part);
  }
}  // namespace fc_oracle
