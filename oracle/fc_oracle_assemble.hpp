// ORACLE — TEST INFRASTRUCTURE ONLY. Never linked into or called by the product library.
//
// Restatement of the per-rank evaluate + assemble loop (paths relative to /root/reference):
//   Discretization::set_state (row -> column import)   src/core/fem/src/discretization/4C_fem_discretization.cpp:503-558
//   Discretization::evaluate element loop              src/core/fem/src/discretization/4C_fem_discretization_evaluate.cpp:65-103
//   Element::location_vector                           src/core/fem/src/general/element/4C_fem_general_element.cpp:474-541
//   SparseMatrix::assemble (local-index fast branch)   src/core/linalg/src/sparse/4C_linalg_sparsematrix.cpp:444-576
//   LinAlg::assemble(Vector)                           src/core/linalg/src/sparse/4C_linalg_utils_sparse_algebra_assemble.cpp:72-92
//   sparsity pattern = full element connectivity "zeros included"  4C_linalg_sparsematrix.cpp:578-611
//
// Third-party ordering (Epetra @ Trilinos 06db4c850654feacabdaed61ee8308219266b6a5, not under
// /root/reference): after FillComplete rows are sorted by LOCAL column id; the local column map
// lists owned GIDs first (row-map order) and then remote GIDs grouped by owner rank.  This
// ordering is not pinned by any reference test ("CSR-position parity unpinned"); comparisons use
// global (row GID, col GID) keys.
#pragma once
#include <thread>

#include "fc_oracle_eas.hpp"
#include "fc_oracle_shell.hpp"
#include "fc_oracle_element.hpp"
#include "fc_oracle_mesh.hpp"

namespace fc_oracle
{
  struct RankSystem
  {
    int rank = 0;
    std::vector<int> col_node_gid;  // owned nodes (ascending) then ghosts sorted by (owner,gid)
    std::unordered_map<int, int> col_node_lid;
    int n_row_nodes = 0;
    std::vector<int> row_dof_gid;  // 3 per owned node
    std::vector<int> col_dof_gid;  // 3 per column node
    std::unordered_map<int, int> row_dof_lid, col_dof_lid;
    std::vector<int64_t> rowptr;
    std::vector<int> colind;  // local column ids, ascending per row
    std::vector<char> dbc_row;  // per local row: skip in assemble (sparsematrix.cpp:480)
  };

  inline void build_rank_system(const Mesh& mesh, const Partition& part, int rank, RankSystem& rs)
  {
    rs.rank = rank;
    const auto& rown = part.row_nodes[rank];
    rs.n_row_nodes = (int)rown.size();
    // column nodes = nodes of all column elements
    std::vector<std::pair<int, int>> ghosts;  // (owner, gid)
    {
      std::set<int> seen(rown.begin(), rown.end());
      for (int eg : part.col_eles[rank])
      {
        const int e = mesh.ele_index.at(eg);
        for (int a = 0; a < mesh.nen; ++a)
        {
          const int g = mesh.ele_nodes[(size_t)e * mesh.nen + a];
          if (seen.insert(g).second) ghosts.push_back({part.node_owner[mesh.node_index.at(g)], g});
        }
      }
    }
    std::sort(ghosts.begin(), ghosts.end());
    rs.col_node_gid = rown;
    for (auto& g : ghosts) rs.col_node_gid.push_back(g.second);
    rs.col_node_lid.clear();
    for (size_t i = 0; i < rs.col_node_gid.size(); ++i) rs.col_node_lid[rs.col_node_gid[i]] = (int)i;
    rs.row_dof_gid.clear();
    rs.col_dof_gid.clear();
    for (size_t i = 0; i < rs.col_node_gid.size(); ++i)
      for (int j = 0; j < 3; ++j)
      {
        const int d = mesh.dof_gid(rs.col_node_gid[i], j);
        rs.col_dof_gid.push_back(d);
        if ((int)i < rs.n_row_nodes) rs.row_dof_gid.push_back(d);
      }
    rs.row_dof_lid.clear();
    rs.col_dof_lid.clear();
    for (size_t i = 0; i < rs.row_dof_gid.size(); ++i) rs.row_dof_lid[rs.row_dof_gid[i]] = (int)i;
    for (size_t i = 0; i < rs.col_dof_gid.size(); ++i) rs.col_dof_lid[rs.col_dof_gid[i]] = (int)i;

    // node adjacency of owned nodes through column elements
    std::vector<std::vector<int>> adj(rs.n_row_nodes);
    for (int eg : part.col_eles[rank])
    {
      const int e = mesh.ele_index.at(eg);
      const int* en = &mesh.ele_nodes[(size_t)e * mesh.nen];
      for (int a = 0; a < mesh.nen; ++a)
      {
        const int la = rs.col_node_lid.at(en[a]);
        if (la >= rs.n_row_nodes) continue;
        for (int b = 0; b < mesh.nen; ++b) adj[la].push_back(rs.col_node_lid.at(en[b]));
      }
    }
    rs.rowptr.assign(rs.row_dof_gid.size() + 1, 0);
    rs.colind.clear();
    for (int n = 0; n < rs.n_row_nodes; ++n)
    {
      auto& v = adj[n];
      std::sort(v.begin(), v.end());
      v.erase(std::unique(v.begin(), v.end()), v.end());
      for (int d = 0; d < 3; ++d)
      {
        for (int m : v)
          for (int j = 0; j < 3; ++j) rs.colind.push_back(3 * m + j);
        rs.rowptr[3 * n + d + 1] = (int64_t)rs.colind.size();
      }
    }
    rs.dbc_row.assign(rs.row_dof_gid.size(), 0);
  }

  // set_state: import row-layout vectors of all ranks into this rank's column layout
  inline void import_column_state(const Mesh& mesh, const Partition& part,
      const std::vector<RankSystem>& all, int rank, const std::vector<std::vector<double>>& u_row,
      std::vector<double>& u_col)
  {
    const RankSystem& rs = all[rank];
    u_col.resize(rs.col_dof_gid.size());
    for (size_t i = 0; i < rs.col_node_gid.size(); ++i)
    {
      const int owner = part.node_owner[mesh.node_index.at(rs.col_node_gid[i])];
      const RankSystem& os = all[owner];
      for (int j = 0; j < 3; ++j)
        u_col[3 * i + j] = u_row[owner][os.row_dof_lid.at(rs.col_dof_gid[3 * i + j])];
    }
  }

  // One rank: loop over column elements, evaluate, assemble owned rows.
  // action: 0 = nlnstiff (f+K), 1 = internalforce (f only).  K_values/f_row are ADDED to.
  // returns number of elements with an error; first_bad = GID of the first one.
  inline int evaluate_rank(const Mesh& mesh, const Partition& part, const RankSystem& rs, int form,
      const MaterialParams& mp, int action, const double* u_col, double* f_row, double* K_values,
      int* first_bad, const double* ele_fiber = nullptr /* [n_ele][3] FIBER1 in mesh element order, or [n_ele][ngp][3] */,
      int fiber_gp_stride = 0 /* 3: ele_fiber holds one fibre per Gauss point */,
      const double* mulf_hist = nullptr /* FORM_MULF: [n_ele][ngp][18] in mesh element order */,
      double* eas_hist = nullptr /* FORM_EAS_*: [n_ele][eas_history_doubles] in mesh element order; invKaa, Kda, s are written */)
  {
    const int nen = mesh.nen, nd = 3 * nen;
    const int neas = eas_modes(form);
    std::vector<int> lm(nd), lmowner(nd), localcol(nd);
    std::vector<double> X(nd), u(nd), fe(nd), Ke((size_t)nd * nd);
    int nbad = 0;
    for (int eg : part.col_eles[rs.rank])
    {
      const int e = mesh.ele_index.at(eg);
      const int* en = &mesh.ele_nodes[(size_t)e * nen];
      const NurbsScope nurbs(mesh, (size_t)e);
      // location vector: node-major x,y,z; owner of a dof = owner of its node (element.cpp:474-541)
      for (int a = 0; a < nen; ++a)
      {
        const int ni = mesh.node_index.at(en[a]);
        for (int j = 0; j < 3; ++j)
        {
          lm[3 * a + j] = mesh.dof_gid(en[a], j);
          lmowner[3 * a + j] = part.node_owner[ni];
          X[3 * a + j] = mesh.node_X[3 * (size_t)ni + j];
        }
      }
      // extract_values: GID -> column LID lookup per dof (extract_values.hpp:54-63)
      for (int i = 0; i < nd; ++i) u[i] = u_col[rs.col_dof_lid.at(lm[i])];
      std::fill(fe.begin(), fe.end(), 0.0);
      if (action == 0) std::fill(Ke.begin(), Ke.end(), 0.0);
      const double* fib = ele_fiber ? ele_fiber + (fiber_gp_stride ? (size_t)fiber_gp_stride * num_gp(mesh.ct) : 3) * (size_t)e : nullptr;
      const bool shell = form == FORM_SHELL_ANS || form == FORM_SHELL_EAS_ANS;
      const int nshell = shell_eas_modes(mesh.ct, form);
      const int err = shell ? element_evaluate_shell(mesh.ct, nshell, mp, X.data(), u.data(),
                                  nshell ? eas_hist + (size_t)shell_history_doubles(nshell, nd) * e : nullptr, fe.data(),
                                  action == 0 ? Ke.data() : nullptr, fib, fiber_gp_stride)
                      : form_is_eas(form)
                          ? element_evaluate_eas(neas, mp, X.data(), u.data(), eas_hist + (size_t)eas_history_doubles(neas) * e, fe.data(),
                                action == 0 ? Ke.data() : nullptr, fib, fiber_gp_stride, nullptr, 0, 0, eas_linear(form))
                          : element_evaluate(mesh.ct, form, mp, X.data(), u.data(), fe.data(), action == 0 ? Ke.data() : nullptr, nullptr,
                                0.0, nullptr, STRESS_CAUCHY, fib, fiber_gp_stride,
                                mulf_hist ? mulf_hist + (size_t)18 * mulf_records(mesh.ct, form) * e : nullptr);
      if (err)
      {
        if (nbad == 0 && first_bad) *first_bad = eg;
        ++nbad;
        continue;
      }
      // matrix assemble (sparsematrix.cpp:444-576)
      if (action == 0 && K_values)
      {
        bool doit = false;
        for (int r = 0; r < nd; ++r)
          if (lmowner[r] == rs.rank)
          {
            doit = true;
            break;
          }
        if (doit)
        {
          for (int c = 0; c < nd; ++c) localcol[c] = rs.col_dof_lid.at(lm[c]);
          for (int lrow = 0; lrow < nd; ++lrow)
          {
            if (lmowner[lrow] != rs.rank) continue;
            const int rlid = rs.row_dof_lid.at(lm[lrow]);
            if (rs.dbc_row[rlid]) continue;
            const int64_t rb = rs.rowptr[rlid];
            const int length = (int)(rs.rowptr[rlid + 1] - rb);
            const int* indices = &rs.colind[rb];
            double* valview = &K_values[rb];
            int dofcount = 0, pos = 0;
            for (int node = 0; node < nen; ++node)
            {
              if (pos >= length || indices[pos] != localcol[dofcount])
                pos = (int)(std::lower_bound(indices, indices + length, localcol[dofcount]) - indices);
              for (int j = 0; j < 3; ++j) valview[pos++] += Ke[lrow + (size_t)nd * dofcount++];
            }
          }
        }
      }
      // vector assemble (sparse_algebra_assemble.cpp:72-92)
      if (f_row)
        for (int lrow = 0; lrow < nd; ++lrow)
        {
          if (lmowner[lrow] != rs.rank) continue;
          f_row[rs.row_dof_lid.at(lm[lrow])] += fe[lrow];
        }
    }
    return nbad;
  }

  // Whole job: all ranks, one thread per rank (mirrors mpirun -np nranks), capped at max_threads.
  struct Job
  {
    Mesh mesh;
    Partition part;
    std::vector<RankSystem> sys;
    int form = FORM_DISP;
    MaterialParams mp{MAT_NEOHOOKE, 10.0, 0.25};
    std::vector<double> ele_fiber;  // [n_ele][3], [n_ele][ngp][3] (fiber_gp_stride = 3) or empty
    int fiber_gp_stride = 0;
    std::vector<double> mulf_hist;  // [n_ele][ngp][18] when form == FORM_MULF
    mutable std::vector<double> eas_hist;  // [n_ele][eas_history_doubles] when form == FORM_EAS_MILD / FORM_EAS_FULL
  };

  inline int evaluate_job(const Job& job, int action, const std::vector<std::vector<double>>& u_row,
      std::vector<std::vector<double>>& f_row, std::vector<std::vector<double>>& K_values,
      int max_threads, int* first_bad)
  {
    const int nr = job.part.nranks;
    std::vector<int> nbad(nr, 0), fb(nr, -1);
    auto work = [&](int r)
    {
      std::vector<double> u_col;
      import_column_state(job.mesh, job.part, job.sys, r, u_row, u_col);
      nbad[r] = evaluate_rank(job.mesh, job.part, job.sys[r], job.form, job.mp, action,
          u_col.data(), f_row[r].data(), action == 0 ? K_values[r].data() : nullptr, &fb[r],
          job.ele_fiber.empty() ? nullptr : job.ele_fiber.data(), job.fiber_gp_stride,
          job.mulf_hist.empty() ? nullptr : job.mulf_hist.data(), job.eas_hist.empty() ? nullptr : job.eas_hist.data());
    };
    // EAS: every rank that holds an element writes its iteration data (the same values) - one rank at a time
    if (max_threads <= 1 || nr == 1 || !job.eas_hist.empty())
      for (int r = 0; r < nr; ++r) work(r);
    else
    {
      for (int r0 = 0; r0 < nr; r0 += max_threads)
      {
        std::vector<std::thread> th;
        for (int r = r0; r < std::min(nr, r0 + max_threads); ++r) th.emplace_back(work, r);
        for (auto& t : th) t.join();
      }
    }
    int total = 0;
    for (int r = 0; r < nr; ++r)
    {
      if (nbad[r] && total == 0 && first_bad) *first_bad = fb[r];
      total += nbad[r];
    }
    return total;
  }
}  // namespace fc_oracle
