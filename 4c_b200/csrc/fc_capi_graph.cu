// CSR graph of the rank's rows and the per-element scatter slot map.
//
// The reference builds the pattern implicitly: the first assembly into an unfilled matrix inserts
// every element coupling "zeros included" (src/core/linalg/src/sparse/4C_linalg_sparsematrix.cpp:
// 578-611), complete() freezes it (:843-865) and every later zero() re-creates the matrix on the
// saved graph (:353-378).  Here the pattern is built once from the connectivity as a node-block
// graph (every DOF row of a node couples to all 3 DOFs of each neighbour node, ascending local id)
// and the scatter positions are precomputed on the device.
#include <algorithm>
#include <climits>
#include <cstdlib>
#include <thread>

#include "fc_brick_plan.hpp"
#include "fc_host_util.hpp"
#include "fc_kernels_setup.cuh"

void fc_free_brick_plan(fc_ctx* ctx);
void fc_mg_free(fc_ctx* ctx);  // fc_capi_mg.cu

namespace
{
  // node-block adjacency of the owned nodes through all column elements
  void build_node_graph(const fc_ctx* ctx, std::vector<int64_t>& nodeptr, std::vector<int>& nodeadj)
  {
    const int nen = ctx->nen, nrow = ctx->n_row_nodes, ne = ctx->n_ele;
    const int* conn = ctx->h_conn.data();
    // owned node -> elements (CSR)
    std::vector<int64_t> n2e_ptr((size_t)nrow + 1, 0);
    for (size_t i = 0; i < (size_t)ne * nen; ++i)
      if (conn[i] < nrow) n2e_ptr[conn[i] + 1]++;
    for (int n = 0; n < nrow; ++n) n2e_ptr[n + 1] += n2e_ptr[n];
    std::vector<int> n2e((size_t)n2e_ptr[nrow]);
    {
      std::vector<int64_t> cur(n2e_ptr.begin(), n2e_ptr.end() - 1);
      for (int e = 0; e < ne; ++e)
        for (int a = 0; a < nen; ++a)
        {
          const int n = conn[(size_t)e * nen + a];
          if (n < nrow) n2e[cur[n]++] = e;
        }
    }
    unsigned nt = std::max(1u, std::min(std::thread::hardware_concurrency(), 32u));
    if (nrow < 4096) nt = 1;
    std::vector<std::vector<int>> part_adj(nt);
    std::vector<int> cnt((size_t)nrow, 0);
    auto work = [&](unsigned t)
    {
      const int lo = (int)((int64_t)nrow * t / nt), hi = (int)((int64_t)nrow * (t + 1) / nt);
      std::vector<int>& out = part_adj[t];
      std::vector<int> tmp;
      for (int n = lo; n < hi; ++n)
      {
        tmp.clear();
        for (int64_t k = n2e_ptr[n]; k < n2e_ptr[n + 1]; ++k)
        {
          const int* en = conn + (size_t)n2e[k] * nen;
          tmp.insert(tmp.end(), en, en + nen);
        }
        std::sort(tmp.begin(), tmp.end());
        tmp.erase(std::unique(tmp.begin(), tmp.end()), tmp.end());
        cnt[n] = (int)tmp.size();
        out.insert(out.end(), tmp.begin(), tmp.end());
      }
    };
    if (nt == 1)
      work(0);
    else
    {
      std::vector<std::thread> th;
      for (unsigned t = 0; t < nt; ++t) th.emplace_back(work, t);
      for (auto& t : th) t.join();
    }
    nodeptr.assign((size_t)nrow + 1, 0);
    for (int n = 0; n < nrow; ++n) nodeptr[n + 1] = nodeptr[n] + cnt[n];
    nodeadj.resize((size_t)nodeptr[nrow]);
    size_t off = 0;
    for (unsigned t = 0; t < nt; ++t)
    {
      std::copy(part_adj[t].begin(), part_adj[t].end(), nodeadj.begin() + off);
      off += part_adj[t].size();
    }
  }

  // hex8: units of node lines for the owner-computed assembly (fc_brick_plan.hpp); a mesh that is not a lattice of
  // hex cells has no plan and keeps the per-element reductions
  int build_brick_plan(fc_ctx* ctx)
  {
    if (ctx->brick_tried) return FC_OK;
    ctx->brick_tried = true;
    if (ctx->celltype != FC_HEX8 || ctx->n_ele == 0 || ctx->n_row_nodes == 0 || !ctx->have_graph) return FC_OK;
    int maxnb = 0;
    for (int n = 0; n < ctx->n_row_nodes; ++n) maxnb = std::max(maxnb, (int)(ctx->h_nodeptr[n + 1] - ctx->h_nodeptr[n]));
    std::vector<double> X((size_t)3 * ctx->n_col_nodes);
    FC_CUDA(cudaMemcpyAsync(X.data(), ctx->d_X, sizeof(double) * X.size(), cudaMemcpyDeviceToHost, ctx->stream));
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    fc_host::BrickPlanHost plan;
    if (!fc_host::build_brick_plan_host(ctx->n_col_nodes, X.data(), ctx->n_ele, ctx->h_conn.data(), ctx->n_row_nodes, maxnb, 0, plan)) return FC_OK;
    static_assert(sizeof(long long) == sizeof(int64_t), "");
    FC_CUDA(dev_upload(ctx->d_brick_planes, plan.unit_planes.data(), plan.unit_planes.size(), ctx->stream));
    FC_CUDA(dev_upload(ctx->d_brick_node_off, plan.unit_node_off.data(), plan.unit_node_off.size(), ctx->stream));
    FC_CUDA(dev_upload(ctx->d_brick_ele_off, plan.unit_ele_off.data(), plan.unit_ele_off.size(), ctx->stream));
    FC_CUDA(dev_upload(ctx->d_brick_nodes, plan.node_tab.data(), plan.node_tab.size(), ctx->stream));
    FC_CUDA(dev_upload(ctx->d_brick_eles, plan.ele_tab.data(), plan.ele_tab.size(), ctx->stream));
    FC_CUDA(dev_upload(ctx->d_brick_rows, plan.rowidx_tab.data(), plan.rowidx_tab.size(), ctx->stream));
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->n_brick_units = plan.n_units;
    ctx->n_brick_interior = plan.n_interior_units;
    return FC_OK;
  }

  int finalize_graph(fc_ctx* ctx)
  {
    const int nrow = ctx->n_row_nodes, ncol = ctx->n_col_nodes;
    // whatever was derived from the previous graph goes before anything can fail
    ctx->have_graph = false;
    fc_free_brick_plan(ctx);
    fc_mg_free(ctx);
    dev_free(ctx->d_Dinv);
    ctx->Dinv_cap = 0;
    ctx->nnz = 9 * ctx->h_nodeptr[nrow];
    ctx->n_row_dof = 3 * nrow;
    ctx->n_col_dof = 3 * ncol;
    std::vector<int64_t> rowbase((size_t)ncol, -1);
    std::vector<int> rowlen((size_t)ncol, 0);
    int maxnb = 0;
    for (int n = 0; n < nrow; ++n)
    {
      const int L = (int)(ctx->h_nodeptr[n + 1] - ctx->h_nodeptr[n]);
      rowbase[n] = 9 * ctx->h_nodeptr[n];
      rowlen[n] = 3 * L;
      maxnb = std::max(maxnb, L);
    }
    if (maxnb > 254) return fail(ctx, FC_ERR_UNSUPPORTED, "a node couples to more than 254 nodes (8-bit slot map)");
    FC_CUDA(cudaSetDevice(ctx->device));
    FC_CUDA(dev_upload(ctx->d_node_rowbase, rowbase.data(), rowbase.size(), ctx->stream));
    FC_CUDA(dev_upload(ctx->d_node_rowlen, rowlen.data(), rowlen.size(), ctx->stream));
    static_assert(sizeof(long long) == sizeof(int64_t), "");
    FC_CUDA(dev_upload(ctx->d_nodeptr, (const long long*)ctx->h_nodeptr.data(), ctx->h_nodeptr.size(), ctx->stream));
    FC_CUDA(dev_upload(ctx->d_nodeadj, ctx->h_nodeadj.data(), ctx->h_nodeadj.size(), ctx->stream));
    long long* const d_nodeptr = ctx->d_nodeptr;
    int* const d_nodeadj = ctx->d_nodeadj;
    {
      // position of every owned node in its own neighbour list (diagonal block: apply_dirichlet, block Jacobi)
      std::vector<int> diagpos((size_t)nrow, 0);
      for (int n = 0; n < nrow; ++n)
      {
        const int* b = ctx->h_nodeadj.data() + ctx->h_nodeptr[n];
        const int* e = ctx->h_nodeadj.data() + ctx->h_nodeptr[n + 1];
        const int* p = std::lower_bound(b, e, n);
        diagpos[n] = (p != e && *p == n) ? (int)(p - b) : -1;
      }
      FC_CUDA(dev_upload(ctx->d_node_diagpos, diagpos.data(), diagpos.size(), ctx->stream));
      FC_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    dev_free(ctx->d_toggle_rows);  // refers to rows of the previous graph
    ctx->n_toggle = 0;
    ctx->h_toggle_rows.clear();
    ctx->toggle_version++;
    dev_free(ctx->d_slots);
    const size_t nslots = (size_t)ctx->n_ele * ctx->nen * ctx->nen;
    int bad = 0;
    if (nslots > 0)
    {
      FC_CUDA(cudaMalloc((void**)&ctx->d_slots, nslots));
      ctx->h_err_pinned[0] = 0;
      ctx->h_err_pinned[1] = INT_MAX;
      FC_CUDA(cudaMemcpyAsync(ctx->d_err, ctx->h_err_pinned, 2 * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
      const long long nthr = (long long)ctx->n_ele * ctx->nen;
      const int bs = 256;
      fc::k_build_slots<<<(unsigned)((nthr + bs - 1) / bs), bs, 0, ctx->stream>>>(
          ctx->d_conn, ctx->n_ele, ctx->nen, nrow, d_nodeptr, d_nodeadj, ctx->d_slots, ctx->d_err);
      ctx->launches++;
      FC_CUDA(cudaGetLastError());
      FC_CUDA(cudaMemcpyAsync(ctx->h_err_pinned, ctx->d_err, 2 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
      FC_CUDA(cudaStreamSynchronize(ctx->stream));
      bad = ctx->h_err_pinned[0];
    }
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    // Dirichlet flags refer to rows of the previous graph
    dev_free(ctx->d_dbc);
    ctx->have_dbc = false;
    if (bad > 0) return fail(ctx, FC_ERR_GRAPH, "elements couple DOFs that are missing from the CSR graph");
    ctx->have_graph = true;
    return FC_OK;
  }
}  // namespace

int fc_build_brick_plan(fc_ctx* ctx) { return build_brick_plan(ctx); }

void fc_free_brick_plan(fc_ctx* ctx)
{
  dev_free(ctx->d_brick_planes);
  dev_free(ctx->d_brick_node_off);
  dev_free(ctx->d_brick_ele_off);
  dev_free(ctx->d_brick_nodes);
  dev_free(ctx->d_brick_eles);
  dev_free(ctx->d_brick_rows);
  ctx->n_brick_units = ctx->n_brick_interior = 0;
  ctx->brick_tried = false;
  ctx->split_store = false;
}

extern "C"
{
  int fc_solid_build_graph(fc_ctx* ctx)
  {
    if (!ctx || ctx->celltype < 0) return fail(ctx, FC_ERR_INVALID, "set_mesh must precede build_graph");
    build_node_graph(ctx, ctx->h_nodeptr, ctx->h_nodeadj);
    return finalize_graph(ctx);
  }

  int fc_solid_set_graph(fc_ctx* ctx, int n_row_dof, const int64_t* rowptr, const int* colind, int n_col_dof)
  {
    if (!ctx || ctx->celltype < 0 || !rowptr || !colind) return fail(ctx, FC_ERR_INVALID, "set_mesh must precede set_graph");
    if (n_row_dof != 3 * ctx->n_row_nodes || n_col_dof != 3 * ctx->n_col_nodes)
      return fail(ctx, FC_ERR_INVALID, "graph dimensions do not match the mesh (3 DOFs per node)");
    const int nrow = ctx->n_row_nodes;
    std::vector<int64_t> nodeptr((size_t)nrow + 1, 0);
    std::vector<int> nodeadj;
    nodeadj.reserve((size_t)(rowptr[3 * (size_t)nrow] / 9));
    for (int n = 0; n < nrow; ++n)
    {
      const int64_t b0 = rowptr[3 * n], len = rowptr[3 * n + 1] - b0;
      if (len % 3 != 0 || rowptr[3 * n + 2] - rowptr[3 * n + 1] != len || rowptr[3 * n + 3] - rowptr[3 * n + 2] != len)
        return fail(ctx, FC_ERR_GRAPH, "graph is not node-blocked: the 3 DOF rows of a node differ in length");
      for (int64_t k = 0; k < len; k += 3)
      {
        const int c = colind[b0 + k];
        if (c % 3 != 0 || colind[b0 + k + 1] != c + 1 || colind[b0 + k + 2] != c + 2 || c < 0 || c + 2 >= n_col_dof)
          return fail(ctx, FC_ERR_GRAPH, "graph is not node-blocked: a row does not list complete node triples");
        if (k > 0 && colind[b0 + k - 1] >= c) return fail(ctx, FC_ERR_GRAPH, "column ids must ascend within a row");
        if (colind[b0 + len + k] != c || colind[b0 + 2 * len + k] != c)
          return fail(ctx, FC_ERR_GRAPH, "graph is not node-blocked: the 3 DOF rows of a node differ");
        nodeadj.push_back(c / 3);
      }
      nodeptr[n + 1] = (int64_t)nodeadj.size();
    }
    ctx->h_nodeptr.swap(nodeptr);
    ctx->h_nodeadj.swap(nodeadj);
    return finalize_graph(ctx);
  }

  int fc_solid_graph_sizes(const fc_ctx* ctx, int* n_row_dof, int* n_col_dof, int64_t* nnz)
  {
    if (!ctx || !ctx->have_graph) return FC_ERR_INVALID;
    if (n_row_dof) *n_row_dof = ctx->n_row_dof;
    if (n_col_dof) *n_col_dof = ctx->n_col_dof;
    if (nnz) *nnz = ctx->nnz;
    return FC_OK;
  }

  int fc_solid_get_graph(const fc_ctx* ctx, int64_t* rowptr, int* colind)
  {
    if (!ctx || !ctx->have_graph || !rowptr || !colind) return FC_ERR_INVALID;
    const int nrow = ctx->n_row_nodes;
    for (int n = 0; n < nrow; ++n)
    {
      const int64_t p0 = ctx->h_nodeptr[n], L = ctx->h_nodeptr[n + 1] - p0;
      for (int i = 0; i < 3; ++i)
      {
        const int64_t rb = 9 * p0 + (int64_t)i * 3 * L;
        rowptr[3 * n + i] = rb;
        for (int64_t k = 0; k < L; ++k)
        {
          const int c = 3 * ctx->h_nodeadj[p0 + k];
          colind[rb + 3 * k + 0] = c;
          colind[rb + 3 * k + 1] = c + 1;
          colind[rb + 3 * k + 2] = c + 2;
        }
      }
    }
    rowptr[3 * nrow] = ctx->nnz;
    return FC_OK;
  }

  int fc_solid_set_dirichlet_rows(fc_ctx* ctx, int n, const int* row_lids)
  {
    if (!ctx || !ctx->have_graph || n < 0 || (n > 0 && !row_lids)) return fail(ctx, FC_ERR_INVALID, "graph required");
    FC_CUDA(cudaSetDevice(ctx->device));
    dev_free(ctx->d_dbc);
    ctx->have_dbc = false;
    if (n == 0) return FC_OK;
    std::vector<unsigned char> flags((size_t)ctx->n_row_dof, 0);
    for (int i = 0; i < n; ++i)
    {
      if (row_lids[i] < 0 || row_lids[i] >= ctx->n_row_dof) return fail(ctx, FC_ERR_INVALID, "Dirichlet row out of range");
      flags[row_lids[i]] = 1;
    }
    FC_CUDA(dev_upload(ctx->d_dbc, flags.data(), flags.size(), ctx->stream));
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->have_dbc = true;
    return FC_OK;
  }
}
