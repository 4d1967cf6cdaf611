// libfourc_b200.so — C ABI implementation (host side).  See include/fourc_b200.h.
// There is no CPU fallback anywhere in this file: every compute entry point launches CUDA
// kernels on the context's device or fails.
#include <algorithm>
#include <cmath>
#include <climits>
#include <cstdio>
#include <cstring>
#include <thread>

#include "../../include/fourc_b200.h"
#include "fc_context.hpp"
#include "fc_host_util.hpp"
#include "fc_kernels_setup.cuh"
#include "fc_sweep_order.hpp"

void fc_free_brick_plan(fc_ctx* ctx);  // fc_capi_graph.cu
void fc_mg_free(fc_ctx* ctx);          // fc_capi_mg.cu
int fc_mulf_init(fc_ctx* ctx);         // fc_capi_eval.cu

namespace
{
  // nodal fibres -> Gauss points of the stiffness rule, normalised (project_fibers_to_gauss_points,
  // src/core/fem/src/general/node/4C_fem_general_fiber_node_utils.cpp:18-75)
  template <int CT>
  int project_nodal_fibers(const fc_ctx* ctx, const double* node_fiber, std::vector<double>& out)
  {
    constexpr int NEN = fc::Cell<CT>::nen, NGP = fc::Cell<CT>::ngp;
    double N[NGP][NEN];
    for (int g = 0; g < NGP; ++g)
    {
      double xi[3], w;
      fc::gauss_point<CT>(g, xi, &w);
      fc::shape_function<CT>(xi, N[g]);
    }
    out.assign((size_t)3 * NGP * ctx->n_ele, 0.0);
    for (int e = 0; e < ctx->n_ele; ++e)
      for (int g = 0; g < NGP; ++g)
      {
        double v[3] = {0.0, 0.0, 0.0};
        for (int a = 0; a < NEN; ++a)
          for (int i = 0; i < 3; ++i) v[i] += N[g][a] * node_fiber[3 * (size_t)ctx->h_conn[(size_t)e * NEN + a] + i];
        const double nrm = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        if (!(nrm > 0.0)) return 1;
        for (int i = 0; i < 3; ++i) out[((size_t)e * NGP + g) * 3 + i] = v[i] * (1.0 / nrm);
      }
    return 0;
  }
}  // namespace

extern "C"
{
  const char* fc_version(void) { return "fourc_b200 0.1 (sm_100a)"; }

  const char* fc_last_error(const fc_ctx* ctx) { return ctx ? ctx->last_error.c_str() : g_create_error.c_str(); }

  int fc_solid_create(fc_ctx** out, int device, int rank, int nranks)
  {
    if (!out || nranks < 1 || rank < 0 || rank >= nranks) return FC_ERR_INVALID;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    {
      cudaGetLastError();
      g_create_error = "no CUDA device visible: libfourc_b200 has no CPU fallback";
      return FC_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= ndev) return FC_ERR_INVALID;
    if (cudaSetDevice(device) != cudaSuccess) return FC_ERR_CUDA;
    fc_ctx* ctx = new fc_ctx;
    ctx->device = device;
    ctx->rank = rank;
    ctx->nranks = nranks;
    if (cudaMalloc((void**)&ctx->d_err, 3 * sizeof(int)) != cudaSuccess ||
        cudaMallocHost((void**)&ctx->h_err_pinned, 2 * sizeof(int)) != cudaSuccess)
    {
      delete ctx;
      return FC_ERR_CUDA;
    }
    for (auto& t : ctx->ev_hist)
      for (auto& e : t) cudaEventCreate(&e);
    *out = ctx;
    return FC_OK;
  }

  void fc_solid_destroy(fc_ctx* ctx)
  {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    dev_free(ctx->d_X);
    dev_free(ctx->d_conn);
    dev_free(ctx->d_ele_gid);
    dev_free(ctx->d_dNtab);
    dev_free(ctx->d_gpw);
    dev_free(ctx->d_nurbs_knots);
    dev_free(ctx->d_nurbs_weights);
    dev_free(ctx->d_ele_bad);
    dev_free(ctx->d_geom);
    dev_free(ctx->d_fiber);
    dev_free(ctx->d_mulf);
    dev_free(ctx->d_eas);
    dev_free(ctx->d_ele_perm);
    dev_free(ctx->d_node_rowbase);
    dev_free(ctx->d_node_rowlen);
    dev_free(ctx->d_slots);
    fc_free_brick_plan(ctx);
    fc_mg_free(ctx);
    dev_free(ctx->d_dbc);
    dev_free(ctx->d_nodeptr);
    dev_free(ctx->d_nodeadj);
    dev_free(ctx->d_node_diagpos);
    dev_free(ctx->d_toggle_rows);
    dev_free(ctx->d_solver_ws);
    dev_free(ctx->d_Dinv);
    dev_free(ctx->d_dot);
    if (ctx->h_dot_pinned) cudaFreeHost(ctx->h_dot_pinned);
    dev_free(ctx->d_send_lids);
    dev_free(ctx->d_halo_sendbuf);
    dev_free(ctx->d_stage_u);
    dev_free(ctx->d_stage_f);
    dev_free(ctx->d_stage_K);
    dev_free(ctx->d_err);
    if (ctx->h_err_pinned) cudaFreeHost(ctx->h_err_pinned);
    for (auto& t : ctx->ev_hist)
      for (auto& e : t)
        if (e) cudaEventDestroy(e);
    delete ctx;
  }

  int fc_solid_set_stream(fc_ctx* ctx, void* cuda_stream)
  {
    if (!ctx) return FC_ERR_INVALID;
    ctx->stream = (cudaStream_t)cuda_stream;
    return FC_OK;
  }

  int fc_solid_set_material(fc_ctx* ctx, int kind, const double* params, int n_params)
  {
    if (!ctx || !params || n_params < 2) return FC_ERR_INVALID;
    if (kind != FC_MAT_STVENANTKIRCHHOFF && kind != FC_MAT_ELASTHYPER_COUPNEOHOOKE)
      return fail(ctx, FC_ERR_UNSUPPORTED, "material kind not on the path");
    const double young = params[0], nue = params[1];
    // parameter checks of the reference's input spec: YOUNG > 0, NUE in (-1, 0.5)
    if (!(young > 0.0) || !(nue > -1.0) || !(nue < 0.5)) return fail(ctx, FC_ERR_INVALID, "YOUNG must be > 0 and NUE in (-1, 0.5)");
    ctx->mat.kind = kind == FC_MAT_STVENANTKIRCHHOFF ? fc::MAT_SVK : fc::MAT_NEOHOOKE;
    ctx->mat.young = young;
    ctx->mat.nue = nue;
    ctx->have_material = true;
    return FC_OK;
  }

  int fc_solid_set_material_elasthyper(fc_ctx* ctx, int n_summands, const int* kinds, const double* params)
  {
    if (!ctx || !kinds || !params || n_summands < 1) return FC_ERR_INVALID;
    if (n_summands > fc::max_summands) return fail(ctx, FC_ERR_UNSUPPORTED, "MAT_ElastHyper with more than 4 summands");
    fc::MatParams m = ctx->mat;
    m.kind = fc::MAT_ELASTHYPER;
    m.n_summands = n_summands;
    bool iso = false;
    for (int k = 0; k < n_summands; ++k)
    {
      const double* p = params + 4 * k;
      if (kinds[k] == FC_SUMMAND_COUPNEOHOOKE || kinds[k] == FC_SUMMAND_COUPSVK)
      {
        // YOUNG > 0; NUE in (-1, 0.5) for CoupNeoHooke (beta = nue / (1 - 2 nue)), (-1, 0.5] for CoupSVK (coupSaintVenantKirchhoff.cpp:21-28)
        const bool nue_ok = p[1] > -1.0 && (kinds[k] == FC_SUMMAND_COUPSVK ? p[1] <= 0.5 : p[1] < 0.5);
        if (!(p[0] > 0.0) || !nue_ok) return fail(ctx, FC_ERR_INVALID, "summand: YOUNG must be > 0 and NUE inside its range");
        iso = true;
      }
      else if (kinds[k] == FC_SUMMAND_COUPANISOEXPO)
      {
        if (p[0] < 0.0 || p[2] < 0.0) return fail(ctx, FC_ERR_INVALID, "ELAST_CoupAnisoExpo: K1 and K1COMP must be >= 0");
      }
      else
        return fail(ctx, FC_ERR_UNSUPPORTED, "summand kind not on the path");
      m.sum_kind[k] = kinds[k];
      for (int j = 0; j < 4; ++j) m.sum_par[k][j] = p[j];
    }
    if (!iso) return fail(ctx, FC_ERR_UNSUPPORTED, "MAT_ElastHyper needs at least one isotropic summand here");
    // one CoupNeoHooke summand is the material of the closed-form kernels
    if (n_summands == 1 && kinds[0] == FC_SUMMAND_COUPNEOHOOKE)
    {
      m.kind = fc::MAT_NEOHOOKE;
      m.young = params[0];
      m.nue = params[1];
    }
    ctx->mat = m;
    ctx->have_material = true;
    return FC_OK;
  }

  int fc_solid_set_fibers(fc_ctx* ctx, const double* fiber1)
  {
    if (!ctx || ctx->celltype < 0) return fail(ctx, FC_ERR_INVALID, "set_mesh must precede set_fibers");
    FC_CUDA(cudaSetDevice(ctx->device));
    dev_free(ctx->d_fiber);
    ctx->fiber_gp = 0;
    if (!fiber1) return FC_OK;
    std::vector<double> f((size_t)3 * ctx->n_ele);
    for (int e = 0; e < ctx->n_ele; ++e)
    {
      // read_anisotropy_fiber (src/mat/4C_mat_anisotropy_utils.cpp:19-44): normalised, a zero vector is an error
      const double* v = fiber1 + 3 * (size_t)e;
      const double nrm = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
      if (nrm < 1e-9) return fail(ctx, FC_ERR_INVALID, "The given fiber is not a vector but zero.");
      for (int i = 0; i < 3; ++i) f[3 * (size_t)e + i] = v[i] / nrm;
    }
    FC_CUDA(dev_upload(ctx->d_fiber, f.data(), f.size(), ctx->stream));
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    return FC_OK;
  }

  int fc_solid_set_nodal_fibers(fc_ctx* ctx, const double* node_fiber1)
  {
    if (!ctx || ctx->celltype < 0) return fail(ctx, FC_ERR_INVALID, "set_mesh must precede set_nodal_fibers");
    FC_CUDA(cudaSetDevice(ctx->device));
    dev_free(ctx->d_fiber);
    ctx->fiber_gp = 0;
    if (!node_fiber1) return FC_OK;
    if (ctx->celltype == FC_NURBS27) return fail(ctx, FC_ERR_UNSUPPORTED, "nodal fibres on nurbs27");
    std::vector<double> f;
    int bad = 0;
    switch (ctx->celltype)
    {
      case FC_HEX8: bad = project_nodal_fibers<fc::HEX8>(ctx, node_fiber1, f); break;
      case FC_HEX27: bad = project_nodal_fibers<fc::HEX27>(ctx, node_fiber1, f); break;
      case FC_TET10: bad = project_nodal_fibers<fc::TET10>(ctx, node_fiber1, f); break;
      case FC_HEX20: bad = project_nodal_fibers<fc::HEX20>(ctx, node_fiber1, f); break;
      case FC_TET4: bad = project_nodal_fibers<fc::TET4>(ctx, node_fiber1, f); break;
      case FC_WEDGE6: bad = project_nodal_fibers<fc::WEDGE6>(ctx, node_fiber1, f); break;
      case FC_HEX18: bad = project_nodal_fibers<fc::HEX18>(ctx, node_fiber1, f); break;
      default: bad = project_nodal_fibers<fc::PYRAMID5>(ctx, node_fiber1, f); break;
    }
    if (bad) return fail(ctx, FC_ERR_INVALID, "nodal fibres interpolate to a zero vector at a Gauss point");
    FC_CUDA(dev_upload(ctx->d_fiber, f.data(), f.size(), ctx->stream));
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->fiber_gp = 3;
    return FC_OK;
  }

  int fc_solid_set_mesh(fc_ctx* ctx, int celltype, int formulation, int n_col_nodes, const double* X,
      const int* node_gid, const int* node_owner, int n_col_ele, const int* conn, const int* ele_gid)
  {
    if (!ctx || !X || !node_owner || !conn || !ele_gid || n_col_nodes <= 0 || n_col_ele < 0) return FC_ERR_INVALID;
    (void)node_gid;
    if (celltype < FC_HEX8 || celltype > FC_NURBS27) return fail(ctx, FC_ERR_UNSUPPORTED, "cell type not on the path");
    if (formulation != FC_FORM_DISPLACEMENT && formulation != FC_FORM_LINEAR && formulation != FC_FORM_MULF &&
        !(formulation == FC_FORM_FBAR && (celltype == FC_HEX8 || celltype == FC_PYRAMID5)) &&
        !(formulation == FC_FORM_MULF_FBAR && (celltype == FC_HEX8 || celltype == FC_PYRAMID5)) &&
        !((formulation == FC_FORM_SHELL_ANS || formulation == FC_FORM_SHELL_EAS_ANS) && (celltype == FC_HEX8 || celltype == FC_WEDGE6)) &&
        !((formulation == FC_FORM_EAS_MILD || formulation == FC_FORM_EAS_FULL || formulation == FC_FORM_EAS_SHELL ||
              formulation == FC_FORM_EAS_MILD_LINEAR || formulation == FC_FORM_EAS_FULL_LINEAR) &&
            celltype == FC_HEX8))
      return fail(ctx, FC_ERR_UNSUPPORTED, "F-bar exists for hex8 and pyramid5, EAS for hex8, solid shells for hex8 and wedge6 only (4C_solid_3D_ele_factory.hpp:49-75)");
    FC_CUDA(cudaSetDevice(ctx->device));
    const int nen = fc::cell_nen(celltype);
    // owned nodes first (Epetra column-map convention)
    int n_row = 0;
    while (n_row < n_col_nodes && node_owner[n_row] == ctx->rank) ++n_row;
    for (int i = n_row; i < n_col_nodes; ++i)
    {
      if (node_owner[i] == ctx->rank) return fail(ctx, FC_ERR_INVALID, "owned nodes must precede ghost nodes");
      if (node_owner[i] < 0 || node_owner[i] >= ctx->nranks) return fail(ctx, FC_ERR_INVALID, "node owner out of range");
      // the halo import lands rank p's message behind the messages of the ranks below p (fc_solid_halo_exchange)
      if (i > n_row && node_owner[i] < node_owner[i - 1]) return fail(ctx, FC_ERR_INVALID, "ghost nodes must be grouped by ascending owner rank");
    }
    for (size_t i = 0; i < (size_t)n_col_ele * nen; ++i)
      if (conn[i] < 0 || conn[i] >= n_col_nodes) return fail(ctx, FC_ERR_INVALID, "connectivity entry out of range");
    ctx->celltype = celltype;
    ctx->formulation = formulation;
    ctx->nen = nen;
    ctx->n_col_nodes = n_col_nodes;
    ctx->n_row_nodes = n_row;
    ctx->n_ele = n_col_ele;
    ctx->h_conn.assign(conn, conn + (size_t)n_col_ele * nen);
    ctx->h_node_owner.assign(node_owner, node_owner + n_col_nodes);
    ctx->h_ele_gid.assign(ele_gid, ele_gid + n_col_ele);
    // everything that describes the previous mesh goes: graph, halo plan, solver state, plans
    ctx->have_graph = false;
    ctx->n_send = 0;
    ctx->halo_send_count.clear();
    ctx->halo_recv_count.clear();
    dev_free(ctx->d_Dinv);
    ctx->Dinv_cap = 0;
    ctx->n_toggle = 0;
    ctx->h_toggle_rows.clear();
    ctx->toggle_version++;
    fc_free_brick_plan(ctx);
    fc_mg_free(ctx);
    ctx->n_col_dof = 3 * n_col_nodes;
    ctx->n_row_dof = 3 * n_row;
    FC_CUDA(dev_upload(ctx->d_X, X, (size_t)3 * n_col_nodes, ctx->stream));
    FC_CUDA(dev_upload(ctx->d_conn, conn, (size_t)n_col_ele * nen, ctx->stream));
    FC_CUDA(dev_upload(ctx->d_ele_gid, ele_gid, (size_t)n_col_ele, ctx->stream));
    std::vector<double> dN, w;
    ctx->nurbs_ready = false;
    dev_free(ctx->d_nurbs_knots);
    dev_free(ctx->d_nurbs_weights);
    switch (celltype)
    {
      case FC_NURBS27:
        // Gauss weights only: the shape tables are per element and come with fc_solid_set_nurbs
        w.resize(27);
        for (int g = 0; g < 27; ++g)
        {
          double xi[3];
          fc::gauss_point<fc::NURBS27>(g, xi, &w[g]);
        }
        break;
      case FC_HEX8: fill_tables<fc::HEX8>(dN, w); break;
      case FC_HEX27: fill_tables<fc::HEX27>(dN, w); break;
      case FC_TET10: fill_tables<fc::TET10>(dN, w); break;
      case FC_HEX20: fill_tables<fc::HEX20>(dN, w); break;
      case FC_TET4: fill_tables<fc::TET4>(dN, w); break;
      case FC_WEDGE6: fill_tables<fc::WEDGE6>(dN, w); break;
      case FC_HEX18: fill_tables<fc::HEX18>(dN, w); break;
      default: fill_tables<fc::PYRAMID5>(dN, w); break;
    }
    FC_CUDA(dev_upload(ctx->d_dNtab, dN.data(), dN.size(), ctx->stream));
    FC_CUDA(dev_upload(ctx->d_gpw, w.data(), w.size(), ctx->stream));
    dev_free(ctx->d_ele_bad);
    dev_free(ctx->d_geom);
    dev_free(ctx->d_fiber);  // fibres belong to the elements of the previous mesh
    ctx->fiber_gp = 0;
    dev_free(ctx->d_ele_perm);
    {
      // launch-slot order of the tensor-core kernels.  Large meshes: visit the elements so that a row's contributions arrive
      // while the row is still in L2; a layer of the sweep may write about 24 MB of K values (hex8 2 KB, tet10 2.8 KB, hex27
      // 37 KB per element).  More than one rank: elements without ghost nodes first, so that they can be evaluated while the
      // halo import is in flight (FC_EVAL_INTERIOR, then FC_EVAL_BOUNDARY).
      std::vector<int> perm;
      const long long budget = celltype == FC_HEX8 ? 12000 : (celltype == FC_TET10 ? 8500 : (celltype == FC_HEX27 || celltype == FC_NURBS27 ? 650 : 2000));
      fc_host::l2_sweep_order(n_col_nodes, X, n_col_ele, nen, conn, budget, perm);
      ctx->n_interior_ele = n_col_ele;
      if (n_row < n_col_nodes)
      {
        if (perm.empty())
        {
          perm.resize((size_t)n_col_ele);
          for (int e = 0; e < n_col_ele; ++e) perm[e] = e;
        }
        auto interior = [&](int e) {
          for (int a = 0; a < nen; ++a)
            if (conn[(size_t)e * nen + a] >= n_row) return false;
          return true;
        };
        ctx->n_interior_ele = (int)(std::stable_partition(perm.begin(), perm.end(), interior) - perm.begin());
      }
      if (!perm.empty()) FC_CUDA(dev_upload(ctx->d_ele_perm, perm.data(), perm.size(), ctx->stream));
    }
    ctx->geom_bad_count = 0;
    ctx->geom_first_bad_gid = -1;
    if (n_col_ele > 0 && celltype == FC_NURBS27)
    {
      // node-Jacobian check and prestress history need the element tables: fc_solid_set_nurbs
      FC_CUDA(cudaMalloc((void**)&ctx->d_ele_bad, (size_t)n_col_ele));
      FC_CUDA(cudaMemsetAsync(ctx->d_ele_bad, 0, (size_t)n_col_ele, ctx->stream));
      ctx->h_err_pinned[0] = 0;
    }
    else if (n_col_ele > 0)
    {
      FC_CUDA(cudaMalloc((void**)&ctx->d_ele_bad, (size_t)n_col_ele));
      ctx->h_err_pinned[0] = 0;
      ctx->h_err_pinned[1] = INT_MAX;
      FC_CUDA(cudaMemcpyAsync(ctx->d_err, ctx->h_err_pinned, 2 * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
      const int bs = 128, gs = (n_col_ele + bs - 1) / bs;
#define FC_CHECK_NODE_JACOBIANS(CT) \
  fc::k_check_node_jacobians<CT><<<gs, bs, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, ctx->d_ele_gid, n_col_ele, ctx->d_dNtab, ctx->d_ele_bad, ctx->d_err)
      switch (celltype)
      {
        case FC_NURBS27: break;  // not reached
        case FC_HEX8: FC_CHECK_NODE_JACOBIANS(fc::HEX8); break;
        case FC_HEX27: FC_CHECK_NODE_JACOBIANS(fc::HEX27); break;
        case FC_TET10: FC_CHECK_NODE_JACOBIANS(fc::TET10); break;
        case FC_HEX20: FC_CHECK_NODE_JACOBIANS(fc::HEX20); break;
        case FC_TET4: FC_CHECK_NODE_JACOBIANS(fc::TET4); break;
        case FC_WEDGE6: FC_CHECK_NODE_JACOBIANS(fc::WEDGE6); break;
        case FC_HEX18: FC_CHECK_NODE_JACOBIANS(fc::HEX18); break;
        default: FC_CHECK_NODE_JACOBIANS(fc::PYRAMID5); break;
      }
#undef FC_CHECK_NODE_JACOBIANS
      ctx->launches++;
      FC_CUDA(cudaGetLastError());
      FC_CUDA(cudaMemcpyAsync(ctx->h_err_pinned, ctx->d_err, 2 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (int rc = fc_mulf_init(ctx)) return rc;
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    if (n_col_ele > 0 && ctx->h_err_pinned[0] > 0)
    {
      ctx->geom_bad_count = ctx->h_err_pinned[0];
      ctx->geom_first_bad_gid = ctx->h_err_pinned[1];
    }
    return FC_OK;
  }

  // Knot spans and control point weights of a nurbs27 mesh (what Core::FE::Nurbs::get_my_nurbs_knots_and_weights hands to every
  // element evaluation, 4C_solid_3D_ele_calc_lib.hpp:193-200): builds the per-element shape tables, then runs what set_mesh
  // postponed - the node-Jacobian check and the initial prestress history.
  int fc_solid_set_nurbs(fc_ctx* ctx, const double* ele_knots, const double* node_weights)
  {
    if (!ctx || ctx->celltype != FC_NURBS27) return fail(ctx, FC_ERR_INVALID, "set_mesh with FC_NURBS27 must precede set_nurbs");
    if (!ele_knots || !node_weights) return fail(ctx, FC_ERR_INVALID, "knots and weights are required");
    for (int i = 0; i < ctx->n_col_nodes; ++i)
      if (!(node_weights[i] > 0.0)) return fail(ctx, FC_ERR_INVALID, "control point weights must be positive");
    FC_CUDA(cudaSetDevice(ctx->device));
    ctx->nurbs_ready = false;
    const int n_ele = ctx->n_ele;
    FC_CUDA(dev_upload(ctx->d_nurbs_knots, ele_knots, (size_t)18 * n_ele, ctx->stream));
    FC_CUDA(dev_upload(ctx->d_nurbs_weights, node_weights, (size_t)ctx->n_col_nodes, ctx->stream));
    dev_free(ctx->d_dNtab);
    dev_free(ctx->d_geom);  // geometry cache of the tensor-core kernel: belongs to the previous tables
    ctx->geom_bad_count = 0;
    ctx->geom_first_bad_gid = -1;
    if (n_ele > 0)
    {
      FC_CUDA(cudaMalloc((void**)&ctx->d_dNtab, sizeof(double) * (size_t)fc::NurbsTab::stride * n_ele));
      FC_CUDA(cudaMemsetAsync(ctx->d_dNtab, 0, sizeof(double) * (size_t)fc::NurbsTab::stride * n_ele, ctx->stream));
      FC_CUDA(cudaMemsetAsync(ctx->d_ele_bad, 0, (size_t)n_ele, ctx->stream));
      unsigned char* const d_zero = ctx->d_ele_bad;  // scratch for the zero-size flags; the node-Jacobian check below rewrites it
      const long long nt = (long long)n_ele * (fc::NurbsTab::ngp + fc::NurbsTab::nen + 1);
      fc::k_nurbs_tables<<<(unsigned)((nt + 127) / 128), 128, 0, ctx->stream>>>(ctx->d_nurbs_knots, ctx->d_nurbs_weights, ctx->d_conn, n_ele,
          ctx->d_dNtab, d_zero);
      std::vector<unsigned char> zero((size_t)n_ele);
      cudaError_t e = cudaGetLastError();
      if (e == cudaSuccess) e = cudaMemcpyAsync(zero.data(), d_zero, (size_t)n_ele, cudaMemcpyDeviceToHost, ctx->stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
      FC_CUDA(e);
      for (int el = 0; el < n_ele; ++el)
        if (zero[el]) return fail(ctx, FC_ERR_INVALID, "nurbs27 element with a zero-sized knot span (the reference throws, 4C_solid_3D_ele_calc_lib.hpp:196-199)");
      ctx->h_err_pinned[0] = 0;
      ctx->h_err_pinned[1] = INT_MAX;
      FC_CUDA(cudaMemcpyAsync(ctx->d_err, ctx->h_err_pinned, 2 * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
      fc::k_check_node_jacobians<fc::NURBS27><<<(n_ele + 127) / 128, 128, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, ctx->d_ele_gid, n_ele, ctx->d_dNtab,
          ctx->d_ele_bad, ctx->d_err);
      ctx->launches += 2;
      FC_CUDA(cudaGetLastError());
      FC_CUDA(cudaMemcpyAsync(ctx->h_err_pinned, ctx->d_err, 2 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
      FC_CUDA(cudaStreamSynchronize(ctx->stream));
      if (ctx->h_err_pinned[0] > 0)
      {
        ctx->geom_bad_count = ctx->h_err_pinned[0];
        ctx->geom_first_bad_gid = ctx->h_err_pinned[1];
      }
    }
    ctx->nurbs_ready = true;
    if (int rc = fc_mulf_init(ctx)) return rc;
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    return FC_OK;
  }
}
