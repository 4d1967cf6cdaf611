// fc_solid_neumann_*: Neumann loads of TYPE Live on the device, the external force of
// Solid::ModelEvaluator::Structure::apply_force_external -> Discretization::evaluate_neumann
// (src/structure_new/src/model_evaluator/4C_structure_new_model_evaluator_structure.cpp:368-396).
#include "fc_host_util.hpp"
#include "fc_kernels_neumann.cuh"

namespace
{
  // device copy of a host|device array of n elements; tmp is set when a staging buffer was allocated
  template <class T>
  int stage_in(fc_ctx* ctx, const T* p, size_t n, const T** dev, T** tmp)
  {
    *tmp = nullptr;
    *dev = p;
    if (!p || n == 0 || is_device_ptr(p)) return FC_OK;
    FC_CUDA(cudaMalloc((void**)tmp, n * sizeof(T)));
    FC_CUDA(cudaMemcpyAsync(*tmp, p, n * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    *dev = *tmp;
    return FC_OK;
  }

  int check_val(fc_ctx* ctx, const int* onoff, const double* val, fc::NeumannVal& V)
  {
    if (!onoff || !val) return fail(ctx, FC_ERR_INVALID, "ONOFF and VAL are required");
    for (int i = 0; i < 3; ++i)
    {
      V.onoff[i] = onoff[i];
      V.val[i] = val[i];
    }
    return FC_OK;
  }

  // shared body: mode 0 = add the load into f_ext, 1 = hand out the Gauss-point reference coordinates
  int neumann_run(fc_ctx* ctx, bool surface, int face_type, int n_items, const int* items, const int* onoff, const double* val,
      const double* gp_factor, double* out, bool coords, const int* surface_id = nullptr /* faces of nurbs27: items = parent elements */,
      bool ortho = false, const double* u_col = nullptr /* orthopressure: the configuration is X + u */)
  {
    const bool nurbs_face = surface && surface_id != nullptr;
    if (!ctx || ctx->celltype < 0) return fail(ctx, FC_ERR_INVALID, "set_mesh must precede the Neumann evaluation");
    if (n_items < 0 || (n_items > 0 && !items) || !out) return fail(ctx, FC_ERR_INVALID, "item list and output are required");
    if (surface && !nurbs_face && (face_type < FC_FACE_QUAD4 || face_type > FC_FACE_QUAD6))
      return fail(ctx, FC_ERR_UNSUPPORTED, "surface cell type");
    if ((nurbs_face || (!surface && ctx->celltype == FC_NURBS27)) && (ctx->celltype != FC_NURBS27 || !ctx->nurbs_ready))
      return fail(ctx, FC_ERR_INVALID, "nurbs27 mesh with fc_solid_set_nurbs required");
    if (surface && !nurbs_face && ctx->celltype == FC_NURBS27)
      return fail(ctx, FC_ERR_INVALID, "faces of nurbs27 elements: fc_solid_neumann_surface_nurbs");
    fc::NeumannVal V{};
    if (!coords)
      if (int rc = check_val(ctx, onoff, val, V)) return rc;
    FC_CUDA(cudaSetDevice(ctx->device));
    const int per = surface && !nurbs_face ? fc::face_nen(face_type) : 1;
    const int ngp = nurbs_face ? 9 : (surface ? fc::face_ngp(face_type) : fc::cell_ngp(ctx->celltype));
    for (size_t i = 0; i < (size_t)n_items * per; ++i)
      if (items[i] < 0 || items[i] >= (surface && !nurbs_face ? ctx->n_col_nodes : ctx->n_ele))
        return fail(ctx, FC_ERR_INVALID, surface && !nurbs_face ? "face node out of range" : "element out of range");
    if (nurbs_face)
      for (int i = 0; i < n_items; ++i)
        if (surface_id[i] < 0 || surface_id[i] > 5) return fail(ctx, FC_ERR_INVALID, "surface number of a nurbs27 element: 0..5");
    const size_t n_out = coords ? (size_t)n_items * ngp * 3 : (size_t)ctx->n_row_dof;
    if (n_items == 0) return FC_OK;
    const int* d_items = nullptr;
    int* t_items = nullptr;
    FC_CUDA(dev_upload(t_items, items, (size_t)n_items * per, ctx->stream));
    d_items = t_items;
    int* t_sid = nullptr;
    if (nurbs_face) FC_CUDA(dev_upload(t_sid, surface_id, (size_t)n_items, ctx->stream));
    const double* d_u = nullptr;
    double* t_u = nullptr;
    const double* d_fac = nullptr;
    double* t_fac = nullptr;
    int rc = ortho && u_col ? stage_in(ctx, u_col, (size_t)ctx->n_col_dof, &d_u, &t_u) : FC_OK;
    if (rc == FC_OK && !coords) rc = stage_in(ctx, gp_factor, (size_t)n_items * ngp * 3, &d_fac, &t_fac);
    double *d_out = out, *t_out = nullptr;
    const bool out_dev = is_device_ptr(out);
    cudaError_t e = cudaSuccess;
    if (rc == FC_OK && !out_dev)
    {
      e = cudaMalloc((void**)&t_out, n_out * sizeof(double));
      if (e == cudaSuccess && !coords) e = cudaMemcpyAsync(t_out, out, n_out * sizeof(double), cudaMemcpyHostToDevice, ctx->stream);  // += semantics
      d_out = t_out;
    }
    if (rc == FC_OK && e == cudaSuccess)
    {
      const int bs = 128;
      double* f = coords ? nullptr : d_out;
      double* xc = coords ? d_out : nullptr;
      if (nurbs_face)
        fc::k_neumann_surface_nurbs<<<(unsigned)((n_items + bs - 1) / bs), bs, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, ctx->d_nurbs_knots,
            ctx->d_nurbs_weights, d_items, t_sid, n_items, V, d_fac, ctx->n_row_nodes, f, xc);
      else if (surface)
      {
        const unsigned gs = (unsigned)((n_items + bs - 1) / bs);
#define FC_NEUMANN_SURFACE(FT)                                                                                                     \
  if (ortho)                                                                                                                       \
    fc::k_neumann_surface<FT, true><<<gs, bs, 0, ctx->stream>>>(ctx->d_X, d_items, n_items, V, d_fac, ctx->n_row_nodes, f, xc, d_u); \
  else                                                                                                                             \
    fc::k_neumann_surface<FT><<<gs, bs, 0, ctx->stream>>>(ctx->d_X, d_items, n_items, V, d_fac, ctx->n_row_nodes, f, xc)
        switch (face_type)
        {
          case FC_FACE_QUAD4: FC_NEUMANN_SURFACE(fc::QUAD4); break;
          case FC_FACE_QUAD9: FC_NEUMANN_SURFACE(fc::QUAD9); break;
          case FC_FACE_TRI6: FC_NEUMANN_SURFACE(fc::TRI6); break;
          case FC_FACE_QUAD8: FC_NEUMANN_SURFACE(fc::QUAD8); break;
          case FC_FACE_QUAD6: FC_NEUMANN_SURFACE(fc::QUAD6); break;
          default: FC_NEUMANN_SURFACE(fc::TRI3); break;
        }
#undef FC_NEUMANN_SURFACE
      }
      else
      {
        const long long nt = (long long)n_items * ngp;
        const unsigned gs = (unsigned)((nt + bs - 1) / bs);
        if (ctx->celltype == FC_HEX8) fc::k_neumann_volume<fc::HEX8><<<gs, bs, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, d_items, n_items, V, d_fac, ctx->n_row_nodes, f, xc);
        if (ctx->celltype == FC_HEX27) fc::k_neumann_volume<fc::HEX27><<<gs, bs, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, d_items, n_items, V, d_fac, ctx->n_row_nodes, f, xc);
        if (ctx->celltype == FC_TET10) fc::k_neumann_volume<fc::TET10><<<gs, bs, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, d_items, n_items, V, d_fac, ctx->n_row_nodes, f, xc);
        if (ctx->celltype == FC_HEX20) fc::k_neumann_volume<fc::HEX20><<<gs, bs, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, d_items, n_items, V, d_fac, ctx->n_row_nodes, f, xc);
        if (ctx->celltype == FC_TET4) fc::k_neumann_volume<fc::TET4><<<gs, bs, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, d_items, n_items, V, d_fac, ctx->n_row_nodes, f, xc);
        if (ctx->celltype == FC_WEDGE6) fc::k_neumann_volume<fc::WEDGE6><<<gs, bs, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, d_items, n_items, V, d_fac, ctx->n_row_nodes, f, xc);
        if (ctx->celltype == FC_HEX18) fc::k_neumann_volume<fc::HEX18><<<gs, bs, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, d_items, n_items, V, d_fac, ctx->n_row_nodes, f, xc);
        if (ctx->celltype == FC_NURBS27)
          fc::k_neumann_volume_nurbs<<<gs, bs, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, ctx->d_dNtab, ctx->d_gpw, d_items, n_items, V, d_fac, ctx->n_row_nodes, f, xc);
        if (ctx->celltype == FC_PYRAMID5) fc::k_neumann_volume<fc::PYRAMID5><<<gs, bs, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, d_items, n_items, V, d_fac, ctx->n_row_nodes, f, xc);
      }
      ctx->launches++;
      e = cudaGetLastError();
      if (e == cudaSuccess && !out_dev) e = cudaMemcpyAsync(out, d_out, n_out * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
    }
    const cudaError_t es = cudaStreamSynchronize(ctx->stream);
    if (e == cudaSuccess) e = es;
    cudaFree(t_items);
    if (t_sid) cudaFree(t_sid);
    if (t_u) cudaFree(t_u);
    if (t_fac) cudaFree(t_fac);
    if (t_out) cudaFree(t_out);
    if (rc != FC_OK) return rc;
    FC_CUDA(e);
    return FC_OK;
  }
}  // namespace

extern "C"
{
  int fc_solid_neumann_surface(fc_ctx* ctx, int face_type, int n_faces, const int* face_nodes, const int* onoff, const double* val,
      const double* gp_factor, double* f_ext)
  {
    return neumann_run(ctx, true, face_type, n_faces, face_nodes, onoff, val, gp_factor, f_ext, false);
  }

  int fc_solid_neumann_volume(fc_ctx* ctx, int n_ele, const int* ele_lids, const int* onoff, const double* val, const double* gp_factor, double* f_ext)
  {
    return neumann_run(ctx, false, 0, n_ele, ele_lids, onoff, val, gp_factor, f_ext, false);
  }

  int fc_solid_neumann_surface_gp_coords(fc_ctx* ctx, int face_type, int n_faces, const int* face_nodes, double* gp_coords)
  {
    return neumann_run(ctx, true, face_type, n_faces, face_nodes, nullptr, nullptr, nullptr, gp_coords, true);
  }

  int fc_solid_neumann_volume_gp_coords(fc_ctx* ctx, int n_ele, const int* ele_lids, double* gp_coords)
  {
    return neumann_run(ctx, false, 0, n_ele, ele_lids, nullptr, nullptr, nullptr, gp_coords, true);
  }

  int fc_solid_neumann_surface_orthopressure(fc_ctx* ctx, int face_type, int n_faces, const int* face_nodes, const double* u_col,
      double pressure, const double* gp_factor, double* f_ext)
  {
    const int onoff[3] = {1, 0, 0};
    const double val[3] = {pressure, 0.0, 0.0};
    return neumann_run(ctx, true, face_type, n_faces, face_nodes, onoff, val, gp_factor, f_ext, false, nullptr, true, u_col);
  }

  int fc_solid_neumann_surface_orthopressure_linearization(fc_ctx* ctx, int face_type, int n_faces, const int* face_nodes, const double* u_col,
      double pressure, const double* gp_factor, double* K_values)
  {
    if (!ctx || ctx->celltype < 0 || !ctx->have_graph) return fail(ctx, FC_ERR_INVALID, "mesh and graph are required");
    if (face_type < FC_FACE_QUAD4 || face_type > FC_FACE_QUAD6) return fail(ctx, FC_ERR_UNSUPPORTED, "surface cell type");
    if (n_faces < 0 || (n_faces > 0 && !face_nodes) || !K_values) return fail(ctx, FC_ERR_INVALID, "faces and K_values are required");
    if (n_faces == 0) return FC_OK;
    FC_CUDA(cudaSetDevice(ctx->device));
    const int per = fc::face_nen(face_type), ngp = fc::face_ngp(face_type);
    for (size_t i = 0; i < (size_t)n_faces * per; ++i)
      if (face_nodes[i] < 0 || face_nodes[i] >= ctx->n_col_nodes) return fail(ctx, FC_ERR_INVALID, "face node out of range");
    int* d_faces = nullptr;
    FC_CUDA(dev_upload(d_faces, face_nodes, (size_t)n_faces * per, ctx->stream));
    const double *d_u = nullptr, *d_fac = nullptr;
    double *t_u = nullptr, *t_fac = nullptr;
    int rc = u_col ? stage_in(ctx, u_col, (size_t)ctx->n_col_dof, &d_u, &t_u) : FC_OK;
    if (rc == FC_OK) rc = stage_in(ctx, gp_factor, (size_t)n_faces * ngp * 3, &d_fac, &t_fac);
    const bool K_dev = is_device_ptr(K_values);
    double* d_K = K_values;
    cudaError_t e = cudaSuccess;
    if (rc == FC_OK && !K_dev)
    {
      e = ensure_cap(ctx->d_stage_K, ctx->stage_K_cap, (size_t)ctx->nnz);
      d_K = ctx->d_stage_K;
      if (e == cudaSuccess) e = cudaMemcpyAsync(d_K, K_values, sizeof(double) * (size_t)ctx->nnz, cudaMemcpyHostToDevice, ctx->stream);  // +=
    }
    int missing = 0;
    if (rc == FC_OK && e == cudaSuccess)
    {
      ctx->h_err_pinned[0] = 0;
      e = cudaMemcpyAsync(ctx->d_err, ctx->h_err_pinned, sizeof(int), cudaMemcpyHostToDevice, ctx->stream);
      const int bs = 64;
      const unsigned gs = (unsigned)((n_faces + bs - 1) / bs);
      const unsigned char* dbc = ctx->have_dbc ? ctx->d_dbc : nullptr;
#define FC_ORTHO_LIN(FT)                                                                                                                 \
  fc::k_orthopressure_linearization<FT><<<gs, bs, 0, ctx->stream>>>(ctx->d_X, d_u, d_faces, n_faces, pressure, d_fac, ctx->n_row_nodes, \
      ctx->d_nodeptr, ctx->d_nodeadj, ctx->d_node_rowbase, ctx->d_node_rowlen, dbc, d_K, ctx->d_err)
      switch (face_type)
      {
        case FC_FACE_QUAD4: FC_ORTHO_LIN(fc::QUAD4); break;
        case FC_FACE_QUAD9: FC_ORTHO_LIN(fc::QUAD9); break;
        case FC_FACE_TRI6: FC_ORTHO_LIN(fc::TRI6); break;
        case FC_FACE_QUAD8: FC_ORTHO_LIN(fc::QUAD8); break;
        case FC_FACE_QUAD6: FC_ORTHO_LIN(fc::QUAD6); break;
        default: FC_ORTHO_LIN(fc::TRI3); break;
      }
#undef FC_ORTHO_LIN
      ctx->launches++;
      if (e == cudaSuccess) e = cudaGetLastError();
      if (e == cudaSuccess) e = cudaMemcpyAsync(ctx->h_err_pinned, ctx->d_err, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream);
      if (e == cudaSuccess && !K_dev) e = cudaMemcpyAsync(K_values, d_K, sizeof(double) * (size_t)ctx->nnz, cudaMemcpyDeviceToHost, ctx->stream);
    }
    const cudaError_t es = cudaStreamSynchronize(ctx->stream);
    if (e == cudaSuccess) e = es;
    missing = ctx->h_err_pinned[0];
    cudaFree(d_faces);
    if (t_u) cudaFree(t_u);
    if (t_fac) cudaFree(t_fac);
    if (rc != FC_OK) return rc;
    FC_CUDA(e);
    if (missing > 0) return fail(ctx, FC_ERR_INVALID, "the graph does not couple all nodes of a loaded face");
    return FC_OK;
  }

  int fc_solid_neumann_surface_nurbs(fc_ctx* ctx, int n_faces, const int* parent_ele, const int* surface_id, const int* onoff,
      const double* val, const double* gp_factor, double* f_ext)
  {
    if (n_faces > 0 && !surface_id) return fail(ctx, FC_ERR_INVALID, "surface_id is required");
    static const int none = 0;
    return neumann_run(ctx, true, -1, n_faces, parent_ele, onoff, val, gp_factor, f_ext, false, surface_id ? surface_id : &none);
  }

  int fc_solid_neumann_surface_nurbs_gp_coords(fc_ctx* ctx, int n_faces, const int* parent_ele, const int* surface_id, double* gp_coords)
  {
    if (n_faces > 0 && !surface_id) return fail(ctx, FC_ERR_INVALID, "surface_id is required");
    static const int none = 0;
    return neumann_run(ctx, true, -1, n_faces, parent_ele, nullptr, nullptr, nullptr, gp_coords, true, surface_id ? surface_id : &none);
  }
}
