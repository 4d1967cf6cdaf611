// Geometric multigrid preconditioner of the device linear solve.
//
// The reference solves the Newton system with Belos + MueLu or a direct solver (src/core/linear_solver); node-block
// Jacobi alone needs ~1000 CG iterations at 10^6 elements.  On a lattice mesh the hierarchy is geometric: every level
// keeps the nodes with even lattice indices, its tangent is the RE-DISCRETISATION of the same elements and material on
// the coarser mesh at the injected displacement state (evaluated by the library's own element kernels, Dirichlet rows
// injected from the fine toggle), the transfers are trilinear.  V(nu, nu) cycle with damped node-block Jacobi.
// One rank, hex8, isotropic materials; other meshes keep FC_SOLVER_CG / FC_SOLVER_BICGSTAB.
#include <algorithm>
#include <cmath>

#include "fc_brick_plan.hpp"
#include "fc_host_util.hpp"
#include "fc_kernels_linsolve.cuh"
#include "fc_kernels_mg.cuh"

void fc_mg_free(fc_ctx* ctx)
{
  for (size_t l = 0; l < ctx->mg.size(); ++l)
  {
    fc_mg_level& L = ctx->mg[l];
    dev_free(L.d_lat2lid);
    dev_free(L.d_lat);
    dev_free(L.d_free);
    dev_free(L.d_K);
    dev_free(L.d_Dinv);
    dev_free(L.d_b);
    dev_free(L.d_x);
    dev_free(L.d_y);
    dev_free(L.d_ev);
    dev_free(L.d_K32);
    dev_free(L.d_Ainv);
    L.have_Ainv = false;
    L.have_ev = false;
    dev_free(L.d_r);
    dev_free(L.d_u);
    if (L.c) fc_solid_destroy(L.c);
  }
  ctx->mg.clear();
  ctx->mg_toggle_version = -1;
  ctx->mg_operators_ready = false;
}

namespace
{
  constexpr int mg_nu_coarsest = 60;          // Jacobi sweeps on a coarsest level too large for the dense inverse
  constexpr int mg_dense_max_dof = 600;       // coarsest level: direct solve up to this many DOFs (2.9 MB of inverse)
  constexpr int mg_power_iterations = 8;       // first estimate of a level, from a random-sign start
  constexpr int mg_power_iterations_warm = 2;  // later estimates start from the previous iterate
  inline int fine_of(int I, int ne) { return std::min(2 * I, ne); }  // fine index of coarse node I on an axis with ne fine elements

  fc::MgLattice lattice_of(const fc_mg_level& L) { return fc::MgLattice{L.nn[0], L.nn[1], L.nn[2], L.d_lat2lid, L.d_lat}; }
  inline unsigned grid_for(long long n, int bs) { return (unsigned)((n + bs - 1) / bs); }

  int alloc_vectors(fc_ctx* ctx, fc_mg_level& L, bool coarse)
  {
    const size_t n3 = 3 * (size_t)L.n_nodes;
    FC_CUDA(cudaMalloc((void**)&L.d_x, n3 * sizeof(double)));
    FC_CUDA(cudaMalloc((void**)&L.d_y, n3 * sizeof(double)));
    FC_CUDA(cudaMalloc((void**)&L.d_ev, n3 * sizeof(double)));
    FC_CUDA(cudaMalloc((void**)&L.d_r, n3 * sizeof(double)));
    FC_CUDA(cudaMalloc((void**)&L.d_free, n3));
    if (coarse)
    {
      FC_CUDA(cudaMalloc((void**)&L.d_b, n3 * sizeof(double)));
      FC_CUDA(cudaMalloc((void**)&L.d_u, n3 * sizeof(double)));
      FC_CUDA(cudaMalloc((void**)&L.d_Dinv, 3 * n3 * sizeof(double)));
    }
    return FC_OK;
  }

  // lattice tables of a level on the device
  int upload_lattice(fc_ctx* ctx, fc_mg_level& L, const std::vector<int>& lat)
  {
    FC_CUDA(dev_upload(L.d_lat2lid, L.h_lat2lid.data(), L.h_lat2lid.size(), ctx->stream));
    FC_CUDA(dev_upload(L.d_lat, lat.data(), lat.size(), ctx->stream));
    return FC_OK;
  }

  // the tangent of a level as the cycle reads it: FP64 values or their single-precision copy
  struct LevelMatrix
  {
    const double* K64;
    const float* K32;  // used when not null
  };

  void smooth(fc_ctx* ctx, const fc_ctx* g, LevelMatrix M, const double* Dinv, double omega, const double* b, const double* x_in, double* x_out)
  {
    const int bs = 128;
    const unsigned gs = grid_for(32LL * g->n_row_nodes, bs);
    if (M.K32)
      fc::k_mg_smooth<float><<<gs, bs, 0, ctx->stream>>>(g->d_nodeptr, g->d_nodeadj, g->n_row_nodes, M.K32, Dinv, omega, b, x_in, x_out);
    else
      fc::k_mg_smooth<double><<<gs, bs, 0, ctx->stream>>>(g->d_nodeptr, g->d_nodeadj, g->n_row_nodes, M.K64, Dinv, omega, b, x_in, x_out);
    ctx->launches++;
  }

  // (re)fill the single-precision copy of a level's tangent; returns what the cycle reads
  int level_matrix(fc_ctx* ctx, fc_mg_level& L, const double* K, long long nnz, bool refresh, LevelMatrix* M)
  {
    M->K64 = K;
    M->K32 = nullptr;
    if (!ctx->mg_fp32) return FC_OK;
    if (!L.d_K32)
    {
      FC_CUDA(cudaMalloc((void**)&L.d_K32, sizeof(float) * (size_t)nnz));
      refresh = true;
    }
    if (refresh)
    {
      fc::k_mg_to_float<<<grid_for(nnz, 256), 256, 0, ctx->stream>>>(nnz, K, L.d_K32);
      ctx->launches++;
    }
    M->K32 = L.d_K32;
    return FC_OK;
  }

  // |v|_2 of a device vector (deterministic block partials, one read-back)
  int norm2(fc_ctx* ctx, int n, const double* v, double* out)
  {
    if (!ctx->d_dot) FC_CUDA(cudaMalloc((void**)&ctx->d_dot, sizeof(double) * (3 * fc::dot_blocks + 3 + 8)));
    fc::k_dot3_partial<<<fc::dot_blocks, fc::dot_threads, 0, ctx->stream>>>(n, v, v, nullptr, nullptr, nullptr, nullptr, ctx->d_dot);
    fc::k_dot3_final<<<3, 256, 0, ctx->stream>>>(ctx->d_dot, fc::dot_blocks, ctx->d_dot + 3 * fc::dot_blocks);
    ctx->launches += 2;
    double s = 0.0;
    FC_CUDA(cudaMemcpyAsync(&s, ctx->d_dot + 3 * fc::dot_blocks, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    *out = std::sqrt(s);
    return FC_OK;
  }

  // Jacobi damping of a level from lambda_max(D^-1 K), a few power iterations on I + D^-1 K (its largest eigenvalue is
  // 1 + lambda_max; the eigenvalues of D^-1 K are real and positive).  Stretched elements push lambda_max past the
  // 2 / omega of a fixed damping, and the smoother then diverges.
  int estimate_omega(fc_ctx* ctx, fc_mg_level& L, const fc_ctx* g, LevelMatrix K, const double* Dinv)
  {
    const int n3 = 3 * L.n_nodes, bs = 256;
    FC_CUDA(cudaMemsetAsync(L.d_r, 0, sizeof(double) * (size_t)n3, ctx->stream));
    // start: the previous estimate's iterate (the tangent of the next Newton iteration has nearly the same dominant
    // eigenvector: two iterations then do what eight do from a random start), zero on the fixed DOFs either way
    const int iterations = L.have_ev ? mg_power_iterations_warm : mg_power_iterations;
    if (L.have_ev)
      FC_CUDA(cudaMemcpyAsync(L.d_x, L.d_ev, sizeof(double) * (size_t)n3, cudaMemcpyDeviceToDevice, ctx->stream));
    else
    {
      fc::k_mg_seed<<<grid_for(n3, bs), bs, 0, ctx->stream>>>(n3, L.d_free, L.d_x);
      ctx->launches++;
    }
    // no normalisation in between: the iterate grows by at most (1 + lambda_max)^iterations, far from overflow,
    // and the host then waits for two norms per level instead of one per iteration
    double *cur = L.d_x, *oth = L.d_y;
    double n_prev = 0.0, n_last = 0.0, lambda = 2.0;
    for (int it = 0; it < iterations; ++it)
    {
      if (it == iterations - 1)
        if (int rc = norm2(ctx, n3, cur, &n_prev)) return rc;
      smooth(ctx, g, K, Dinv, -1.0, L.d_r, cur, oth);  // oth = (I + D^-1 K) cur
      std::swap(cur, oth);
    }
    if (int rc = norm2(ctx, n3, cur, &n_last)) return rc;
    if (n_prev > 0.0) lambda = n_last / n_prev - 1.0;
    if (!(lambda > 0.5) || !std::isfinite(lambda) || !(n_last > 0.0))
    {
      lambda = 2.0;
      L.have_ev = false;
    }
    else
    {
      fc::k_mg_scale_copy<<<grid_for(n3, bs), bs, 0, ctx->stream>>>(n3, 1.0 / n_last, cur, L.d_ev);
      ctx->launches++;
      L.have_ev = true;
    }
    // default 4/3.  The power iteration approaches lambda_max from below (measured at C2: the estimate is at least 15 % low), so
    // omega can exceed 2 / lambda_max on a level; CG-MG then sees r.z <= 0, cuts the damping and restarts (fc_capi_linsolve.cu)
    L.omega = ctx->mg_damping / lambda;
    return FC_OK;
  }
}  // namespace

extern "C"
{
  int fc_solid_mg_setup(fc_ctx* ctx, int max_levels, int* n_levels)
  {
    if (!ctx || !ctx->have_graph) return fail(ctx, FC_ERR_INVALID, "graph required");
    if (ctx->celltype != FC_HEX8 || ctx->nranks != 1 || ctx->n_col_nodes != ctx->n_row_nodes)
      return fail(ctx, FC_ERR_UNSUPPORTED, "the geometric multigrid hierarchy needs a hex8 mesh on one rank");
    if (ctx->d_fiber) return fail(ctx, FC_ERR_UNSUPPORTED, "element fibres are not carried to the coarse levels");
    if (ctx->formulation >= FC_FORM_MULF) return fail(ctx, FC_ERR_UNSUPPORTED, "element histories (MULF, EAS) are not carried to the coarse levels");
    FC_CUDA(cudaSetDevice(ctx->device));
    fc_mg_free(ctx);
    if (max_levels < 2) max_levels = 2;
    const int nn = ctx->n_col_nodes;
    std::vector<double> X((size_t)3 * nn);
    FC_CUDA(cudaMemcpyAsync(X.data(), ctx->d_X, sizeof(double) * X.size(), cudaMemcpyDeviceToHost, ctx->stream));
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    std::vector<int> Lc[3];
    int nl[3];
    for (int d = 0; d < 3; ++d)
      if (!fc_host::lattice_axis(nn, X.data(), d, ctx->n_ele, ctx->h_conn.data(), Lc[d], nl[d])) return fail(ctx, FC_ERR_GRAPH, "the mesh is not a lattice");
    if ((long long)nl[0] * nl[1] * nl[2] != nn) return fail(ctx, FC_ERR_GRAPH, "the mesh does not fill its lattice");
    if ((long long)(nl[0] - 1) * (nl[1] - 1) * (nl[2] - 1) != ctx->n_ele) return fail(ctx, FC_ERR_GRAPH, "the elements do not fill the lattice cells");
    ctx->mg.resize(1);
    {
      fc_mg_level& L = ctx->mg[0];
      L.n_nodes = nn;
      for (int d = 0; d < 3; ++d) L.nn[d] = nl[d];
      L.h_lat2lid.assign((size_t)nn, -1);
      std::vector<int> lat((size_t)3 * nn);
      for (int n = 0; n < nn; ++n)
      {
        int& slot = L.h_lat2lid[Lc[0][n] + (size_t)nl[0] * (Lc[1][n] + (size_t)nl[1] * Lc[2][n])];
        if (slot >= 0)
        {
          fc_mg_free(ctx);
          return fail(ctx, FC_ERR_GRAPH, "two nodes on one lattice point");
        }
        slot = n;
        for (int d = 0; d < 3; ++d) lat[3 * (size_t)n + d] = Lc[d][n];
      }
      if (int rc = upload_lattice(ctx, L, lat)) return rc;
      if (int rc = alloc_vectors(ctx, L, false)) return rc;
    }
    // coarsen while the coarse level keeps at least 2 elements per direction: every second node, plus the last node of a
    // direction with an odd number of elements ((ne + 1) / 2 coarse elements, the last one a single fine element wide)
    std::vector<double> Xf = X;
    while ((int)ctx->mg.size() < max_levels)
    {
      const fc_mg_level& F = ctx->mg.back();
      bool ok = true;
      for (int d = 0; d < 3; ++d) ok = ok && F.nn[d] - 1 >= 3;
      if (!ok) break;
      fc_mg_level C;
      for (int d = 0; d < 3; ++d) C.nn[d] = F.nn[d] / 2 + 1;  // (ne + 1) / 2 elements
      C.n_nodes = C.nn[0] * C.nn[1] * C.nn[2];
      const int cx = C.nn[0], cy = C.nn[1], cz = C.nn[2];
      std::vector<double> Xc((size_t)3 * C.n_nodes);
      std::vector<int> lat((size_t)3 * C.n_nodes), gid((size_t)C.n_nodes), owner((size_t)C.n_nodes, 0);
      C.h_lat2lid.resize((size_t)C.n_nodes);
      for (int k = 0; k < cz; ++k)
        for (int j = 0; j < cy; ++j)
          for (int i = 0; i < cx; ++i)
          {
            const int c = i + cx * (j + cy * k);
            const int f = F.h_lat2lid[fine_of(i, F.nn[0] - 1) + (size_t)F.nn[0] * (fine_of(j, F.nn[1] - 1) + (size_t)F.nn[1] * fine_of(k, F.nn[2] - 1))];
            for (int d = 0; d < 3; ++d) Xc[3 * (size_t)c + d] = Xf[3 * (size_t)f + d];
            lat[3 * (size_t)c] = i;
            lat[3 * (size_t)c + 1] = j;
            lat[3 * (size_t)c + 2] = k;
            gid[c] = c;
            C.h_lat2lid[c] = c;
          }
      const int ex = cx - 1, ey = cy - 1, ez = cz - 1;
      std::vector<int> conn((size_t)8 * ex * ey * ez), egid((size_t)ex * ey * ez);
      for (int k = 0; k < ez; ++k)
        for (int j = 0; j < ey; ++j)
          for (int i = 0; i < ex; ++i)
          {
            const int e = i + ex * (j + ey * k);
            auto nd = [&](int a, int b, int c) { return (i + a) + cx * ((j + b) + cy * (k + c)); };
            const int en[8] = {nd(0, 0, 0), nd(1, 0, 0), nd(1, 1, 0), nd(0, 1, 0), nd(0, 0, 1), nd(1, 0, 1), nd(1, 1, 1), nd(0, 1, 1)};
            std::copy(en, en + 8, conn.begin() + 8 * (size_t)e);
            egid[e] = e;
          }
      int rc = fc_solid_create(&C.c, ctx->device, 0, 1);
      if (rc == FC_OK) rc = fc_solid_set_stream(C.c, (void*)ctx->stream);
      if (rc == FC_OK) rc = fc_solid_set_mesh(C.c, FC_HEX8, ctx->formulation, C.n_nodes, Xc.data(), gid.data(), owner.data(), ex * ey * ez, conn.data(), egid.data());
      if (rc == FC_OK) rc = fc_solid_build_graph(C.c);
      if (rc != FC_OK)
      {
        const std::string msg = C.c ? C.c->last_error : "fc_solid_create failed";
        if (C.c) fc_solid_destroy(C.c);
        fc_mg_free(ctx);
        return fail(ctx, rc, "coarse level: " + msg);
      }
      FC_CUDA(cudaMalloc((void**)&C.d_K, sizeof(double) * (size_t)C.c->nnz));
      if (int r2 = upload_lattice(ctx, C, lat)) return r2;
      if (int r2 = alloc_vectors(ctx, C, true)) return r2;
      ctx->mg.push_back(std::move(C));
      Xf.swap(Xc);
    }
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    if (ctx->mg.size() < 2)
    {
      fc_mg_free(ctx);
      return fail(ctx, FC_ERR_UNSUPPORTED, "the lattice cannot be coarsened (fewer than 3 elements in a direction)");
    }
    if (n_levels) *n_levels = (int)ctx->mg.size();
    return FC_OK;
  }

  int fc_solid_mg_set_sweeps(fc_ctx* ctx, int nu)
  {
    if (!ctx || nu < 1 || nu > 8) return FC_ERR_INVALID;
    ctx->mg_nu = nu;
    return FC_OK;
  }

  int fc_solid_mg_set_damping(fc_ctx* ctx, double factor)
  {
    if (!ctx || !(factor > 0.0) || !(factor < 2.0)) return FC_ERR_INVALID;
    ctx->mg_damping = factor;
    ctx->mg_operators_ready = false;
    return FC_OK;
  }

  int fc_solid_mg_set_precision(fc_ctx* ctx, int single_precision_cycle)
  {
    if (!ctx) return FC_ERR_INVALID;
    ctx->mg_fp32 = single_precision_cycle != 0;
    ctx->mg_operators_ready = false;  // the copies are (re)made by fc_solid_mg_update
    return FC_OK;
  }

  int fc_solid_mg_level_info(fc_ctx* ctx, int level, int* n_nodes, double* omega)
  {
    if (!ctx || level < 0 || level >= (int)ctx->mg.size()) return FC_ERR_INVALID;
    if (n_nodes) *n_nodes = ctx->mg[level].n_nodes;
    if (omega) *omega = ctx->mg[level].omega;
    return FC_OK;
  }

  // coarse tangents at the injected state u_col_dev, coarse Dirichlet maps from the toggle, smoother blocks
  int fc_solid_mg_update(fc_ctx* ctx, const double* u_col_dev)
  {
    if (!ctx || ctx->mg.size() < 2) return fail(ctx, FC_ERR_INVALID, "fc_solid_mg_setup must precede fc_solid_mg_update");
    if (!u_col_dev || !is_device_ptr(u_col_dev)) return fail(ctx, FC_ERR_INVALID, "u_col_dev must be a device buffer");
    FC_CUDA(cudaSetDevice(ctx->device));
    const int bs = 256;
    if (ctx->mg_toggle_version != ctx->toggle_version)
    {
      // free-DOF masks: the fine one from the toggle, a coarse DOF is fixed where the coincident fine DOF is
      fc_mg_level& L0 = ctx->mg[0];
      L0.h_free.assign((size_t)3 * L0.n_nodes, 1);
      for (int r : ctx->h_toggle_rows) L0.h_free[r] = 0;
      for (size_t l = 1; l < ctx->mg.size(); ++l)
      {
        fc_mg_level &F = ctx->mg[l - 1], &C = ctx->mg[l];
        C.h_free.assign((size_t)3 * C.n_nodes, 1);
        std::vector<int> rows;
        for (int k = 0; k < C.nn[2]; ++k)
          for (int j = 0; j < C.nn[1]; ++j)
            for (int i = 0; i < C.nn[0]; ++i)
            {
              const int c = C.h_lat2lid[i + (size_t)C.nn[0] * (j + (size_t)C.nn[1] * k)];
              const int f = F.h_lat2lid[fine_of(i, F.nn[0] - 1) + (size_t)F.nn[0] * (fine_of(j, F.nn[1] - 1) + (size_t)F.nn[1] * fine_of(k, F.nn[2] - 1))];
              for (int d = 0; d < 3; ++d)
                if (!F.h_free[3 * (size_t)f + d])
                {
                  C.h_free[3 * (size_t)c + d] = 0;
                  rows.push_back(3 * c + d);
                }
            }
        if (int rc = fc_solid_set_dirichlet_toggle(C.c, (int)rows.size(), rows.data())) return fail(ctx, rc, "coarse level: " + C.c->last_error);
      }
      for (fc_mg_level& L : ctx->mg)
      {
        FC_CUDA(cudaMemcpyAsync(L.d_free, L.h_free.data(), L.h_free.size(), cudaMemcpyHostToDevice, ctx->stream));
        L.have_ev = false;  // the stored iterate is zero on the DOFs that were fixed before
      }
      FC_CUDA(cudaStreamSynchronize(ctx->stream));
      ctx->mg_toggle_version = ctx->toggle_version;
    }
    const double* u_f = u_col_dev;
    for (size_t l = 1; l < ctx->mg.size(); ++l)
    {
      fc_mg_level &F = ctx->mg[l - 1], &C = ctx->mg[l];
      C.c->mat = ctx->mat;
      C.c->have_material = ctx->have_material;
      fc::k_mg_inject<<<grid_for(C.n_nodes, bs), bs, 0, ctx->stream>>>(lattice_of(F), lattice_of(C), C.n_nodes, u_f, C.d_u);
      ctx->launches++;
      int cnt = 0, gid = -1;
      int rc = fc_solid_evaluate(C.c, FC_ACTION_NLNSTIFF, FC_EVAL_ZERO_FIRST, C.d_u, nullptr, C.d_K, &cnt, &gid);
      if (rc == FC_OK && cnt > 0)
      {
        // the injected state folds a coarse element: fall back to the tangent of the undeformed coarse mesh
        FC_CUDA(cudaMemsetAsync(C.d_u, 0, sizeof(double) * 3 * (size_t)C.n_nodes, ctx->stream));
        rc = fc_solid_evaluate(C.c, FC_ACTION_NLNSTIFF, FC_EVAL_ZERO_FIRST, C.d_u, nullptr, C.d_K, &cnt, &gid);
      }
      if (rc == FC_OK) rc = fc_solid_apply_dirichlet(C.c, C.d_K, 1);
      if (rc != FC_OK) return fail(ctx, rc, "coarse level: " + C.c->last_error);
      fc::k_block_jacobi_setup<<<grid_for(C.n_nodes, bs), bs, 0, ctx->stream>>>(C.c->d_nodeptr, C.c->d_node_diagpos, C.n_nodes, C.d_K, C.d_Dinv);
      ctx->launches++;
      LevelMatrix Mc;
      if (int r2 = level_matrix(ctx, C, C.d_K, C.c->nnz, true, &Mc)) return r2;
      if (int r2 = estimate_omega(ctx, C, C.c, Mc, C.d_Dinv)) return r2;
      u_f = C.d_u;
    }
    {
      // coarsest level: its inverse replaces the sweeps (one launch instead of sixty per cycle) when the level is small
      fc_mg_level& Cl = ctx->mg.back();
      const int n = 3 * Cl.n_nodes;
      Cl.have_Ainv = false;
      if (n <= mg_dense_max_dof)
      {
        if (!Cl.d_Ainv) FC_CUDA(cudaMalloc((void**)&Cl.d_Ainv, sizeof(double) * ((size_t)n * n + 1)));  // + the pivot flag
        int* d_ok = reinterpret_cast<int*>(Cl.d_Ainv + (size_t)n * n);
        fc::k_mg_dense_fill<<<1, 512, 0, ctx->stream>>>(Cl.c->d_nodeptr, Cl.c->d_nodeadj, Cl.n_nodes, Cl.d_K, Cl.d_Ainv);
        fc::k_mg_dense_invert<<<1, 1024, 0, ctx->stream>>>(n, Cl.d_Ainv, d_ok);
        ctx->launches += 2;
        int ok = 0;
        FC_CUDA(cudaMemcpyAsync(&ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        FC_CUDA(cudaStreamSynchronize(ctx->stream));
        Cl.have_Ainv = ok != 0;
      }
    }
    FC_CUDA(cudaGetLastError());
    ctx->mg_operators_ready = true;
    ctx->mg_fine_omega_stale = true;
    return FC_OK;
  }
}

// omega *= f on every level (the solver's answer to an indefinite cycle); holds until the next fc_solid_mg_update
void fc_mg_scale_damping(fc_ctx* ctx, double f)
{
  for (fc_mg_level& L : ctx->mg) L.omega *= f;
}

// z = M r: one V(nu, nu) cycle; K = the fine tangent (Dirichlet applied), ctx->d_Dinv = its inverse diagonal blocks
int fc_mg_vcycle(fc_ctx* ctx, const double* K, const double* r, double* z)
{
  const int nlev = (int)ctx->mg.size();
  const int bs = 256;
  // the fine tangent and its diagonal blocks are the caller's: known at the first cycle of a solve
  LevelMatrix M0;
  if (int rc = level_matrix(ctx, ctx->mg[0], K, ctx->nnz, ctx->mg_fine_omega_stale, &M0)) return rc;
  if (ctx->mg_fine_omega_stale)
  {
    if (int rc = estimate_omega(ctx, ctx->mg[0], ctx, M0, ctx->d_Dinv)) return rc;
    ctx->mg_fine_omega_stale = false;
  }
  // down
  for (int l = 0; l < nlev; ++l)
  {
    fc_mg_level& L = ctx->mg[l];
    const fc_ctx* g = l == 0 ? ctx : L.c;
    const LevelMatrix Kl = l == 0 ? M0 : LevelMatrix{L.d_K, ctx->mg_fp32 ? L.d_K32 : nullptr};
    const double* Di = l == 0 ? ctx->d_Dinv : L.d_Dinv;
    const double* b = l == 0 ? r : L.d_b;
    const int mg_nu = ctx->mg_nu;
    if (l == nlev - 1 && L.have_Ainv)
    {
      const int n = 3 * L.n_nodes;
      fc::k_mg_dense_apply<<<grid_for(32LL * n, 128), 128, 0, ctx->stream>>>(n, L.d_Ainv, b, L.d_x);
      ctx->launches++;
      break;
    }
    const int sweeps = l == nlev - 1 ? mg_nu_coarsest : mg_nu;
    double *cur = L.d_x, *oth = L.d_y;
    smooth(ctx, g, Kl, Di, L.omega, b, nullptr, cur);
    for (int s = 1; s < sweeps; ++s)
    {
      smooth(ctx, g, Kl, Di, L.omega, b, cur, oth);
      std::swap(cur, oth);
    }
    if (cur != L.d_x) std::swap(L.d_x, L.d_y);  // the iterate of the level lives in d_x
    if (l == nlev - 1) break;
    smooth(ctx, g, Kl, nullptr, 0.0, b, L.d_x, L.d_r);  // residual
    fc_mg_level& C = ctx->mg[l + 1];
    fc::k_mg_restrict<<<grid_for(C.n_nodes, bs), bs, 0, ctx->stream>>>(lattice_of(L), lattice_of(C), C.n_nodes, L.d_r, C.d_free, C.d_b);
    ctx->launches++;
  }
  // up
  for (int l = nlev - 2; l >= 0; --l)
  {
    fc_mg_level &L = ctx->mg[l], &C = ctx->mg[l + 1];
    const fc_ctx* g = l == 0 ? ctx : L.c;
    const LevelMatrix Kl = l == 0 ? M0 : LevelMatrix{L.d_K, ctx->mg_fp32 ? L.d_K32 : nullptr};
    const double* Di = l == 0 ? ctx->d_Dinv : L.d_Dinv;
    const double* b = l == 0 ? r : L.d_b;
    fc::k_mg_prolong_add<<<grid_for(L.n_nodes, bs), bs, 0, ctx->stream>>>(lattice_of(L), lattice_of(C), L.n_nodes, C.d_x, L.d_free, L.d_x);
    ctx->launches++;
    for (int s = 0; s < ctx->mg_nu; ++s)
    {
      smooth(ctx, g, Kl, Di, L.omega, b, L.d_x, L.d_y);
      std::swap(L.d_x, L.d_y);
    }
  }
  FC_CUDA(cudaMemcpyAsync(z, ctx->mg[0].d_x, sizeof(double) * 3 * (size_t)ctx->mg[0].n_nodes, cudaMemcpyDeviceToDevice, ctx->stream));
  FC_CUDA(cudaGetLastError());
  return FC_OK;
}
