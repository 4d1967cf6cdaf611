// The hot path: fc_solid_evaluate and friends.  Kernel selection:
//   hex8 + displacement-based + CoupNeoHooke -> closed-form kernel on FP64 tensor cores with per-element reductions
//                                               (fc_kernels_hex8nh_mma.cuh); with FC_EVAL_OWNER_COMPUTED on a lattice
//                                               mesh owner-computed rows with plain stores (fc_kernels_hex8nh_brick.cuh);
//                                               or the FP64-FMA sibling with FC_EVAL_CLOSED_FORM_FMA
//   everything else (and FC_EVAL_FORCE_GENERIC) -> generic B-operator kernel (fc_kernels_generic.cuh)
#include <climits>

#include "fc_host_util.hpp"
#include "fc_kernels_generic.cuh"
#include "fc_kernels_hex8nh.cuh"
#include "fc_kernels_hex8eas.cuh"
#include "fc_kernels_shell.cuh"
#include "fc_kernels_hex8nh_brick.cuh"
#include "fc_kernels_hex8nh_mma.cuh"
#include "fc_kernels_hex8_mma_ext.cuh"
#include "fc_kernels_nh_mma_gen.cuh"
#include "fc_kernels_setup.cuh"
#include "fc_kernels_tet10nh.cuh"

int fc_build_brick_plan(fc_ctx* ctx);  // fc_capi_graph.cu

namespace
{
  template <int CT, int FORM, bool WANT_K>
  cudaError_t launch_generic(fc_ctx* ctx, const fc::EvalArgs& A)
  {
    using Cfg = fc::GenCfg<CT, FORM>;
    auto kern = fc::k_generic<CT, FORM, WANT_K>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::smem_bytes);
    if (e != cudaSuccess) return e;
    const int grid = (A.n_ele + Cfg::epb - 1) / Cfg::epb;
    kern<<<grid, Cfg::threads, Cfg::smem_bytes, ctx->stream>>>(A, ctx->d_ele_bad);
    ctx->launches++;
    return cudaGetLastError();
  }

  template <int NEAS, bool WANT_K, bool LINEAR = false>
  cudaError_t launch_hex8_eas(fc_ctx* ctx, const fc::EvalArgs& A)
  {
    using Cfg = fc::EasCfg<NEAS>;
    auto kern = fc::k_hex8_eas<NEAS, WANT_K, LINEAR>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::smem_bytes);
    if (e != cudaSuccess) return e;
    kern<<<(A.n_ele + Cfg::epb - 1) / Cfg::epb, Cfg::threads, Cfg::smem_bytes, ctx->stream>>>(A, ctx->d_ele_bad);
    ctx->launches++;
    return cudaGetLastError();
  }

  template <int CT, int NEAS, bool WANT_K>
  cudaError_t launch_shell(fc_ctx* ctx, const fc::EvalArgs& A)
  {
    using Cfg = fc::ShellCfg<CT, NEAS>;
    auto kern = fc::k_shell<CT, NEAS, WANT_K>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::smem_bytes);
    if (e != cudaSuccess) return e;
    kern<<<(A.n_ele + Cfg::epb - 1) / Cfg::epb, Cfg::threads, Cfg::smem_bytes, ctx->stream>>>(A, ctx->d_ele_bad);
    ctx->launches++;
    return cudaGetLastError();
  }

  cudaError_t ensure_hex8_geom(fc_ctx* ctx)
  {
    if (ctx->d_geom) return cudaSuccess;  // geometry cache: once per mesh
    cudaError_t eg = cudaMalloc((void**)&ctx->d_geom, sizeof(double) * fc::Hex8MmaCfg::geom_doubles * (size_t)ctx->n_ele);
    if (eg != cudaSuccess) return eg;
    fc::k_hex8_geom_cache<<<ctx->n_ele, 64, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, ctx->n_ele, ctx->d_geom);
    ctx->launches++;
    return cudaGetLastError();
  }

  template <bool WANT_K>
  cudaError_t launch_hex8nh(fc_ctx* ctx, const fc::EvalArgs& A, bool fma_variant)
  {
    if (fma_variant)
    {
      using Cfg = fc::Hex8NHCfg;
      auto kern = fc::k_hex8_neohooke<WANT_K>;
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::smem_bytes);
      if (e != cudaSuccess) return e;
      kern<<<(A.n_ele + Cfg::epb - 1) / Cfg::epb, Cfg::threads, Cfg::smem_bytes, ctx->stream>>>(A, ctx->d_ele_bad);
    }
    else
    {
      using Cfg = fc::Hex8MmaCfg;
      cudaError_t e = ensure_hex8_geom(ctx);
      if (e != cudaSuccess) return e;
      fc::EvalArgs Ag = A;
      Ag.geom = ctx->d_geom;
      auto kern = Ag.ele_perm ? (A.dbc ? fc::k_hex8_neohooke_mma<WANT_K, true, true> : fc::k_hex8_neohooke_mma<WANT_K, false, true>)
                              : (A.dbc ? fc::k_hex8_neohooke_mma<WANT_K, true, false> : fc::k_hex8_neohooke_mma<WANT_K, false, false>);
      e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::smem_bytes);
      if (e != cudaSuccess) return e;
      kern<<<(A.n_ele - A.ele_offset + Cfg::epb - 1) / Cfg::epb, Cfg::threads, Cfg::smem_bytes, ctx->stream>>>(Ag, ctx->d_ele_bad);
    }
    ctx->launches++;
    return cudaGetLastError();
  }

  // owner-computed assembly of the units [u0, u1) of the brick plan; store = the rows are written (FC_EVAL_ZERO_FIRST),
  // otherwise added to what the caller's K and f hold
  cudaError_t launch_hex8_brick(fc_ctx* ctx, const fc::EvalArgs& A, int u0, int u1, bool store)
  {
    using Cfg = fc::Hex8BrickCfg;
    if (u1 <= u0) return cudaSuccess;
    cudaError_t e = ensure_hex8_geom(ctx);
    if (e != cudaSuccess) return e;
    fc::EvalArgs Ag = A;
    Ag.geom = ctx->d_geom;
    fc::BrickPlanDev P{ctx->d_brick_planes, ctx->d_brick_node_off, ctx->d_brick_ele_off, ctx->d_brick_nodes, ctx->d_brick_eles, ctx->d_brick_rows, u0};
    auto kern = store ? (A.dbc ? fc::k_hex8_nh_brick<true, true> : fc::k_hex8_nh_brick<true, false>)
                      : (A.dbc ? fc::k_hex8_nh_brick<false, true> : fc::k_hex8_nh_brick<false, false>);
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::smem_bytes);
    if (e != cudaSuccess) return e;
    kern<<<u1 - u0, Cfg::threads, Cfg::smem_bytes, ctx->stream>>>(Ag, P, ctx->d_ele_bad);
    ctx->launches++;
    return cudaGetLastError();
  }

  template <int VAR, bool WANT_K>
  cudaError_t launch_hex8_ext(fc_ctx* ctx, const fc::EvalArgs& A)
  {
    using Cfg = fc::Hex8MmaExtCfg;
    auto kern = A.ele_perm ? (A.dbc ? fc::k_hex8_mma_ext<VAR, WANT_K, true, true> : fc::k_hex8_mma_ext<VAR, WANT_K, false, true>)
                           : (A.dbc ? fc::k_hex8_mma_ext<VAR, WANT_K, true, false> : fc::k_hex8_mma_ext<VAR, WANT_K, false, false>);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::smem_bytes);
    if (e != cudaSuccess) return e;
    kern<<<(A.n_ele - A.ele_offset + Cfg::epb - 1) / Cfg::epb, Cfg::threads, Cfg::smem_bytes, ctx->stream>>>(A, ctx->d_ele_bad);
    ctx->launches++;
    return cudaGetLastError();
  }

  template <int CT, bool WANT_K>
  cudaError_t launch_nh_gen(fc_ctx* ctx, const fc::EvalArgs& A)
  {
    using Cfg = fc::NhGenCfg<CT>;
    fc::EvalArgs Ag = A;
    if (Cfg::geom)
    {
      if (!ctx->d_geom)  // geometry cache: once per mesh
      {
        cudaError_t eg = cudaMalloc((void**)&ctx->d_geom, sizeof(double) * Cfg::geom_doubles * (size_t)ctx->n_ele);
        if (eg != cudaSuccess) return eg;
        fc::k_gen_geom_cache<CT><<<ctx->n_ele, 128, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, ctx->n_ele, ctx->d_dNtab, ctx->d_gpw, ctx->d_geom);
        ctx->launches++;
      }
      Ag.geom = ctx->d_geom;
    }
    auto kern = A.dbc ? fc::k_nh_mma_gen<CT, WANT_K, true> : fc::k_nh_mma_gen<CT, WANT_K, false>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::smem_bytes);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    kern<<<(A.n_ele - A.ele_offset + Cfg::epb - 1) / Cfg::epb, Cfg::threads, Cfg::smem_bytes, ctx->stream>>>(Ag, ctx->d_ele_bad);
    ctx->launches++;
    return cudaGetLastError();
  }

  // tet10 Neo-Hooke, one warp per element
  template <bool WANT_K>
  cudaError_t launch_tet10_nh(fc_ctx* ctx, const fc::EvalArgs& A)
  {
    using Cfg = fc::Tet10NhCfg;
    if (!ctx->d_geom)  // geometry cache: once per mesh
    {
      cudaError_t eg = cudaMalloc((void**)&ctx->d_geom, sizeof(double) * Cfg::geom_doubles * (size_t)ctx->n_ele);
      if (eg != cudaSuccess) return eg;
      fc::k_tet10_geom_cache<<<(ctx->n_ele + 127) / 128, 128, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, ctx->n_ele, ctx->d_dNtab, ctx->d_gpw, ctx->d_geom);
      ctx->launches++;
    }
    fc::EvalArgs Ag = A;
    Ag.geom = ctx->d_geom;
    const bool ele_out = A.Ke_out || A.fe_out;
    auto kern = ele_out ? (A.dbc ? fc::k_tet10_nh<WANT_K, true, true> : fc::k_tet10_nh<WANT_K, false, true>)
                        : (A.dbc ? fc::k_tet10_nh<WANT_K, true, false> : fc::k_tet10_nh<WANT_K, false, false>);
    kern<<<(A.n_ele - A.ele_offset + Cfg::epb - 1) / Cfg::epb, Cfg::threads, 0, ctx->stream>>>(Ag, ctx->d_ele_bad);
    ctx->launches++;
    return cudaGetLastError();
  }

  // true if dispatch() picks a kernel that visits the elements through EvalArgs::ele_perm and the launch-slot range
  bool follows_launch_order(const fc_ctx* ctx, int flags)
  {
    if (flags & FC_EVAL_FORCE_GENERIC) return false;
    if (ctx->formulation >= FC_FORM_LINEAR) return false;  // linear, MULF, EAS: k_generic / k_hex8_eas
    const bool fbar = ctx->formulation == FC_FORM_FBAR, nh = ctx->mat.kind == fc::MAT_NEOHOOKE;
    if (ctx->celltype == FC_HEX8) return nh ? (fbar || !(flags & FC_EVAL_CLOSED_FORM_FMA)) : (ctx->mat.kind == fc::MAT_SVK && !fbar);  // k_hex8_neohooke_mma, k_hex8_mma_ext
    if (ctx->celltype != FC_HEX27 && ctx->celltype != FC_TET10 && ctx->celltype != FC_NURBS27) return false;  // k_generic
    return nh && !fbar;                                                                            // k_nh_mma_gen
  }

  template <bool WANT_K>
  cudaError_t dispatch(fc_ctx* ctx, const fc::EvalArgs& A, int flags)
  {
    if (ctx->formulation == FC_FORM_LINEAR && ctx->celltype == FC_HEX8 && ctx->mat.kind == fc::MAT_SVK && !(flags & FC_EVAL_FORCE_GENERIC))
      return launch_hex8_ext<fc::VAR_SVK_LINEAR, WANT_K>(ctx, A);  // tensor-core closed form
    if (ctx->formulation == FC_FORM_LINEAR)  // KINEM linear: the generic B-operator kernel for every other combination
      switch (ctx->celltype)
      {
        case FC_HEX8: return launch_generic<fc::HEX8, fc::FORM_LINEAR, WANT_K>(ctx, A);
        case FC_HEX27: return launch_generic<fc::HEX27, fc::FORM_LINEAR, WANT_K>(ctx, A);
        case FC_TET10: return launch_generic<fc::TET10, fc::FORM_LINEAR, WANT_K>(ctx, A);
        case FC_HEX20: return launch_generic<fc::HEX20, fc::FORM_LINEAR, WANT_K>(ctx, A);
        case FC_TET4: return launch_generic<fc::TET4, fc::FORM_LINEAR, WANT_K>(ctx, A);
        case FC_WEDGE6: return launch_generic<fc::WEDGE6, fc::FORM_LINEAR, WANT_K>(ctx, A);
        case FC_HEX18: return launch_generic<fc::HEX18, fc::FORM_LINEAR, WANT_K>(ctx, A);
        case FC_NURBS27: return launch_generic<fc::NURBS27, fc::FORM_LINEAR, WANT_K>(ctx, A);
        default: return launch_generic<fc::PYRAMID5, fc::FORM_LINEAR, WANT_K>(ctx, A);
      }
    if (ctx->formulation == FC_FORM_EAS_MILD) return launch_hex8_eas<9, WANT_K>(ctx, A);
    if (ctx->formulation == FC_FORM_EAS_FULL) return launch_hex8_eas<21, WANT_K>(ctx, A);
    if (ctx->formulation == FC_FORM_EAS_SHELL) return launch_hex8_eas<7, WANT_K>(ctx, A);
    if (ctx->formulation == FC_FORM_EAS_MILD_LINEAR) return launch_hex8_eas<9, WANT_K, true>(ctx, A);
    if (ctx->formulation == FC_FORM_EAS_FULL_LINEAR) return launch_hex8_eas<21, WANT_K, true>(ctx, A);
    if (ctx->formulation == FC_FORM_SHELL_ANS)
      return ctx->celltype == FC_HEX8 ? launch_shell<fc::HEX8, 0, WANT_K>(ctx, A) : launch_shell<fc::WEDGE6, 0, WANT_K>(ctx, A);
    if (ctx->formulation == FC_FORM_SHELL_EAS_ANS)
      return ctx->celltype == FC_HEX8 ? launch_shell<fc::HEX8, 7, WANT_K>(ctx, A) : launch_shell<fc::WEDGE6, 1, WANT_K>(ctx, A);
    if (ctx->formulation == FC_FORM_MULF_FBAR)
      return ctx->celltype == FC_HEX8 ? launch_generic<fc::HEX8, fc::FORM_MULF_FBAR, WANT_K>(ctx, A)
                                      : launch_generic<fc::PYRAMID5, fc::FORM_MULF_FBAR, WANT_K>(ctx, A);
    if (ctx->formulation == FC_FORM_MULF)  // PRESTRESS_TECH mulf: same kernel, F from the Gauss-point history
      switch (ctx->celltype)
      {
        case FC_HEX8: return launch_generic<fc::HEX8, fc::FORM_MULF, WANT_K>(ctx, A);
        case FC_HEX27: return launch_generic<fc::HEX27, fc::FORM_MULF, WANT_K>(ctx, A);
        case FC_TET10: return launch_generic<fc::TET10, fc::FORM_MULF, WANT_K>(ctx, A);
        case FC_HEX20: return launch_generic<fc::HEX20, fc::FORM_MULF, WANT_K>(ctx, A);
        case FC_TET4: return launch_generic<fc::TET4, fc::FORM_MULF, WANT_K>(ctx, A);
        case FC_WEDGE6: return launch_generic<fc::WEDGE6, fc::FORM_MULF, WANT_K>(ctx, A);
        case FC_HEX18: return launch_generic<fc::HEX18, fc::FORM_MULF, WANT_K>(ctx, A);
        case FC_NURBS27: return launch_generic<fc::NURBS27, fc::FORM_MULF, WANT_K>(ctx, A);
        default: return launch_generic<fc::PYRAMID5, fc::FORM_MULF, WANT_K>(ctx, A);
      }
    const bool fbar = ctx->formulation == FC_FORM_FBAR;
    if (ctx->celltype == FC_HEX8 && !fbar && ctx->mat.kind == fc::MAT_NEOHOOKE && !(flags & FC_EVAL_FORCE_GENERIC))
      return launch_hex8nh<WANT_K>(ctx, A, (flags & FC_EVAL_CLOSED_FORM_FMA) != 0);
    if (ctx->celltype == FC_HEX8 && fbar && ctx->mat.kind == fc::MAT_NEOHOOKE && !(flags & FC_EVAL_FORCE_GENERIC))
      return launch_hex8_ext<fc::VAR_NH_FBAR, WANT_K>(ctx, A);
    if (ctx->celltype == FC_HEX8 && !fbar && ctx->mat.kind == fc::MAT_SVK && !(flags & FC_EVAL_FORCE_GENERIC))
      return launch_hex8_ext<fc::VAR_SVK_DISP, WANT_K>(ctx, A);
    if (ctx->celltype == FC_HEX8) return fbar ? launch_generic<fc::HEX8, fc::FORM_FBAR, WANT_K>(ctx, A) : launch_generic<fc::HEX8, fc::FORM_DISP, WANT_K>(ctx, A);
    // SURVEY 8f F4 cell types: the generic B-operator kernel
    if (ctx->celltype == FC_HEX20) return launch_generic<fc::HEX20, fc::FORM_DISP, WANT_K>(ctx, A);
    if (ctx->celltype == FC_TET4) return launch_generic<fc::TET4, fc::FORM_DISP, WANT_K>(ctx, A);
    if (ctx->celltype == FC_WEDGE6) return launch_generic<fc::WEDGE6, fc::FORM_DISP, WANT_K>(ctx, A);
    if (ctx->celltype == FC_HEX18) return launch_generic<fc::HEX18, fc::FORM_DISP, WANT_K>(ctx, A);
    if (ctx->celltype == FC_NURBS27)  // Neo-Hooke: the tensor-core closed form of hex27 on the element's own shape tables
      return ctx->mat.kind == fc::MAT_NEOHOOKE && !(flags & FC_EVAL_FORCE_GENERIC) ? launch_nh_gen<fc::NURBS27, WANT_K>(ctx, A)
                                                                                  : launch_generic<fc::NURBS27, fc::FORM_DISP, WANT_K>(ctx, A);
    if (ctx->celltype == FC_PYRAMID5)
      return fbar ? launch_generic<fc::PYRAMID5, fc::FORM_FBAR, WANT_K>(ctx, A) : launch_generic<fc::PYRAMID5, fc::FORM_DISP, WANT_K>(ctx, A);
    const bool nh_gen = ctx->mat.kind == fc::MAT_NEOHOOKE && !fbar && !(flags & FC_EVAL_FORCE_GENERIC);
    if (ctx->celltype == FC_HEX27) return nh_gen ? launch_nh_gen<fc::HEX27, WANT_K>(ctx, A) : launch_generic<fc::HEX27, fc::FORM_DISP, WANT_K>(ctx, A);
    if (nh_gen && !(flags & FC_EVAL_DMMA_TILES)) return launch_tet10_nh<WANT_K>(ctx, A);
    return nh_gen ? launch_nh_gen<fc::TET10, WANT_K>(ctx, A) : launch_generic<fc::TET10, fc::FORM_DISP, WANT_K>(ctx, A);
  }

  // MULF history records per element: the Gauss points, with F-bar also the centroid; 0 without MULF
  int mulf_records(const fc_ctx* ctx)
  {
    if (ctx->formulation != FC_FORM_MULF && ctx->formulation != FC_FORM_MULF_FBAR) return 0;
    return fc::mulf_records(fc::cell_ngp(ctx->celltype), ctx->formulation == FC_FORM_MULF_FBAR ? fc::FORM_MULF_FBAR : fc::FORM_MULF);
  }

  // k_mulf_history over all column elements: u == nullptr writes the initial history
  cudaError_t launch_mulf_history(fc_ctx* ctx, const double* d_u)
  {
    const int nrec = mulf_records(ctx);
    const long long nt = (long long)ctx->n_ele * nrec;
    if (nt == 0) return cudaSuccess;
    const int bs = 128;
    const unsigned gs = (unsigned)((nt + bs - 1) / bs);
#define FC_MULF_HISTORY(CT) fc::k_mulf_history<CT><<<gs, bs, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, ctx->n_ele, ctx->d_dNtab, d_u, ctx->d_mulf, nrec)
    switch (ctx->celltype)
    {
      case FC_HEX8: FC_MULF_HISTORY(fc::HEX8); break;
      case FC_HEX27: FC_MULF_HISTORY(fc::HEX27); break;
      case FC_TET10: FC_MULF_HISTORY(fc::TET10); break;
      case FC_HEX20: FC_MULF_HISTORY(fc::HEX20); break;
      case FC_TET4: FC_MULF_HISTORY(fc::TET4); break;
      case FC_WEDGE6: FC_MULF_HISTORY(fc::WEDGE6); break;
      case FC_HEX18: FC_MULF_HISTORY(fc::HEX18); break;
      case FC_NURBS27: FC_MULF_HISTORY(fc::NURBS27); break;
      default: FC_MULF_HISTORY(fc::PYRAMID5); break;
    }
#undef FC_MULF_HISTORY
    ctx->launches++;
    return cudaGetLastError();
  }

  fc::EvalArgs base_args(const fc_ctx* ctx)
  {
    fc::EvalArgs A{};
    A.X = ctx->d_X;
    A.conn = ctx->d_conn;
    A.ele_gid = ctx->d_ele_gid;
    A.n_ele = ctx->n_ele;
    A.n_row_nodes = ctx->n_row_nodes;
    A.dNtab = ctx->d_dNtab;
    A.gpw = ctx->d_gpw;
    A.node_rowbase = ctx->d_node_rowbase;
    A.node_rowlen = ctx->d_node_rowlen;
    A.slots = ctx->d_slots;
    A.dbc = ctx->have_dbc ? ctx->d_dbc : nullptr;
    A.err = ctx->d_err;
    A.mat = ctx->mat;
    A.fiber = ctx->d_fiber;
    A.fiber_gp = ctx->fiber_gp;
    A.fiber_ele = ctx->fiber_gp ? ctx->fiber_gp * fc::cell_ngp(ctx->celltype) : 3;
    A.mulf = ctx->d_mulf;
    A.eas = ctx->d_eas;
    return A;
  }

  int reset_err(fc_ctx* ctx)
  {
    ctx->h_err_pinned[0] = 0;
    ctx->h_err_pinned[1] = INT_MAX;
    FC_CUDA(cudaMemcpyAsync(ctx->d_err, ctx->h_err_pinned, 2 * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    return FC_OK;
  }

  int read_err(fc_ctx* ctx, int* count, int* first_gid)
  {
    FC_CUDA(cudaMemcpyAsync(ctx->h_err_pinned, ctx->d_err, 2 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    if (count) *count = ctx->h_err_pinned[0];
    if (first_gid) *first_gid = ctx->h_err_pinned[0] > 0 ? ctx->h_err_pinned[1] : -1;
    return FC_OK;
  }
}  // namespace

static int eas_doubles(const fc_ctx* ctx)
{
  switch (ctx->formulation)
  {
    case FC_FORM_EAS_MILD:
    case FC_FORM_EAS_MILD_LINEAR: return fc::EasCfg<9>::hist_doubles;
    case FC_FORM_EAS_FULL:
    case FC_FORM_EAS_FULL_LINEAR: return fc::EasCfg<21>::hist_doubles;
    case FC_FORM_EAS_SHELL: return fc::EasCfg<7>::hist_doubles;
    case FC_FORM_SHELL_EAS_ANS: return ctx->celltype == FC_HEX8 ? fc::ShellCfg<fc::HEX8, 7>::hist_doubles : fc::ShellCfg<fc::WEDGE6, 1>::hist_doubles;
    default: return 0;
  }
}

// fc_solid_set_mesh: allocate and initialise the element history of an FC_FORM_MULF / FC_FORM_EAS_* mesh (the stream is
// synchronised by the caller)
int fc_mulf_init(fc_ctx* ctx)
{
  dev_free(ctx->d_mulf);
  dev_free(ctx->d_eas);
  if (eas_doubles(ctx) && ctx->n_ele > 0)
  {
    const long long n = (long long)ctx->n_ele * eas_doubles(ctx);
    FC_CUDA(cudaMalloc((void**)&ctx->d_eas, sizeof(double) * (size_t)n));
    fc::k_eas_history_init<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->n_ele, eas_doubles(ctx), ctx->d_eas);
    ctx->launches++;
    FC_CUDA(cudaGetLastError());
  }
  if (!mulf_records(ctx) || ctx->n_ele == 0) return FC_OK;
  if (ctx->celltype == FC_NURBS27 && !ctx->nurbs_ready) return FC_OK;  // the history needs the element tables: fc_solid_set_nurbs calls again
  FC_CUDA(cudaMalloc((void**)&ctx->d_mulf, sizeof(double) * 18 * (size_t)ctx->n_ele * mulf_records(ctx)));
  FC_CUDA(launch_mulf_history(ctx, nullptr));
  return FC_OK;
}

extern "C"
{
  int fc_solid_zero(fc_ctx* ctx, double* f_row_dev, double* K_values_dev)
  {
    if (!ctx || !ctx->have_graph) return fail(ctx, FC_ERR_INVALID, "graph required");
    FC_CUDA(cudaSetDevice(ctx->device));
    if (f_row_dev) FC_CUDA(cudaMemsetAsync(f_row_dev, 0, sizeof(double) * (size_t)ctx->n_row_dof, ctx->stream));
    if (K_values_dev) FC_CUDA(cudaMemsetAsync(K_values_dev, 0, sizeof(double) * (size_t)ctx->nnz, ctx->stream));
    ctx->launches += (f_row_dev ? 1 : 0) + (K_values_dev ? 1 : 0);
    return FC_OK;
  }

  int fc_solid_evaluate(fc_ctx* ctx, int action, int flags, const double* u_col, double* f_row,
      double* K_values, int* ele_error_count, int* first_bad_ele_gid)
  {
    if (!ctx || ctx->celltype < 0) return fail(ctx, FC_ERR_INVALID, "set_mesh must precede evaluate");
    if (!ctx->have_material) return fail(ctx, FC_ERR_INVALID, "set_material must precede evaluate");
    if (ctx->celltype == FC_NURBS27 && !ctx->nurbs_ready) return fail(ctx, FC_ERR_INVALID, "nurbs27: fc_solid_set_nurbs must follow set_mesh");
    if (!u_col) return fail(ctx, FC_ERR_INVALID, "u_col is required");
    if (action != FC_ACTION_NLNSTIFF && action != FC_ACTION_INTERNALFORCE) return fail(ctx, FC_ERR_UNSUPPORTED, "action not on the path");
    const bool want_K = action == FC_ACTION_NLNSTIFF;
    if (want_K && !K_values) return fail(ctx, FC_ERR_INVALID, "K_values is required for struct_calc_nlnstiff");
    if (want_K && !ctx->have_graph) return fail(ctx, FC_ERR_INVALID, "set_graph/build_graph must precede struct_calc_nlnstiff");
    FC_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    // host callers: stage through device buffers (copies are part of the call)
    const bool u_dev = is_device_ptr(u_col), f_dev = !f_row || is_device_ptr(f_row), K_dev = !want_K || is_device_ptr(K_values);
    const double* d_u = u_col;
    double* d_f = f_row;
    double* d_K = want_K ? K_values : nullptr;
    if (!u_dev)
    {
      FC_CUDA(ensure_cap(ctx->d_stage_u, ctx->stage_u_cap, (size_t)ctx->n_col_dof));
      FC_CUDA(cudaMemcpyAsync(ctx->d_stage_u, u_col, sizeof(double) * (size_t)ctx->n_col_dof, cudaMemcpyHostToDevice, s));
      d_u = ctx->d_stage_u;
    }
    if (!f_dev)
    {
      FC_CUDA(ensure_cap(ctx->d_stage_f, ctx->stage_f_cap, (size_t)ctx->n_row_dof));
      if (!(flags & FC_EVAL_ZERO_FIRST))
        FC_CUDA(cudaMemcpyAsync(ctx->d_stage_f, f_row, sizeof(double) * (size_t)ctx->n_row_dof, cudaMemcpyHostToDevice, s));
      d_f = ctx->d_stage_f;
    }
    if (!K_dev)
    {
      FC_CUDA(ensure_cap(ctx->d_stage_K, ctx->stage_K_cap, (size_t)ctx->nnz));
      if (!(flags & FC_EVAL_ZERO_FIRST))
        FC_CUDA(cudaMemcpyAsync(ctx->d_stage_K, K_values, sizeof(double) * (size_t)ctx->nnz, cudaMemcpyHostToDevice, s));
      d_K = ctx->d_stage_K;
    }

    // owner-computed assembly (brick kernel): hex8 displacement-based Neo-Hooke on a lattice mesh.  It writes every row of
    // K and f itself, so FC_EVAL_ZERO_FIRST costs no memset; without the flag it adds to the caller's values.
    bool brick = (flags & FC_EVAL_OWNER_COMPUTED) && want_K && ctx->celltype == FC_HEX8 && ctx->formulation == FC_FORM_DISPLACEMENT &&
                 ctx->mat.kind == fc::MAT_NEOHOOKE && !(flags & (FC_EVAL_FORCE_GENERIC | FC_EVAL_CLOSED_FORM_FMA));
    if (brick)
    {
      if (int rc = fc_build_brick_plan(ctx)) return rc;
      brick = ctx->n_brick_units > 0;
    }
    const bool only_interior = (flags & FC_EVAL_INTERIOR) && !(flags & FC_EVAL_BOUNDARY);
    const bool only_boundary = (flags & FC_EVAL_BOUNDARY) && !(flags & FC_EVAL_INTERIOR);
    // the boundary half completes an evaluation whose interior half stored its rows: it stores the remaining rows
    const bool store = (flags & FC_EVAL_ZERO_FIRST) || (brick && only_boundary && ctx->split_store);
    ctx->ev = ctx->ev_hist[ctx->ev_count++ % fc_ctx::ev_ring];
    FC_CUDA(cudaEventRecord(ctx->ev[0], s));
    if ((flags & FC_EVAL_ZERO_FIRST) && !brick)
    {
      if (d_f) FC_CUDA(cudaMemsetAsync(d_f, 0, sizeof(double) * (size_t)ctx->n_row_dof, s));
      if (d_K) FC_CUDA(cudaMemsetAsync(d_K, 0, sizeof(double) * (size_t)ctx->nnz, s));
      ctx->launches += (d_f ? 1 : 0) + (d_K ? 1 : 0);
    }
    FC_CUDA(cudaEventRecord(ctx->ev[1], s));
    if (int rc = reset_err(ctx)) return rc;
    fc::EvalArgs A = base_args(ctx);
    A.u = d_u;
    A.f = d_f;
    A.K = d_K;
    A.ele_perm = ctx->d_ele_perm;
    if (brick)
    {
      const int u0 = only_boundary ? ctx->n_brick_interior : 0, u1 = only_interior ? ctx->n_brick_interior : ctx->n_brick_units;
      FC_CUDA(launch_hex8_brick(ctx, A, u0, u1, store));
      ctx->split_store = only_interior && (flags & FC_EVAL_ZERO_FIRST);
    }
    else
    {
      if (flags & (FC_EVAL_INTERIOR | FC_EVAL_BOUNDARY))
      {
        // elements without / with ghost nodes: the interior needs no halo import and can run while it is in flight.
        // The split lives in the launch-slot order (ele_perm), which only the tensor-core kernels follow.
        if (!follows_launch_order(ctx, flags)) return fail(ctx, FC_ERR_UNSUPPORTED, "FC_EVAL_INTERIOR/BOUNDARY need one of the tensor-core kernels");
        if (only_interior) A.n_ele = ctx->n_interior_ele;
        if (only_boundary) A.ele_offset = ctx->n_interior_ele;
      }
      if (A.n_ele > A.ele_offset) FC_CUDA(want_K ? dispatch<true>(ctx, A, flags) : dispatch<false>(ctx, A, flags));
    }
    FC_CUDA(cudaEventRecord(ctx->ev[2], s));
    ctx->ev_valid = true;

    if (!f_dev) FC_CUDA(cudaMemcpyAsync(f_row, d_f, sizeof(double) * (size_t)ctx->n_row_dof, cudaMemcpyDeviceToHost, s));
    if (!K_dev) FC_CUDA(cudaMemcpyAsync(K_values, d_K, sizeof(double) * (size_t)ctx->nnz, cudaMemcpyDeviceToHost, s));
    const bool must_sync = !f_dev || !K_dev || ele_error_count || first_bad_ele_gid;
    if (must_sync)
    {
      int cnt = 0, gid = -1;
      if (int rc = read_err(ctx, &cnt, &gid)) return rc;
      if (ele_error_count) *ele_error_count = cnt;
      if (first_bad_ele_gid) *first_bad_ele_gid = gid;
    }
    // the node-Jacobian check of set_mesh is a host value: reported by every evaluate, synchronous or not
    if (ctx->geom_bad_count > 0) return fail(ctx, FC_ERR_NEGATIVE_JACOBIAN, "determinant of jacobian is <= 0 at one node of an element");
    return FC_OK;
  }

  int fc_solid_error_state(fc_ctx* ctx, int* ele_error_count, int* first_bad_ele_gid)
  {
    if (!ctx) return FC_ERR_INVALID;
    FC_CUDA(cudaSetDevice(ctx->device));
    if (int rc = read_err(ctx, ele_error_count, first_bad_ele_gid)) return rc;
    if (ctx->geom_bad_count > 0) return fail(ctx, FC_ERR_NEGATIVE_JACOBIAN, "determinant of jacobian is <= 0 at one node of an element");
    return FC_OK;
  }

  int fc_solid_evaluate_elements(fc_ctx* ctx, const double* u_col, double* fe, double* Ke)
  {
    if (!ctx || ctx->celltype < 0 || !ctx->have_material || !u_col || !fe) return fail(ctx, FC_ERR_INVALID, "mesh, material, u_col and fe are required");
    if (ctx->celltype == FC_NURBS27 && !ctx->nurbs_ready) return fail(ctx, FC_ERR_INVALID, "nurbs27: fc_solid_set_nurbs must follow set_mesh");
    FC_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    const size_t nd = 3 * (size_t)ctx->nen, ne = (size_t)ctx->n_ele;
    const bool u_dev = is_device_ptr(u_col), fe_dev = is_device_ptr(fe), Ke_dev = !Ke || is_device_ptr(Ke);
    double *d_u = nullptr, *d_fe = nullptr, *d_Ke = nullptr;
    if (!u_dev)
    {
      FC_CUDA(cudaMalloc((void**)&d_u, sizeof(double) * (size_t)ctx->n_col_dof));
      FC_CUDA(cudaMemcpyAsync(d_u, u_col, sizeof(double) * (size_t)ctx->n_col_dof, cudaMemcpyHostToDevice, s));
    }
    if (!fe_dev) FC_CUDA(cudaMalloc((void**)&d_fe, sizeof(double) * ne * nd));
    if (!Ke_dev) FC_CUDA(cudaMalloc((void**)&d_Ke, sizeof(double) * ne * nd * nd));
    if (int rc = reset_err(ctx)) return rc;
    fc::EvalArgs A = base_args(ctx);
    A.u = u_dev ? u_col : d_u;
    A.node_rowbase = nullptr;
    A.node_rowlen = nullptr;
    A.slots = nullptr;
    A.dbc = nullptr;
    A.fe_out = fe_dev ? fe : d_fe;
    A.Ke_out = Ke ? (Ke_dev ? Ke : d_Ke) : nullptr;
    // failing elements leave their slots untouched: define them as zero
    FC_CUDA(cudaMemsetAsync(A.fe_out, 0, sizeof(double) * ne * nd, s));
    if (A.Ke_out) FC_CUDA(cudaMemsetAsync(A.Ke_out, 0, sizeof(double) * ne * nd * nd, s));
    const int flags = ctx->eval_elements_flags;
    if (ne > 0) FC_CUDA(Ke ? dispatch<true>(ctx, A, flags) : dispatch<false>(ctx, A, flags));
    if (!fe_dev) FC_CUDA(cudaMemcpyAsync(fe, d_fe, sizeof(double) * ne * nd, cudaMemcpyDeviceToHost, s));
    if (!Ke_dev) FC_CUDA(cudaMemcpyAsync(Ke, d_Ke, sizeof(double) * ne * nd * nd, cudaMemcpyDeviceToHost, s));
    FC_CUDA(cudaStreamSynchronize(s));
    if (d_u) cudaFree(d_u);
    if (d_fe) cudaFree(d_fe);
    if (d_Ke) cudaFree(d_Ke);
    return FC_OK;
  }

  int fc_solid_set_element_kernel_flags(fc_ctx* ctx, int flags)
  {
    if (!ctx) return FC_ERR_INVALID;
    ctx->eval_elements_flags = flags;
    return FC_OK;
  }

  int fc_solid_halo_set_send(fc_ctx* ctx, int n_send, const int* send_row_lids)
  {
    if (!ctx || n_send < 0 || (n_send > 0 && !send_row_lids)) return FC_ERR_INVALID;
    FC_CUDA(cudaSetDevice(ctx->device));
    for (int i = 0; i < n_send; ++i)
      if (send_row_lids[i] < 0 || send_row_lids[i] >= 3 * ctx->n_row_nodes) return fail(ctx, FC_ERR_INVALID, "send DOF is not an owned row");
    FC_CUDA(dev_upload(ctx->d_send_lids, send_row_lids, (size_t)n_send, ctx->stream));
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->n_send = n_send;
    return FC_OK;
  }

  int fc_solid_halo_pack(fc_ctx* ctx, const double* u_dev, double* sendbuf_dev)
  {
    if (!ctx || !u_dev || (ctx->n_send > 0 && !sendbuf_dev)) return FC_ERR_INVALID;
    if (ctx->n_send == 0) return FC_OK;
    FC_CUDA(cudaSetDevice(ctx->device));
    const int bs = 256;
    fc::k_halo_pack<<<(ctx->n_send + bs - 1) / bs, bs, 0, ctx->stream>>>(u_dev, ctx->d_send_lids, ctx->n_send, sendbuf_dev);
    ctx->launches++;
    FC_CUDA(cudaGetLastError());
    return FC_OK;
  }

  int fc_solid_last_kernel_ms(fc_ctx* ctx, float* zero_ms, float* element_ms)
  {
    if (!ctx || !ctx->ev_valid) return FC_ERR_INVALID;
    FC_CUDA(cudaEventSynchronize(ctx->ev[2]));
    if (zero_ms) FC_CUDA(cudaEventElapsedTime(zero_ms, ctx->ev[0], ctx->ev[1]));
    if (element_ms) FC_CUDA(cudaEventElapsedTime(element_ms, ctx->ev[1], ctx->ev[2]));
    return FC_OK;
  }

  int fc_solid_kernel_ms_history(fc_ctx* ctx, int n, float* zero_ms, float* element_ms)
  {
    if (!ctx || !ctx->ev_valid || n < 1 || n > fc_ctx::ev_ring || n > ctx->ev_count) return FC_ERR_INVALID;
    FC_CUDA(cudaEventSynchronize(ctx->ev[2]));
    for (int k = 0; k < n; ++k)
    {
      cudaEvent_t* t = ctx->ev_hist[(ctx->ev_count - n + k) % fc_ctx::ev_ring];
      if (zero_ms) FC_CUDA(cudaEventElapsedTime(zero_ms + k, t[0], t[1]));
      if (element_ms) FC_CUDA(cudaEventElapsedTime(element_ms + k, t[1], t[2]));
    }
    return FC_OK;
  }

  int fc_solid_kernel_launch_count(const fc_ctx* ctx, int64_t* launches)
  {
    if (!ctx || !launches) return FC_ERR_INVALID;
    *launches = ctx->launches;
    return FC_OK;
  }

  int fc_solid_update_prestress(fc_ctx* ctx, const double* u_col)
  {
    if (!ctx || ctx->celltype < 0 || !u_col) return fail(ctx, FC_ERR_INVALID, "mesh and u_col are required");
    if (!mulf_records(ctx)) return fail(ctx, FC_ERR_INVALID, "the mesh was not set with FC_FORM_MULF / FC_FORM_MULF_FBAR");
    if (ctx->celltype == FC_NURBS27 && !ctx->nurbs_ready) return fail(ctx, FC_ERR_INVALID, "nurbs27: fc_solid_set_nurbs must follow set_mesh");
    if (ctx->n_ele == 0) return FC_OK;
    FC_CUDA(cudaSetDevice(ctx->device));
    double* t_u = nullptr;
    const double* d_u = u_col;
    if (!is_device_ptr(u_col))
    {
      FC_CUDA(cudaMalloc((void**)&t_u, sizeof(double) * (size_t)ctx->n_col_dof));
      FC_CUDA(cudaMemcpyAsync(t_u, u_col, sizeof(double) * (size_t)ctx->n_col_dof, cudaMemcpyHostToDevice, ctx->stream));
      d_u = t_u;
    }
    cudaError_t e = launch_mulf_history(ctx, d_u);
    const cudaError_t es = cudaStreamSynchronize(ctx->stream);
    if (t_u) cudaFree(t_u);
    FC_CUDA(e);
    FC_CUDA(es);
    return FC_OK;
  }

  int fc_solid_mulf_get_history(fc_ctx* ctx, double* history)
  {
    if (!ctx || ctx->celltype < 0 || !history) return fail(ctx, FC_ERR_INVALID, "mesh and history are required");
    if (!mulf_records(ctx)) return fail(ctx, FC_ERR_INVALID, "the mesh was not set with FC_FORM_MULF / FC_FORM_MULF_FBAR");
    if (ctx->n_ele == 0) return FC_OK;
    if (!ctx->d_mulf) return fail(ctx, FC_ERR_INVALID, "no prestress history yet (nurbs27: fc_solid_set_nurbs first)");
    FC_CUDA(cudaSetDevice(ctx->device));
    FC_CUDA(cudaMemcpyAsync(history, ctx->d_mulf, sizeof(double) * 18 * (size_t)ctx->n_ele * mulf_records(ctx), cudaMemcpyDefault, ctx->stream));
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    return FC_OK;
  }

  int fc_solid_mulf_set_history(fc_ctx* ctx, const double* history)
  {
    if (!ctx || ctx->celltype < 0 || !history) return fail(ctx, FC_ERR_INVALID, "mesh and history are required");
    if (!mulf_records(ctx)) return fail(ctx, FC_ERR_INVALID, "the mesh was not set with FC_FORM_MULF / FC_FORM_MULF_FBAR");
    if (ctx->n_ele == 0) return FC_OK;
    if (!ctx->d_mulf) return fail(ctx, FC_ERR_INVALID, "no prestress history yet (nurbs27: fc_solid_set_nurbs first)");
    FC_CUDA(cudaSetDevice(ctx->device));
    FC_CUDA(cudaMemcpyAsync(ctx->d_mulf, history, sizeof(double) * 18 * (size_t)ctx->n_ele * mulf_records(ctx), cudaMemcpyDefault, ctx->stream));
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    return FC_OK;
  }

  int fc_solid_recover(fc_ctx* ctx, const double* du_col, double step_length)
  {
    if (!ctx || ctx->celltype < 0 || !du_col) return fail(ctx, FC_ERR_INVALID, "mesh and du_col are required");
    if (!eas_doubles(ctx)) return FC_OK;  // no condensed variables: nothing to recover (SolidEleCalc::recover, calc.cpp:244-273)
    if (ctx->n_ele == 0) return FC_OK;
    FC_CUDA(cudaSetDevice(ctx->device));
    double* t_u = nullptr;
    const double* d_u = du_col;
    if (!is_device_ptr(du_col))
    {
      FC_CUDA(cudaMalloc((void**)&t_u, sizeof(double) * (size_t)ctx->n_col_dof));
      FC_CUDA(cudaMemcpyAsync(t_u, du_col, sizeof(double) * (size_t)ctx->n_col_dof, cudaMemcpyHostToDevice, ctx->stream));
      d_u = t_u;
    }
    const unsigned gs = (unsigned)((ctx->n_ele + 7) / 8);
    if (ctx->formulation == FC_FORM_SHELL_EAS_ANS && ctx->celltype == FC_HEX8)
      fc::k_shell_recover<fc::HEX8, 7><<<gs, 256, 0, ctx->stream>>>(ctx->d_conn, ctx->n_ele, d_u, step_length, ctx->d_eas);
    else if (ctx->formulation == FC_FORM_SHELL_EAS_ANS)
      fc::k_shell_recover<fc::WEDGE6, 1><<<gs, 256, 0, ctx->stream>>>(ctx->d_conn, ctx->n_ele, d_u, step_length, ctx->d_eas);
    else if (ctx->formulation == FC_FORM_EAS_SHELL)
      fc::k_hex8_eas_recover<7><<<gs, 256, 0, ctx->stream>>>(ctx->d_conn, ctx->n_ele, d_u, step_length, ctx->d_eas);
    else if (ctx->formulation == FC_FORM_EAS_MILD || ctx->formulation == FC_FORM_EAS_MILD_LINEAR)
      fc::k_hex8_eas_recover<9><<<gs, 256, 0, ctx->stream>>>(ctx->d_conn, ctx->n_ele, d_u, step_length, ctx->d_eas);
    else
      fc::k_hex8_eas_recover<21><<<gs, 256, 0, ctx->stream>>>(ctx->d_conn, ctx->n_ele, d_u, step_length, ctx->d_eas);
    ctx->launches++;
    const cudaError_t e = cudaGetLastError();
    const cudaError_t es = t_u ? cudaStreamSynchronize(ctx->stream) : cudaSuccess;  // device input: stays asynchronous
    if (t_u) cudaFree(t_u);
    FC_CUDA(e);
    FC_CUDA(es);
    return FC_OK;
  }

  int fc_solid_eas_history_doubles(const fc_ctx* ctx, int* doubles_per_element)
  {
    if (!ctx || !doubles_per_element) return FC_ERR_INVALID;
    *doubles_per_element = eas_doubles(ctx);
    return FC_OK;
  }

  int fc_solid_eas_get_history(fc_ctx* ctx, double* history)
  {
    if (!ctx || ctx->celltype < 0 || !history) return fail(ctx, FC_ERR_INVALID, "mesh and history are required");
    if (!eas_doubles(ctx)) return fail(ctx, FC_ERR_INVALID, "the mesh was not set with one of the FC_FORM_EAS_* formulations");
    if (ctx->n_ele == 0) return FC_OK;
    FC_CUDA(cudaSetDevice(ctx->device));
    FC_CUDA(cudaMemcpyAsync(history, ctx->d_eas, sizeof(double) * (size_t)ctx->n_ele * eas_doubles(ctx), cudaMemcpyDefault, ctx->stream));
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    return FC_OK;
  }

  int fc_solid_eas_set_history(fc_ctx* ctx, const double* history)
  {
    if (!ctx || ctx->celltype < 0 || !history) return fail(ctx, FC_ERR_INVALID, "mesh and history are required");
    if (!eas_doubles(ctx)) return fail(ctx, FC_ERR_INVALID, "the mesh was not set with one of the FC_FORM_EAS_* formulations");
    if (ctx->n_ele == 0) return FC_OK;
    FC_CUDA(cudaSetDevice(ctx->device));
    FC_CUDA(cudaMemcpyAsync(ctx->d_eas, history, sizeof(double) * (size_t)ctx->n_ele * eas_doubles(ctx), cudaMemcpyDefault, ctx->stream));
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    return FC_OK;
  }
}
