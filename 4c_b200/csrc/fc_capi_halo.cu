// fc_solid_halo_exchange: the forward import of Discretization::set_state (4C_fem_discretization.cpp:546-551)
// as pack kernel + grouped NCCL point-to-point on the context's stream.  NCCL is not linked: the five entry
// points are resolved from the process image (a 4C build with NCCL, or torch's bundled copy) or libnccl.so.2.
#include <dlfcn.h>

#include "fc_host_util.hpp"
#include "fc_kernels_setup.cuh"

namespace
{
  struct NcclApi
  {
    int (*send)(const void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*recv)(void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*group_start)() = nullptr;
    int (*group_end)() = nullptr;
    int (*all_reduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    const char* (*error_string)(int) = nullptr;
    bool ok = false;
  };

  NcclApi& nccl()
  {
    static NcclApi api = []
    {
      NcclApi a;
      void* h = RTLD_DEFAULT;  // note: RTLD_DEFAULT is a null handle
      if (!dlsym(h, "ncclSend"))
      {
        h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) return a;
      }
      a.send = (decltype(a.send))dlsym(h, "ncclSend");
      a.recv = (decltype(a.recv))dlsym(h, "ncclRecv");
      a.group_start = (decltype(a.group_start))dlsym(h, "ncclGroupStart");
      a.group_end = (decltype(a.group_end))dlsym(h, "ncclGroupEnd");
      a.all_reduce = (decltype(a.all_reduce))dlsym(h, "ncclAllReduce");
      a.error_string = (decltype(a.error_string))dlsym(h, "ncclGetErrorString");
      a.ok = a.send && a.recv && a.group_start && a.group_end;
      return a;
    }();
    return api;
  }
  constexpr int kNcclFloat64 = 8;  // ncclDataType_t::ncclFloat64 (nccl.h), stable across NCCL 2.x
  constexpr int kNcclInt32 = 2;    // ncclDataType_t::ncclInt32
  constexpr int kNcclSum = 0;      // ncclRedOp_t::ncclSum
  constexpr int kNcclMax = 2;      // ncclRedOp_t::ncclMax

  // err[2] = 1 if this rank saw an element failure in the last evaluation (err[0] > 0) or its mesh failed the node-Jacobian check
  __global__ void k_error_flag(int* err, int geom_bad) { err[2] = (err[0] > 0 || geom_bad > 0) ? 1 : 0; }
}  // namespace

// sum of n doubles over the ranks of the context's communicator, in place, on the context's stream (linear solve)
int fc_allreduce_sum(fc_ctx* ctx, double* buf_dev, int n)
{
  if (ctx->nranks == 1) return FC_OK;
  NcclApi& api = nccl();
  if (!api.ok || !api.all_reduce) return fail(ctx, FC_ERR_UNSUPPORTED, "NCCL entry points not found in the process and libnccl.so.2 cannot be loaded");
  if (!ctx->nccl_comm) return fail(ctx, FC_ERR_INVALID, "fc_solid_set_comm must precede a multi-rank linear solve");
  const int rc = api.all_reduce(buf_dev, buf_dev, (size_t)n, kNcclFloat64, kNcclSum, ctx->nccl_comm, ctx->stream);
  if (rc != 0) return fail(ctx, FC_ERR_CUDA, std::string("NCCL: ") + (api.error_string ? api.error_string(rc) : "error"));
  ctx->launches++;
  return FC_OK;
}

extern "C"
{
  int fc_solid_halo_set_counts(fc_ctx* ctx, const int* send_dof_count, const int* recv_dof_count)
  {
    if (!ctx || !send_dof_count || !recv_dof_count) return FC_ERR_INVALID;
    long long ns = 0, nr = 0;
    for (int p = 0; p < ctx->nranks; ++p)
    {
      if (send_dof_count[p] < 0 || recv_dof_count[p] < 0) return fail(ctx, FC_ERR_INVALID, "negative halo count");
      ns += send_dof_count[p];
      nr += recv_dof_count[p];
    }
    if (send_dof_count[ctx->rank] || recv_dof_count[ctx->rank]) return fail(ctx, FC_ERR_INVALID, "halo count for the own rank must be 0");
    if (ns != ctx->n_send) return fail(ctx, FC_ERR_INVALID, "send counts do not add up to the send list (fc_solid_halo_set_send)");
    if (ctx->n_col_nodes > 0 && nr != 3ll * (ctx->n_col_nodes - ctx->n_row_nodes))
      return fail(ctx, FC_ERR_INVALID, "receive counts do not add up to the ghost DOFs of the mesh");
    if (ctx->n_col_nodes > 0)
    {
      // rank p's message lands behind the messages of the ranks below p: it must be exactly p's ghost nodes
      std::vector<long long> ghosts((size_t)ctx->nranks, 0);
      for (int n = ctx->n_row_nodes; n < ctx->n_col_nodes; ++n) ghosts[ctx->h_node_owner[n]] += 3;
      for (int p = 0; p < ctx->nranks; ++p)
        if (ghosts[p] != recv_dof_count[p]) return fail(ctx, FC_ERR_INVALID, "receive count of a rank differs from 3 x the ghost nodes it owns");
    }
    ctx->halo_send_count.assign(send_dof_count, send_dof_count + ctx->nranks);
    ctx->halo_recv_count.assign(recv_dof_count, recv_dof_count + ctx->nranks);
    return FC_OK;
  }

  int fc_solid_set_comm(fc_ctx* ctx, void* nccl_comm)
  {
    if (!ctx) return FC_ERR_INVALID;
    ctx->nccl_comm = nccl_comm;
    return FC_OK;
  }

  // eval_error_check (src/structure_new/src/model_evaluator/4C_structure_new_model_evaluator_generic.cpp:225-258): the
  // element error flag of the last evaluation, maximum over the ranks of the communicator given to fc_solid_set_comm
  // (one 4-byte ncclAllReduce).  Synchronises the stream.  One rank: the local flag.
  int fc_solid_error_allreduce(fc_ctx* ctx, int* any_error)
  {
    if (!ctx || !any_error) return FC_ERR_INVALID;
    FC_CUDA(cudaSetDevice(ctx->device));
    k_error_flag<<<1, 1, 0, ctx->stream>>>(ctx->d_err, ctx->geom_bad_count);
    ctx->launches++;
    FC_CUDA(cudaGetLastError());
    if (ctx->nranks > 1)
    {
      NcclApi& api = nccl();
      if (!api.ok || !api.all_reduce) return fail(ctx, FC_ERR_UNSUPPORTED, "NCCL entry points not found in the process and libnccl.so.2 cannot be loaded");
      if (!ctx->nccl_comm) return fail(ctx, FC_ERR_INVALID, "fc_solid_set_comm must precede fc_solid_error_allreduce on numproc > 1");
      const int rc = api.all_reduce(ctx->d_err + 2, ctx->d_err + 2, 1, kNcclInt32, kNcclMax, ctx->nccl_comm, ctx->stream);
      if (rc != 0) return fail(ctx, FC_ERR_CUDA, std::string("NCCL: ") + (api.error_string ? api.error_string(rc) : "error"));
      ctx->launches++;
    }
    int flag = 0;
    FC_CUDA(cudaMemcpyAsync(&flag, ctx->d_err + 2, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
    *any_error = flag;
    return FC_OK;
  }

  int fc_solid_halo_exchange(fc_ctx* ctx, void* nccl_comm, double* u_col_dev)
  {
    if (!ctx || !u_col_dev) return FC_ERR_INVALID;
    if (ctx->nranks == 1) return FC_OK;
    if (!nccl_comm) return fail(ctx, FC_ERR_INVALID, "nccl_comm is required for numproc > 1");
    if ((int)ctx->halo_send_count.size() != ctx->nranks) return fail(ctx, FC_ERR_INVALID, "fc_solid_halo_set_counts must precede halo_exchange");
    if (!is_device_ptr(u_col_dev)) return fail(ctx, FC_ERR_INVALID, "u_col_dev must be a device pointer");
    NcclApi& api = nccl();
    if (!api.ok) return fail(ctx, FC_ERR_UNSUPPORTED, "NCCL entry points not found in the process and libnccl.so.2 cannot be loaded");
    FC_CUDA(cudaSetDevice(ctx->device));
    FC_CUDA(ensure_cap(ctx->d_halo_sendbuf, ctx->halo_sendbuf_cap, (size_t)ctx->n_send));
    if (ctx->n_send > 0)
    {
      fc::k_halo_pack<<<(ctx->n_send + 255) / 256, 256, 0, ctx->stream>>>(u_col_dev, ctx->d_send_lids, ctx->n_send, ctx->d_halo_sendbuf);
      ctx->launches++;
      FC_CUDA(cudaGetLastError());
    }
    int rc = api.group_start();
    size_t soff = 0, roff = 3 * (size_t)ctx->n_row_nodes;
    for (int p = 0; p < ctx->nranks && rc == 0; ++p)
    {
      const size_t ns = (size_t)ctx->halo_send_count[p], nr = (size_t)ctx->halo_recv_count[p];
      if (ns) rc = api.send(ctx->d_halo_sendbuf + soff, ns, kNcclFloat64, p, nccl_comm, ctx->stream);
      if (nr && rc == 0) rc = api.recv(u_col_dev + roff, nr, kNcclFloat64, p, nccl_comm, ctx->stream);
      soff += ns;
      roff += nr;
    }
    const int rc_end = api.group_end();
    if (rc == 0) rc = rc_end;
    if (rc != 0) return fail(ctx, FC_ERR_CUDA, std::string("NCCL: ") + (api.error_string ? api.error_string(rc) : "error"));
    ctx->launches++;  // the grouped send/recv kernel
    return FC_OK;
  }
}
