// fc_solid_evaluate_mass: the consistent mass matrix of struct_calc_nlnstiffmass in the CSR layout of the
// stiffness graph.  One-off per mesh (MASSLIN none): not on the per-Newton-iteration path.
#include "fc_host_util.hpp"
#include "fc_kernels_mass.cuh"

namespace
{
  template <int CT>
  int mass_impl(fc_ctx* ctx, double density, double* d_M)
  {
    constexpr int NEN = fc::Cell<CT>::nen, NGP = fc::MassRule<CT>::ngp, REC = 1 + 4 * NEN;
    std::vector<double> tab((size_t)NGP * REC);
    for (int g = 0; g < NGP; ++g)
    {
      double xi[3];
      fc::mass_gauss_point<CT>(g, xi, &tab[(size_t)g * REC]);
      fc::shape_function<CT>(xi, &tab[(size_t)g * REC + 1]);
      fc::shape_deriv1<CT>(xi, &tab[(size_t)g * REC + 1 + NEN]);
    }
    double* d_tab = nullptr;
    FC_CUDA(dev_upload(d_tab, tab.data(), tab.size(), ctx->stream));
    const long long nthreads = (long long)ctx->n_ele * NEN;
    if (nthreads > 0)
    {
      fc::k_mass<CT><<<(unsigned)((nthreads + 127) / 128), 128, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, ctx->n_ele,
          ctx->n_row_nodes, d_tab, ctx->d_node_rowbase, ctx->d_node_rowlen, ctx->d_slots, density, d_M);
      ctx->launches++;
    }
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_tab);
    FC_CUDA(e);
    return FC_OK;
  }
}  // namespace

extern "C" int fc_solid_evaluate_mass(fc_ctx* ctx, double density, int flags, double* M_values)
{
  if (!ctx || ctx->celltype < 0) return fail(ctx, FC_ERR_INVALID, "set_mesh must precede evaluate_mass");
  if (!ctx->have_graph) return fail(ctx, FC_ERR_INVALID, "set_graph/build_graph must precede evaluate_mass");
  if (!M_values) return fail(ctx, FC_ERR_INVALID, "M_values is required");
  if (!(density > 0.0)) return fail(ctx, FC_ERR_INVALID, "density must be positive");
  FC_CUDA(cudaSetDevice(ctx->device));
  const bool dev = is_device_ptr(M_values);
  double* d_M = M_values;
  if (!dev)
  {
    FC_CUDA(ensure_cap(ctx->d_stage_K, ctx->stage_K_cap, (size_t)ctx->nnz));
    d_M = ctx->d_stage_K;
    if (!(flags & FC_EVAL_ZERO_FIRST))
      FC_CUDA(cudaMemcpyAsync(d_M, M_values, sizeof(double) * (size_t)ctx->nnz, cudaMemcpyHostToDevice, ctx->stream));
  }
  if (flags & FC_EVAL_ZERO_FIRST)
  {
    FC_CUDA(cudaMemsetAsync(d_M, 0, sizeof(double) * (size_t)ctx->nnz, ctx->stream));
    ctx->launches++;
  }
  int rc = FC_ERR_UNSUPPORTED;
  switch (ctx->celltype)
  {
    case FC_HEX8: rc = mass_impl<fc::HEX8>(ctx, density, d_M); break;
    case FC_HEX27: rc = mass_impl<fc::HEX27>(ctx, density, d_M); break;
    case FC_TET10: rc = mass_impl<fc::TET10>(ctx, density, d_M); break;
    case FC_HEX20: rc = mass_impl<fc::HEX20>(ctx, density, d_M); break;
    case FC_TET4: rc = mass_impl<fc::TET4>(ctx, density, d_M); break;
    case FC_WEDGE6: rc = mass_impl<fc::WEDGE6>(ctx, density, d_M); break;
    case FC_PYRAMID5: rc = mass_impl<fc::PYRAMID5>(ctx, density, d_M); break;
    case FC_HEX18: rc = mass_impl<fc::HEX18>(ctx, density, d_M); break;
    case FC_NURBS27:
    {
      if (!ctx->nurbs_ready) return fail(ctx, FC_ERR_INVALID, "nurbs27: fc_solid_set_nurbs must follow set_mesh");
      const long long nthreads = (long long)ctx->n_ele * 27;
      if (nthreads > 0)
      {
        fc::k_mass_nurbs<<<(unsigned)((nthreads + 127) / 128), 128, 0, ctx->stream>>>(ctx->d_X, ctx->d_conn, ctx->n_ele, ctx->n_row_nodes,
            ctx->d_dNtab, ctx->d_gpw, ctx->d_node_rowbase, ctx->d_node_rowlen, ctx->d_slots, density, d_M);
        ctx->launches++;
      }
      FC_CUDA(cudaGetLastError());
      FC_CUDA(cudaStreamSynchronize(ctx->stream));
      rc = FC_OK;
      break;
    }
  }
  if (rc != FC_OK) return rc;
  if (!dev)
  {
    FC_CUDA(cudaMemcpyAsync(M_values, d_M, sizeof(double) * (size_t)ctx->nnz, cudaMemcpyDeviceToHost, ctx->stream));
    FC_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  return FC_OK;
}
