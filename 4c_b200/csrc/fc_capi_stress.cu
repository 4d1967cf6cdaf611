// fc_solid_evaluate_stress: struct_calc_stress post-processing.
#include <cmath>

#include "fc_host_util.hpp"
#include "fc_kernels_stress.cuh"

namespace
{
  // in-place inverse of a dense n x n matrix (row-major), Gauss-Jordan with partial pivoting
  bool invert_dense(std::vector<double>& M, int n)
  {
    std::vector<double> I((size_t)n * n, 0.0);
    for (int i = 0; i < n; ++i) I[(size_t)i * n + i] = 1.0;
    for (int c = 0; c < n; ++c)
    {
      int piv = c;
      for (int r = c + 1; r < n; ++r)
        if (std::fabs(M[(size_t)r * n + c]) > std::fabs(M[(size_t)piv * n + c])) piv = r;
      if (M[(size_t)piv * n + c] == 0.0) return false;
      if (piv != c)
        for (int k = 0; k < n; ++k)
        {
          std::swap(M[(size_t)piv * n + k], M[(size_t)c * n + k]);
          std::swap(I[(size_t)piv * n + k], I[(size_t)c * n + k]);
        }
      const double d = 1.0 / M[(size_t)c * n + c];
      for (int k = 0; k < n; ++k)
      {
        M[(size_t)c * n + k] *= d;
        I[(size_t)c * n + k] *= d;
      }
      for (int r = 0; r < n; ++r)
        if (r != c)
        {
          const double f = M[(size_t)r * n + c];
          if (f != 0.0)
            for (int k = 0; k < n; ++k)
            {
              M[(size_t)r * n + k] -= f * M[(size_t)c * n + k];
              I[(size_t)r * n + k] -= f * I[(size_t)c * n + k];
            }
        }
    }
    M = I;
    return true;
  }

  // Gauss points -> nodes (evaluate_gauss_points_to_nodes_extrapolation_matrix,
  // src/core/fem/src/general/utils/4C_fem_general_utils_gauss_point_extrapolation.cpp:262-292): the shape functions of a BASE
  // cell chosen by the number of Gauss points (:23-85) are evaluated at the Gauss points and inverted (:147-180), then
  // interpolated to the nodes of the cell (:211-260).  hex8 / 8 gp -> hex8, hex27 and hex20 / 27 gp -> hex27, tet10 / 4 gp ->
  // tet4, tet4 / 1 gp -> point1 (every node takes the Gauss-point value), wedge6 / 6 gp -> wedge6; pyramid5 / 8 gp uses the
  // reference's fixed 5 x 8 matrix (:299-378).
  template <int CT>
  bool extrapolation_matrix(std::vector<double>& E)
  {
    constexpr int NEN = fc::Cell<CT>::nen, NGP = fc::Cell<CT>::ngp;
    E.assign((size_t)NEN * NGP, 0.0);
    if (CT == fc::TET4)
    {
      for (int a = 0; a < NEN; ++a) E[a] = 1.0;
      return true;
    }
    if (CT == fc::PYRAMID5)
    {
      // entries of the reference's table by their pattern: rows 0-3 rotate with the base node, row 4 is the apex
      const double lo[4] = {2.408235313815748, -0.6452847075210328, 0.1729035162684118, -0.6452847075210328};
      const double hi[4] = {-0.542209910031327, 0.1452847075210439, -0.03892892005285509, 0.1452847075210439};
      for (int i = 0; i < 4; ++i)
        for (int k = 0; k < 4; ++k)
        {
          E[(size_t)i * 8 + k] = lo[(k - i + 4) % 4];
          E[(size_t)i * 8 + 4 + k] = hi[(k - i + 4) % 4];
        }
      for (int k = 0; k < 4; ++k)
      {
        E[(size_t)4 * 8 + k] = -0.2702847075210531;
        E[(size_t)4 * 8 + 4 + k] = 0.520284707521053;
      }
      return true;
    }
    if (CT == fc::HEX18)
    {
      // 18 Gauss points of a hex cell -> base cell hex8 (:36-44); more points than base nodes: least squares
      // P = (M^T M)^-1 M^T with M[gp][c] = N^hex8_c(xi_gp) (evaluate_projection_gauss_points_to_base_distype, :150-208), then the
      // hex8 functions at the 18 nodes
      std::vector<double> M((size_t)NGP * 8), A(64, 0.0), P((size_t)8 * NGP, 0.0);
      for (int g = 0; g < NGP; ++g)
      {
        double xi[3], w;
        fc::gauss_point<CT>(g, xi, &w);
        fc::shape_function<fc::HEX8>(xi, &M[(size_t)g * 8]);
      }
      for (int c = 0; c < 8; ++c)
        for (int d = 0; d < 8; ++d)
          for (int g = 0; g < NGP; ++g) A[(size_t)c * 8 + d] += M[(size_t)g * 8 + c] * M[(size_t)g * 8 + d];
      if (!invert_dense(A, 8)) return false;
      for (int c = 0; c < 8; ++c)
        for (int g = 0; g < NGP; ++g)
          for (int d = 0; d < 8; ++d) P[(size_t)c * NGP + g] += A[(size_t)c * 8 + d] * M[(size_t)g * 8 + d];
      for (int a = 0; a < NEN; ++a)
      {
        double xi[3], N8[8];
        fc::node_param_coords(CT, a, xi);
        fc::shape_function<fc::HEX8>(xi, N8);
        for (int g = 0; g < NGP; ++g)
          for (int c = 0; c < 8; ++c) E[(size_t)a * NGP + g] += N8[c] * P[(size_t)c * NGP + g];
      }
      return true;
    }
    constexpr int BASE = CT == fc::HEX20 ? fc::HEX27 : CT;  // tet10 -> tet4 is written out below
    constexpr int NB = CT == fc::TET10 ? 4 : fc::Cell<BASE>::nen;
    static_assert(NB == NGP || CT == fc::TET4 || CT == fc::PYRAMID5, "square base system");
    std::vector<double> M((size_t)NGP * NB);
    auto base_shape = [](const double* xi, double* N)
    {
      if (CT == fc::TET10)
      {
        N[0] = 1.0 - xi[0] - xi[1] - xi[2];
        N[1] = xi[0];
        N[2] = xi[1];
        N[3] = xi[2];
      }
      else
        fc::shape_function<BASE>(xi, N);
    };
    for (int g = 0; g < NGP; ++g)
    {
      double xi[3], w;
      fc::gauss_point<CT>(g, xi, &w);
      base_shape(xi, &M[(size_t)g * NB]);
    }
    if (!invert_dense(M, NB)) return false;  // M^-1: [base node][gp]
    for (int a = 0; a < NEN; ++a)
    {
      double xi[3], Nb[NB > 0 ? NB : 1];
      if (CT == fc::TET10)
      {
        fc::node_param_coords(CT, a, xi);
        base_shape(xi, Nb);
      }
      else  // node a of the cell is node a of the base cell (hex20: the first 20 nodes of hex27)
        for (int c = 0; c < NB; ++c) Nb[c] = c == a ? 1.0 : 0.0;
      for (int g = 0; g < NGP; ++g)
      {
        double v = 0.0;
        for (int c = 0; c < NB; ++c) v += Nb[c] * M[(size_t)c * NGP + g];
        E[(size_t)a * NGP + g] = v;
      }
    }
    return true;
  }

  template <int CT, int FORM>
  int stress_impl(fc_ctx* ctx, const fc::EvalArgs& A, int stress_type, double* d_gp, double* d_nodal)
  {
    constexpr int NEN = fc::Cell<CT>::nen, NGP = fc::Cell<CT>::ngp;
    const long long n1 = (long long)ctx->n_ele * NGP;
    if (n1 > 0)
    {
      const unsigned gs = (unsigned)((n1 + 63) / 64);
      if (CT == fc::HEX8 && ctx->formulation == FC_FORM_EAS_MILD)
        fc::k_gp_stress_eas<9><<<gs, 64, 0, ctx->stream>>>(A, stress_type, d_gp);
      else if (CT == fc::HEX8 && ctx->formulation == FC_FORM_EAS_FULL)
        fc::k_gp_stress_eas<21><<<gs, 64, 0, ctx->stream>>>(A, stress_type, d_gp);
      else if (CT == fc::HEX8 && ctx->formulation == FC_FORM_EAS_SHELL)
        fc::k_gp_stress_eas<7><<<gs, 64, 0, ctx->stream>>>(A, stress_type, d_gp);
      else if (CT == fc::HEX8 && ctx->formulation == FC_FORM_EAS_MILD_LINEAR)
        fc::k_gp_stress_eas<9, true><<<gs, 64, 0, ctx->stream>>>(A, stress_type, d_gp);
      else if (CT == fc::HEX8 && ctx->formulation == FC_FORM_EAS_FULL_LINEAR)
        fc::k_gp_stress_eas<21, true><<<gs, 64, 0, ctx->stream>>>(A, stress_type, d_gp);
      else if (CT == fc::HEX8 && ctx->formulation == FC_FORM_SHELL_ANS)
        fc::k_gp_stress_shell<fc::HEX8, 0><<<gs, 64, 0, ctx->stream>>>(A, stress_type, d_gp);
      else if (CT == fc::HEX8 && ctx->formulation == FC_FORM_SHELL_EAS_ANS)
        fc::k_gp_stress_shell<fc::HEX8, 7><<<gs, 64, 0, ctx->stream>>>(A, stress_type, d_gp);
      else if (CT == fc::WEDGE6 && ctx->formulation == FC_FORM_SHELL_ANS)
        fc::k_gp_stress_shell<fc::WEDGE6, 0><<<gs, 64, 0, ctx->stream>>>(A, stress_type, d_gp);
      else if (CT == fc::WEDGE6 && ctx->formulation == FC_FORM_SHELL_EAS_ANS)
        fc::k_gp_stress_shell<fc::WEDGE6, 1><<<gs, 64, 0, ctx->stream>>>(A, stress_type, d_gp);
      else
        fc::k_gp_stress<CT, FORM><<<gs, 64, 0, ctx->stream>>>(A, stress_type, d_gp);
      ctx->launches++;
      FC_CUDA(cudaGetLastError());
    }
    if (!d_nodal) return FC_OK;
    if constexpr (CT == fc::NURBS27)
    {
      std::vector<int> count((size_t)ctx->n_col_nodes, 0);
      for (int n : ctx->h_conn) count[n]++;
      int* d_count = nullptr;
      FC_CUDA(dev_upload(d_count, count.data(), count.size(), ctx->stream));
      FC_CUDA(cudaMemsetAsync(d_nodal, 0, sizeof(double) * 6 * (size_t)ctx->n_row_nodes, ctx->stream));
      if (ctx->n_ele > 0)
      {
        fc::k_extrapolate_nurbs27<<<(ctx->n_ele + 3) / 4, 128, 0, ctx->stream>>>(ctx->d_conn, ctx->n_ele, ctx->n_row_nodes, ctx->d_dNtab, d_count,
            d_gp, d_nodal);
        ctx->launches += 2;
      }
      cudaError_t e = cudaGetLastError();
      if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
      cudaFree(d_count);
      FC_CUDA(e);
      return FC_OK;
    }
    else
    {
    std::vector<double> E;
    if (!extrapolation_matrix<CT>(E)) return fail(ctx, FC_ERR_INVALID, "singular Gauss point extrapolation matrix");
    std::vector<int> count((size_t)ctx->n_col_nodes, 0);
    for (int n : ctx->h_conn) count[n]++;
    double* d_E = nullptr;
    int* d_count = nullptr;
    FC_CUDA(dev_upload(d_E, E.data(), E.size(), ctx->stream));
    FC_CUDA(dev_upload(d_count, count.data(), count.size(), ctx->stream));
    FC_CUDA(cudaMemsetAsync(d_nodal, 0, sizeof(double) * 6 * (size_t)ctx->n_row_nodes, ctx->stream));
    const long long n2 = (long long)ctx->n_ele * NEN;
    if (n2 > 0)
    {
      fc::k_extrapolate_to_nodes<CT><<<(unsigned)((n2 + 127) / 128), 128, 0, ctx->stream>>>(ctx->d_conn, ctx->n_ele, ctx->n_row_nodes, d_E,
          d_count, d_gp, d_nodal);
      ctx->launches += 2;
    }
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_E);
    cudaFree(d_count);
    FC_CUDA(e);
    return FC_OK;
    }
  }
}  // namespace

// Gauss-point output of kind fc::GpOutput and its extrapolation to the nodes
static int gp_output_run(fc_ctx* ctx, const double* u_col, int stress_type, double* gp_stress, double* nodal_stress)
{
  if (!ctx || ctx->celltype < 0) return fail(ctx, FC_ERR_INVALID, "set_mesh must precede evaluate_stress");
  if (!ctx->have_material) return fail(ctx, FC_ERR_INVALID, "set_material must precede evaluate_stress");
  if (!u_col || (!gp_stress && !nodal_stress)) return fail(ctx, FC_ERR_INVALID, "u_col and one output are required");
  if (ctx->celltype == FC_NURBS27 && !ctx->nurbs_ready) return fail(ctx, FC_ERR_INVALID, "nurbs27: fc_solid_set_nurbs must follow set_mesh");
  FC_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->stream;
  const int ngp = fc::cell_ngp(ctx->celltype);
  const size_t n_gp = 6 * (size_t)ctx->n_ele * ngp, n_nod = 6 * (size_t)ctx->n_row_nodes;
  const bool u_dev = is_device_ptr(u_col), g_dev = gp_stress && is_device_ptr(gp_stress), n_dev = nodal_stress && is_device_ptr(nodal_stress);
  const double* d_u = u_col;
  if (!u_dev)
  {
    FC_CUDA(ensure_cap(ctx->d_stage_u, ctx->stage_u_cap, (size_t)ctx->n_col_dof));
    FC_CUDA(cudaMemcpyAsync(ctx->d_stage_u, u_col, sizeof(double) * (size_t)ctx->n_col_dof, cudaMemcpyHostToDevice, s));
    d_u = ctx->d_stage_u;
  }
  double *d_gp = g_dev ? gp_stress : nullptr, *d_nodal = n_dev ? nodal_stress : nullptr;
  double *tmp_gp = nullptr, *tmp_nodal = nullptr;
  if (!d_gp)
  {
    FC_CUDA(cudaMalloc((void**)&tmp_gp, sizeof(double) * (n_gp ? n_gp : 1)));
    d_gp = tmp_gp;
  }
  if (nodal_stress && !d_nodal)
  {
    FC_CUDA(cudaMalloc((void**)&tmp_nodal, sizeof(double) * (n_nod ? n_nod : 1)));
    d_nodal = tmp_nodal;
  }
  fc::EvalArgs A{};
  A.X = ctx->d_X;
  A.conn = ctx->d_conn;
  A.n_ele = ctx->n_ele;
  A.dNtab = ctx->d_dNtab;
  A.u = d_u;
  A.mat = ctx->mat;
  A.fiber = ctx->d_fiber;
  A.fiber_gp = ctx->fiber_gp;
  A.fiber_ele = ctx->fiber_gp ? ctx->fiber_gp * fc::cell_ngp(ctx->celltype) : 3;
  A.mulf = ctx->d_mulf;
  A.eas = ctx->d_eas;
  const bool fbar = ctx->formulation == FC_FORM_FBAR, lin = ctx->formulation == FC_FORM_LINEAR, mulf = ctx->formulation == FC_FORM_MULF;
  const bool mulf_fbar = ctx->formulation == FC_FORM_MULF_FBAR;
  int rc = FC_ERR_UNSUPPORTED;
#define FC_STRESS(CT) \
  rc = lin    ? stress_impl<CT, fc::FORM_LINEAR>(ctx, A, stress_type, d_gp, d_nodal) \
       : mulf ? stress_impl<CT, fc::FORM_MULF>(ctx, A, stress_type, d_gp, d_nodal)   \
              : stress_impl<CT, fc::FORM_DISP>(ctx, A, stress_type, d_gp, d_nodal)
  switch (ctx->celltype)
  {
    case FC_HEX8:
      if (mulf_fbar)
        rc = stress_impl<fc::HEX8, fc::FORM_MULF_FBAR>(ctx, A, stress_type, d_gp, d_nodal);
      else if (fbar)
        rc = stress_impl<fc::HEX8, fc::FORM_FBAR>(ctx, A, stress_type, d_gp, d_nodal);
      else
        FC_STRESS(fc::HEX8);
      break;
    case FC_HEX27: FC_STRESS(fc::HEX27); break;
    case FC_TET10: FC_STRESS(fc::TET10); break;
    case FC_HEX20: FC_STRESS(fc::HEX20); break;
    case FC_TET4: FC_STRESS(fc::TET4); break;
    case FC_WEDGE6: FC_STRESS(fc::WEDGE6); break;
    case FC_HEX18: FC_STRESS(fc::HEX18); break;
    case FC_NURBS27: FC_STRESS(fc::NURBS27); break;
    case FC_PYRAMID5:
      if (mulf_fbar)
        rc = stress_impl<fc::PYRAMID5, fc::FORM_MULF_FBAR>(ctx, A, stress_type, d_gp, d_nodal);
      else if (fbar)
        rc = stress_impl<fc::PYRAMID5, fc::FORM_FBAR>(ctx, A, stress_type, d_gp, d_nodal);
      else
        FC_STRESS(fc::PYRAMID5);
      break;
  }
#undef FC_STRESS
  if (rc == FC_OK && gp_stress && !g_dev) FC_CUDA(cudaMemcpyAsync(gp_stress, d_gp, sizeof(double) * n_gp, cudaMemcpyDeviceToHost, s));
  if (rc == FC_OK && nodal_stress && !n_dev) FC_CUDA(cudaMemcpyAsync(nodal_stress, d_nodal, sizeof(double) * n_nod, cudaMemcpyDeviceToHost, s));
  cudaError_t e = cudaStreamSynchronize(s);
  if (tmp_gp) cudaFree(tmp_gp);
  if (tmp_nodal) cudaFree(tmp_nodal);
  if (rc != FC_OK) return rc;
  FC_CUDA(e);
  return FC_OK;
}

extern "C" int fc_solid_evaluate_stress(fc_ctx* ctx, const double* u_col, int stress_type, double* gp_stress, double* nodal_stress)
{
  if (stress_type != FC_STRESS_2PK && stress_type != FC_STRESS_CAUCHY) return fail(ctx, FC_ERR_UNSUPPORTED, "stress type");
  return gp_output_run(ctx, u_col, stress_type == FC_STRESS_2PK ? fc::OUT_PK2 : fc::OUT_CAUCHY, gp_stress, nodal_stress);
}

extern "C" int fc_solid_evaluate_strain(fc_ctx* ctx, const double* u_col, int strain_type, double* gp_strain, double* nodal_strain)
{
  if (strain_type != FC_STRAIN_GL && strain_type != FC_STRAIN_EA && strain_type != FC_STRAIN_LOG)
    return fail(ctx, FC_ERR_UNSUPPORTED, "strain type");
  return gp_output_run(ctx, u_col, fc::OUT_STRAIN_GL + strain_type, gp_strain, nodal_strain);
}
