// fc_solid_evaluate_stress_noalloc: the nodal-stress post-processing of fc_solid_evaluate_stress (fc_capi_stress.cu) for
// callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  No counterpart
// in the reference.
//
// The existing entry point allocates its scratch per call (Gauss-point stresses, extrapolation matrix, elements-per-node
// counts).  Here that scratch is ONE file-scope device array of a fixed, modest size that every GPU loading the library
// holds until exit.  Host code must not hand the array's NAME to a kernel (it is a host-side stand-in, not a device
// address): the device address is taken once per device with cudaGetSymbolAddress, with that device current, and passed
// on.  Nothing here calls cudaMalloc, cudaFree, cudaMallocHost or any allocating library routine, and nothing synchronises:
// the work is enqueued on the context's stream.
//
// Every call shares the array and finds the previous call's contents in it: the node counts are cleared at the start of
// each call, the Gauss-point stresses and the extrapolation matrix are fully overwritten before they are read.  Calls
// that overlap on the GPU (two streams, two contexts on one device) race on it.
#include "fc_host_util.hpp"
#include "fc_kernels_stress.cuh"

namespace
{
  constexpr int max_gp = FC_NOALLOC_MAX_GAUSS_POINTS;  // n_col_ele * ngp
  constexpr int max_nodes = FC_NOALLOC_MAX_COL_NODES;
  constexpr int max_extrap = 27 * 27;                  // hex27: [nen][ngp]

  // layout in doubles: Gauss-point stresses [max_gp][6] | extrapolation matrix [27*27] | node counts (int) [max_nodes]
  constexpr size_t off_gp = 0;
  constexpr size_t off_E = off_gp + 6 * (size_t)max_gp;
  constexpr size_t off_count = off_E + ((max_extrap + 31) / 32) * 32;
  constexpr size_t scratch_doubles = off_count + (max_nodes + 1) / 2;
  static __device__ __align__(256) double g_stress_scratch[scratch_doubles];  // 1.7 MB per GPU

  // device address of g_stress_scratch per device, resolved at the first call on that device
  constexpr int max_devices = 64;
  double* g_scratch_addr[max_devices] = {};

  int scratch_address(fc_ctx* ctx, double** out)
  {
    if (ctx->device < 0 || ctx->device >= max_devices) return fail(ctx, FC_ERR_INVALID, "noalloc: device ordinal beyond 63");
    FC_CUDA(cudaSetDevice(ctx->device));
    if (!g_scratch_addr[ctx->device])
    {
      double* p = nullptr;
      FC_CUDA(cudaGetSymbolAddress((void**)&p, g_stress_scratch));
      if (!p) return fail(ctx, FC_ERR_CUDA, "noalloc: cudaGetSymbolAddress returned a null address");
      g_scratch_addr[ctx->device] = p;
    }
    *out = g_scratch_addr[ctx->device];
    return FC_OK;
  }

  // elements per column node (the divisor of the nodal average); count must be zero on entry
  __global__ void k_count_node_elements(const int* __restrict__ conn, long long n_entries, int* __restrict__ count)
  {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n_entries) atomicAdd(count + conn[t], 1);
  }

  // Gauss points -> nodes on the device: E[a][g] = sum_c Nb_c(xi_a) (M^-1)[c][g], M[g][c] = Nb_c(xi_g), the same
  // Gauss-Jordan elimination with partial pivoting as the host code of fc_capi_stress.cu.  One block; thread k owns
  // column k of M and of the identity that becomes the inverse.
  template <int CT>
  __global__ void k_extrapolation_matrix(double* __restrict__ E)
  {
    constexpr int NEN = fc::Cell<CT>::nen, NGP = fc::Cell<CT>::ngp;
    constexpr int NB = CT == fc::TET10 ? 4 : NEN;
    static_assert(NB == NGP && NB <= 32, "square base system, one warp");
    __shared__ double M[NB][NB], I[NB][NB], fcol[NB];
    __shared__ int piv_s;
    const int k = threadIdx.x;
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
        fc::shape_function<CT>(xi, N);
    };
    if (k < NB)
    {
      double xi[3], w, N[NB];
      fc::gauss_point<CT>(k, xi, &w);
      base_shape(xi, N);
      for (int c = 0; c < NB; ++c)
      {
        M[k][c] = N[c];
        I[k][c] = c == k ? 1.0 : 0.0;
      }
    }
    __syncthreads();
    for (int c = 0; c < NB; ++c)
    {
      if (k == 0)
      {
        int piv = c;
        for (int r = c + 1; r < NB; ++r)
          if (fabs(M[r][c]) > fabs(M[piv][c])) piv = r;
        piv_s = M[piv][c] == 0.0 ? -1 : piv;
      }
      __syncthreads();
      const int piv = piv_s;
      if (piv < 0)
      {
        // singular base system (cannot happen for the three cell types): make the result loud
        for (int i = k; i < NEN * NGP; i += blockDim.x) E[i] = nan("");
        return;
      }
      if (k < NB && piv != c)
      {
        double t = M[piv][k];
        M[piv][k] = M[c][k];
        M[c][k] = t;
        t = I[piv][k];
        I[piv][k] = I[c][k];
        I[c][k] = t;
      }
      __syncthreads();
      const double d = 1.0 / M[c][c];
      if (k < NB) fcol[k] = M[k][c];
      __syncthreads();
      if (k < NB)
      {
        M[c][k] *= d;
        I[c][k] *= d;
        for (int r = 0; r < NB; ++r)
        {
          const double f = fcol[r];
          if (r != c && f != 0.0)
          {
            M[r][k] -= f * M[c][k];
            I[r][k] -= f * I[c][k];
          }
        }
      }
      __syncthreads();
    }
    // thread k = Gauss point column g of E
    if (k < NGP)
      for (int a = 0; a < NEN; ++a)
      {
        double xi[3], Nb[NB];
        fc::node_param_coords(CT, a, xi);
        if (CT == fc::TET10)
          base_shape(xi, Nb);
        else
          for (int c = 0; c < NB; ++c) Nb[c] = c == a ? 1.0 : 0.0;
        double v = 0.0;
        for (int c = 0; c < NB; ++c) v += Nb[c] * I[c][k];
        E[a * NGP + k] = v;
      }
  }

  template <int CT, int FORM>
  int stress_noalloc_impl(fc_ctx* ctx, const fc::EvalArgs& A, int stress_type, double* scratch, double* d_nodal)
  {
    constexpr int NEN = fc::Cell<CT>::nen, NGP = fc::Cell<CT>::ngp;
    cudaStream_t s = ctx->stream;
    double* d_gp = scratch + off_gp;
    double* d_E = scratch + off_E;
    int* d_count = (int*)(scratch + off_count);
    // the array keeps the last call's contents: the counts must start at zero
    FC_CUDA(cudaMemsetAsync(d_count, 0, sizeof(int) * (size_t)ctx->n_col_nodes, s));
    FC_CUDA(cudaMemsetAsync(d_nodal, 0, sizeof(double) * 6 * (size_t)ctx->n_row_nodes, s));
    k_extrapolation_matrix<CT><<<1, 32, 0, s>>>(d_E);
    ctx->launches++;
    const long long n_conn = (long long)ctx->n_ele * NEN, n1 = (long long)ctx->n_ele * NGP;
    if (n_conn > 0)
    {
      k_count_node_elements<<<(unsigned)((n_conn + 255) / 256), 256, 0, s>>>(ctx->d_conn, n_conn, d_count);
      fc::k_gp_stress<CT, FORM><<<(unsigned)((n1 + 63) / 64), 64, 0, s>>>(A, stress_type, d_gp);
      fc::k_extrapolate_to_nodes<CT><<<(unsigned)((n_conn + 127) / 128), 128, 0, s>>>(ctx->d_conn, ctx->n_ele, ctx->n_row_nodes, d_E, d_count, d_gp, d_nodal);
      ctx->launches += 3;
    }
    FC_CUDA(cudaGetLastError());
    return FC_OK;
  }
}  // namespace

extern "C" int fc_solid_evaluate_stress_noalloc(fc_ctx* ctx, const double* u_col_dev, int stress_type, double* nodal_stress_dev)
{
  if (!ctx || ctx->celltype < 0) return fail(ctx, FC_ERR_INVALID, "set_mesh must precede evaluate_stress_noalloc");
  if (!ctx->have_material) return fail(ctx, FC_ERR_INVALID, "set_material must precede evaluate_stress_noalloc");
  if (ctx->celltype > FC_TET10 || ctx->formulation >= FC_FORM_LINEAR)  // linear, MULF, EAS
    return fail(ctx, FC_ERR_UNSUPPORTED, "evaluate_stress_noalloc covers hex8, hex27 and tet10 with nonlinear kinematics (use fc_solid_evaluate_stress)");
  if (!u_col_dev || !nodal_stress_dev) return fail(ctx, FC_ERR_INVALID, "u_col_dev and nodal_stress_dev are required");
  if (stress_type != FC_STRESS_2PK && stress_type != FC_STRESS_CAUCHY) return fail(ctx, FC_ERR_UNSUPPORTED, "stress type");
  const long long n_gp = (long long)ctx->n_ele * fc::cell_ngp(ctx->celltype);
  if (n_gp > max_gp)
    return fail(ctx, FC_ERR_INVALID,
        "evaluate_stress_noalloc: " + std::to_string(n_gp) + " Gauss points exceed the fixed scratch of FC_NOALLOC_MAX_GAUSS_POINTS = " +
            std::to_string(max_gp) + " (use fc_solid_evaluate_stress)");
  if (ctx->n_col_nodes > max_nodes)
    return fail(ctx, FC_ERR_INVALID,
        "evaluate_stress_noalloc: " + std::to_string(ctx->n_col_nodes) + " column nodes exceed the fixed scratch of FC_NOALLOC_MAX_COL_NODES = " +
            std::to_string(max_nodes) + " (use fc_solid_evaluate_stress)");
  double* scratch = nullptr;
  if (int rc = scratch_address(ctx, &scratch)) return rc;  // makes ctx->device current
  if (!is_device_ptr(u_col_dev) || !is_device_ptr(nodal_stress_dev))
    return fail(ctx, FC_ERR_INVALID, "evaluate_stress_noalloc works on device buffers (staging a host buffer would allocate)");
  fc::EvalArgs A{};
  A.X = ctx->d_X;
  A.conn = ctx->d_conn;
  A.n_ele = ctx->n_ele;
  A.dNtab = ctx->d_dNtab;
  A.u = u_col_dev;
  A.mat = ctx->mat;
  A.fiber = ctx->d_fiber;
  A.fiber_gp = ctx->fiber_gp;
  A.fiber_ele = ctx->fiber_gp ? ctx->fiber_gp * fc::cell_ngp(ctx->celltype) : 3;
  const bool fbar = ctx->formulation == FC_FORM_FBAR;
  if (ctx->celltype == FC_HEX8)
    return fbar ? stress_noalloc_impl<fc::HEX8, fc::FORM_FBAR>(ctx, A, stress_type, scratch, nodal_stress_dev)
                : stress_noalloc_impl<fc::HEX8, fc::FORM_DISP>(ctx, A, stress_type, scratch, nodal_stress_dev);
  if (ctx->celltype == FC_HEX27) return stress_noalloc_impl<fc::HEX27, fc::FORM_DISP>(ctx, A, stress_type, scratch, nodal_stress_dev);
  if (ctx->celltype == FC_TET10) return stress_noalloc_impl<fc::TET10, fc::FORM_DISP>(ctx, A, stress_type, scratch, nodal_stress_dev);
  return fail(ctx, FC_ERR_UNSUPPORTED, "cell type");
}
