// Geometric multigrid on lattice meshes: smoother, residual and grid transfers of the V-cycle that
// preconditions fc_solid_linear_solve(FC_SOLVER_CG_MG).  Stands where the reference hands the tangent to
// Belos + MueLu (src/core/linear_solver); the coarse operators are re-discretisations of the same elements on the
// coarser lattices, evaluated by the library's own element kernels.
//
// Matrix layout: the node-blocked CSR of fc_kernels_linsolve.cuh.  Lattice: node (i, j, k) of a level with (nx, ny, nz)
// elements; the next coarser level keeps the nodes with even indices (and the last node of a direction with an odd number of
// elements).  Prolongation = trilinear interpolation per
// displacement component, restriction = its transpose (a gather: no atomics, deterministic).
#pragma once
#include <cuda_runtime.h>

namespace fc
{
  // one damped node-block Jacobi sweep, or the residual:
  //   Dinv != nullptr:  x_out = x_in + omega Dinv (b - K x_in)          (x_in == nullptr: x_in = 0, no matrix access)
  //   Dinv == nullptr:  x_out = b - K x_in                              (the residual)
  // One warp per owned node: the three DOF rows share the gathered x.  TK = double, or float: the cycle may run on a single-
  // precision copy of the level's tangent (half the bytes of the pass that bounds it; vectors and sums stay FP64).
  template <typename TK>
  static __global__ void k_mg_smooth(const long long* __restrict__ nodeptr, const int* __restrict__ nodeadj, int n_nodes, const TK* __restrict__ K,
      const double* __restrict__ Dinv, double omega, const double* __restrict__ b, const double* __restrict__ x_in, double* __restrict__ x_out)
  {
    const int a = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (a >= n_nodes) return;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0;
    if (x_in)
    {
      const long long p0 = nodeptr[a];
      const int L3 = 3 * (int)(nodeptr[a + 1] - p0);
      const TK* r0 = K + 9 * p0;
      const int* adj = nodeadj + p0;
      for (int m = lane; m < L3; m += 32)
      {
        const int k = m / 3;
        const double xv = __ldg(x_in + 3 * (long long)__ldg(adj + k) + (m - 3 * k));
        s0 += (double)__ldg(r0 + m) * xv;
        s1 += (double)__ldg(r0 + L3 + m) * xv;
        s2 += (double)__ldg(r0 + 2 * L3 + m) * xv;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1)
      {
        s0 += __shfl_xor_sync(0xFFFFFFFFu, s0, o);
        s1 += __shfl_xor_sync(0xFFFFFFFFu, s1, o);
        s2 += __shfl_xor_sync(0xFFFFFFFFu, s2, o);
      }
    }
    if (lane == 0)
    {
      const long long o = 3 * (long long)a;
      const double r[3] = {b[o] - s0, b[o + 1] - s1, b[o + 2] - s2};
      if (!Dinv)
      {
        x_out[o] = r[0];
        x_out[o + 1] = r[1];
        x_out[o + 2] = r[2];
        return;
      }
      const double* d = Dinv + 9 * (long long)a;
      for (int i = 0; i < 3; ++i)
        x_out[o + i] = (x_in ? x_in[o + i] : 0.0) + omega * (d[3 * i] * r[0] + d[3 * i + 1] * r[1] + d[3 * i + 2] * r[2]);
    }
  }

  struct MgLattice
  {
    int nx, ny, nz;      // nodes per direction of this level
    const int* lat2lid;  // [nx ny nz] node id of lattice point (i + nx (j + ny k))
    const int* lat;      // [n_nodes][3] lattice point of every node
  };

  // Coarsening of one lattice axis with ne fine elements: the coarse nodes are the fine nodes 0, 2, 4, ... plus the last
  // node when ne is odd (the last coarse element is then one fine element wide), (ne + 1) / 2 coarse elements.
  // Fine index of coarse node I
  __device__ __forceinline__ int mg_fine_of(int I, int ne) { return min(2 * I, ne); }
  // is fine node i a coarse node
  __device__ __forceinline__ bool mg_is_coarse(int i, int ne) { return (i & 1) == 0 || i == ne; }
  // coarse nodes and weights that fine node i interpolates from: returns the count (1 or 2)
  __device__ __forceinline__ int mg_parents(int i, int ne, int* c, double* w)
  {
    if ((i & 1) == 0)
    {
      c[0] = i >> 1;
      w[0] = 1.0;
      return 1;
    }
    if (i == ne)
    {
      c[0] = (i + 1) >> 1;
      w[0] = 1.0;
      return 1;
    }
    c[0] = (i - 1) >> 1;
    c[1] = (i + 1) >> 1;
    w[0] = w[1] = 0.5;
    return 2;
  }

  // r_c = P^T r_f on the free coarse DOFs (free_c[dof] != 0), zero elsewhere.  One thread per coarse node: a gather.
  static __global__ void k_mg_restrict(MgLattice F, MgLattice Cc, int n_coarse, const double* __restrict__ r_f, const unsigned char* __restrict__ free_c,
      double* __restrict__ r_c)
  {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_coarse) return;
    const int ne[3] = {F.nx - 1, F.ny - 1, F.nz - 1};
    int f0[3];
    for (int d = 0; d < 3; ++d) f0[d] = mg_fine_of(Cc.lat[3 * c + d], ne[d]);
    double s[3] = {0, 0, 0};
    for (int dk = -1; dk <= 1; ++dk)
      for (int dj = -1; dj <= 1; ++dj)
        for (int di = -1; di <= 1; ++di)
        {
          const int g[3] = {f0[0] + di, f0[1] + dj, f0[2] + dk}, dd[3] = {di, dj, dk};
          bool ok = true;
          double w = 1.0;
          for (int d = 0; d < 3; ++d)
            if (dd[d])
            {
              // a neighbour interpolates from this coarse node only if it is not a coarse node itself
              ok = ok && g[d] >= 0 && g[d] <= ne[d] && !mg_is_coarse(g[d], ne[d]);
              w *= 0.5;
            }
          if (!ok) continue;
          const int n = F.lat2lid[g[0] + F.nx * (g[1] + (long long)F.ny * g[2])];
          if (n < 0) continue;
          for (int d = 0; d < 3; ++d) s[d] += w * r_f[3 * (long long)n + d];
        }
    for (int d = 0; d < 3; ++d) r_c[3 * (long long)c + d] = free_c[3 * (long long)c + d] ? s[d] : 0.0;
  }

  // x_f += P x_c on the free fine DOFs.  One thread per fine node.
  static __global__ void k_mg_prolong_add(MgLattice F, MgLattice Cc, int n_fine, const double* __restrict__ x_c, const unsigned char* __restrict__ free_f,
      double* __restrict__ x_f)
  {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= n_fine) return;
    int ci[2], cj[2], ck[2];
    double wi[2], wj[2], wk[2];
    const int ni = mg_parents(F.lat[3 * n], F.nx - 1, ci, wi), nj = mg_parents(F.lat[3 * n + 1], F.ny - 1, cj, wj),
              nk = mg_parents(F.lat[3 * n + 2], F.nz - 1, ck, wk);
    double s[3] = {0, 0, 0};
    for (int c = 0; c < nk; ++c)
      for (int b = 0; b < nj; ++b)
        for (int a = 0; a < ni; ++a)
        {
          const int m = Cc.lat2lid[ci[a] + Cc.nx * (cj[b] + (long long)Cc.ny * ck[c])];
          if (m < 0) continue;
          const double w = wi[a] * wj[b] * wk[c];
          for (int d = 0; d < 3; ++d) s[d] += w * x_c[3 * (long long)m + d];
        }
    for (int d = 0; d < 3; ++d)
      if (free_f[3 * (long long)n + d]) x_f[3 * (long long)n + d] += s[d];
  }

  // u_c = u_f at the coincident nodes (injection): the state the coarse tangent is evaluated at
  static __global__ void k_mg_inject(MgLattice F, MgLattice Cc, int n_coarse, const double* __restrict__ u_f, double* __restrict__ u_c)
  {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_coarse) return;
    const int i = mg_fine_of(Cc.lat[3 * c], F.nx - 1), j = mg_fine_of(Cc.lat[3 * c + 1], F.ny - 1), k = mg_fine_of(Cc.lat[3 * c + 2], F.nz - 1);
    const int n = F.lat2lid[i + F.nx * (j + (long long)F.ny * k)];
    for (int d = 0; d < 3; ++d) u_c[3 * (long long)c + d] = u_f[3 * (long long)n + d];
  }

  // start vector of the power iteration: pseudo-random signs (the largest eigenvectors of D^-1 K oscillate from node
  // to node, a smooth start would hold little of them), zero on the fixed DOFs
  static __global__ void k_mg_seed(int n, const unsigned char* __restrict__ mask, double* __restrict__ v)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned long long z = (unsigned long long)i * 0x9E3779B97F4A7C15ull + 0x632BE59BD9B4E019ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    v[i] = mask[i] ? ((z & 1) ? 1.0 : -1.0) * (0.5 + (double)((z >> 11) & 0xFFFF) / 131072.0) : 0.0;
  }

  static __global__ void k_mg_to_float(long long n, const double* __restrict__ src, float* __restrict__ dst)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = (float)src[i];
  }
  // dst = a src
  static __global__ void k_mg_scale_copy(int n, double a, const double* __restrict__ src, double* __restrict__ dst)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = a * src[i];
  }

  // ---- direct solve of the coarsest level: dense inverse, one CTA ------------------------------------------------------------
  // A [n][n] (row-major, n = 3 n_nodes) = the level's tangent expanded from the node-blocked CSR
  static __global__ void k_mg_dense_fill(const long long* __restrict__ nodeptr, const int* __restrict__ nodeadj, int n_nodes, const double* __restrict__ K,
      double* __restrict__ A)
  {
    const int n = 3 * n_nodes;
    for (long long i = threadIdx.x; i < (long long)n * n; i += blockDim.x) A[i] = 0.0;
    __syncthreads();
    for (int a = threadIdx.x; a < n_nodes; a += blockDim.x)
    {
      const long long p0 = nodeptr[a];
      const int L = (int)(nodeptr[a + 1] - p0), L3 = 3 * L;
      for (int k = 0; k < L; ++k)
      {
        const int b = nodeadj[p0 + k];
        for (int i = 0; i < 3; ++i)
          for (int j = 0; j < 3; ++j) A[(long long)(3 * a + i) * n + 3 * b + j] = K[9 * p0 + (long long)i * L3 + 3 * k + j];
      }
    }
  }
  // in-place Gauss-Jordan inversion without pivoting (the tangent is symmetric positive definite on the free DOFs, identity on
  // the fixed ones); ok[0] = 0 if a pivot is not positive (the caller then keeps the smoother on this level)
  static __global__ void k_mg_dense_invert(int n, double* __restrict__ A, int* __restrict__ ok)
  {
    __shared__ double s_piv;
    if (threadIdx.x == 0) ok[0] = 1;
    for (int k = 0; k < n; ++k)
    {
      __syncthreads();
      if (threadIdx.x == 0)
      {
        s_piv = A[(long long)k * n + k];
        if (!(s_piv > 0.0)) ok[0] = 0;
      }
      __syncthreads();
      const double ip = 1.0 / s_piv;
      // scale the pivot row (its diagonal becomes 1 / pivot)
      for (int j = threadIdx.x; j < n; j += blockDim.x) A[(long long)k * n + j] = (j == k ? 1.0 : A[(long long)k * n + j]) * ip;
      __syncthreads();
      // eliminate column k from every other row: a warp per row
      for (int i = threadIdx.x >> 5; i < n; i += blockDim.x >> 5)
      {
        if (i == k) continue;
        const double f = A[(long long)i * n + k];
        for (int j = threadIdx.x & 31; j < n; j += 32)
          if (j != k) A[(long long)i * n + j] -= f * A[(long long)k * n + j];
      }
      __syncthreads();
      for (int i = threadIdx.x; i < n; i += blockDim.x)
        if (i != k) A[(long long)i * n + k] = -A[(long long)i * n + k] * ip;
    }
  }
  // x = Ainv b, one warp per row
  static __global__ void k_mg_dense_apply(int n, const double* __restrict__ Ainv, const double* __restrict__ b, double* __restrict__ x)
  {
    const int r = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (r >= n) return;
    double s = 0.0;
    for (int j = lane; j < n; j += 32) s += Ainv[(long long)r * n + j] * b[j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
    if (lane == 0) x[r] = s;
  }

  // y = x restricted to the free DOFs (mask != 0), zero elsewhere
  static __global__ void k_mg_mask(int n, const unsigned char* __restrict__ mask, const double* __restrict__ x, double* __restrict__ y)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = mask[i] ? x[i] : 0.0;
  }
}  // namespace fc
