// Device linear algebra on the node-blocked CSR matrix of the context: Dirichlet application,
// y = K x, node-block Jacobi, the vector kernels of the Krylov solvers.  K never leaves HBM between the element
// evaluation and the solve.
//
// Matrix layout (fc_context.hpp): DOF row 3a+i of owned node a starts at 9 nodeptr[a] + i 3L (L = number of coupled
// nodes) and holds [k][j] = entry for column 3 nodeadj[nodeptr[a]+k] + j.  The column ids are implicit: one int per
// 3x3 block instead of one per entry (Epetra keeps 4 bytes per entry).
#pragma once
#include <cuda_runtime.h>

namespace fc
{
  // SparseMatrix::apply_dirichlet, non-explicit branch (src/core/linalg/src/sparse/4C_linalg_sparsematrix.cpp:1066-1092):
  // memset the row, then 1.0 on the diagonal if diagonalblock.  One warp per Dirichlet row.
  static __global__ void k_apply_dirichlet(const int* __restrict__ rows, int n, const long long* __restrict__ nodeptr,
      const int* __restrict__ diagpos, double* __restrict__ K, int diagonalblock)
  {
    const int w = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (w >= n) return;
    const int r = rows[w], a = r / 3, i = r - 3 * a;
    const long long p0 = nodeptr[a];
    const int L3 = 3 * (int)(nodeptr[a + 1] - p0);
    double* row = K + 9 * p0 + (long long)i * L3;
    for (int m = lane; m < L3; m += 32) row[m] = 0.0;
    __syncwarp();
    if (lane == 0 && diagonalblock && diagpos[a] >= 0) row[3 * diagpos[a] + i] = 1.0;
  }

  // apply_dirichlet_to_system(x, b, dbcval, toggle) (src/core/linalg/src/sparse/4C_linalg_utils_sparse_algebra_assemble.cpp:
  // 126-140): x[i] = b[i] = dbcval[i] on the Dirichlet rows; dbcval == nullptr stands for the zero vector the structure
  // time integration passes (Solid::Dbc::apply_dirichlet_to_local_rhs, 4C_structure_new_dbc.cpp:244-254)
  static __global__ void k_apply_dirichlet_vec(const int* __restrict__ rows, int n, double* x, double* b, const double* __restrict__ dbcval)
  {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int r = rows[t];
    const double v = dbcval ? dbcval[r] : 0.0;
    if (x) x[r] = v;
    if (b) b[r] = v;
  }

  // y = K x.  One warp per owned node: the three DOF rows share the gathered x.
  static __global__ void k_spmv(const long long* __restrict__ nodeptr, const int* __restrict__ nodeadj, int n_row_nodes,
      const double* __restrict__ K, const double* __restrict__ x, double* __restrict__ y)
  {
    const int a = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (a >= n_row_nodes) return;
    const long long p0 = nodeptr[a];
    const int L3 = 3 * (int)(nodeptr[a + 1] - p0);
    const double* r0 = K + 9 * p0;
    const int* adj = nodeadj + p0;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0;
    for (int m = lane; m < L3; m += 32)
    {
      const int k = m / 3;
      const double xv = __ldg(x + 3 * (long long)__ldg(adj + k) + (m - 3 * k));
      s0 += __ldg(r0 + m) * xv;
      s1 += __ldg(r0 + L3 + m) * xv;
      s2 += __ldg(r0 + 2 * L3 + m) * xv;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
      s0 += __shfl_xor_sync(0xFFFFFFFFu, s0, o);
      s1 += __shfl_xor_sync(0xFFFFFFFFu, s1, o);
      s2 += __shfl_xor_sync(0xFFFFFFFFu, s2, o);
    }
    if (lane < 3) y[3 * (long long)a + lane] = lane == 0 ? s0 : (lane == 1 ? s1 : s2);
  }

  // (measured on B200 at C2: this kernel with 128-thread blocks 0.408 ms = 5.21 TB/s; with streaming loads (ld.global.cs) of K
  // 0.428 ms - the 8-byte aligned rows share sectors between passes, L1 keeps them; a variant striding over the 9 L contiguous
  // values of a node: 0.80 ms; a variant issuing the loads of all three passes first: 0.55 ms)

  // inverse of the 3x3 diagonal block of every owned node (node-block Jacobi); a singular block falls back to identity
  static __global__ void k_block_jacobi_setup(const long long* __restrict__ nodeptr, const int* __restrict__ diagpos, int n_row_nodes,
      const double* __restrict__ K, double* __restrict__ Dinv)
  {
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n_row_nodes) return;
    const long long p0 = nodeptr[a];
    const int L3 = 3 * (int)(nodeptr[a + 1] - p0);
    const int dp = diagpos[a];
    const double* b = K + 9 * p0 + 3 * (dp < 0 ? 0 : dp);
    double m[9];
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) m[3 * i + j] = dp < 0 ? 0.0 : b[(long long)i * L3 + j];
    const double c0 = m[4] * m[8] - m[5] * m[7], c1 = m[5] * m[6] - m[3] * m[8], c2 = m[3] * m[7] - m[4] * m[6];
    const double det = m[0] * c0 + m[1] * c1 + m[2] * c2;
    double* o = Dinv + 9 * (long long)a;
    if (!(fabs(det) > 0.0) || !isfinite(det))
    {
      for (int i = 0; i < 9; ++i) o[i] = (i % 4 == 0) ? 1.0 : 0.0;
      return;
    }
    const double id = 1.0 / det;
    o[0] = c0 * id;
    o[1] = (m[2] * m[7] - m[1] * m[8]) * id;
    o[2] = (m[1] * m[5] - m[2] * m[4]) * id;
    o[3] = c1 * id;
    o[4] = (m[0] * m[8] - m[2] * m[6]) * id;
    o[5] = (m[2] * m[3] - m[0] * m[5]) * id;
    o[6] = c2 * id;
    o[7] = (m[1] * m[6] - m[0] * m[7]) * id;
    o[8] = (m[0] * m[4] - m[1] * m[3]) * id;
  }

  // z = Dinv r (node blocks)
  static __global__ void k_block_jacobi_apply(const double* __restrict__ Dinv, int n_row_nodes, const double* __restrict__ r, double* __restrict__ z)
  {
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n_row_nodes) return;
    const double* d = Dinv + 9 * (long long)a;
    const double r0 = r[3 * (long long)a], r1 = r[3 * (long long)a + 1], r2 = r[3 * (long long)a + 2];
    z[3 * (long long)a + 0] = d[0] * r0 + d[1] * r1 + d[2] * r2;
    z[3 * (long long)a + 1] = d[3] * r0 + d[4] * r1 + d[5] * r2;
    z[3 * (long long)a + 2] = d[6] * r0 + d[7] * r1 + d[8] * r2;
  }

  // deterministic dot products: up to 3 pairs in one pass, fixed grid, block partials summed by the last kernel
  constexpr int dot_blocks = 592, dot_threads = 256;  // 4 CTAs per SM on 148 SMs
  static __global__ void __launch_bounds__(dot_threads) k_dot3_partial(int n, const double* __restrict__ a0, const double* __restrict__ b0,
      const double* __restrict__ a1, const double* __restrict__ b1, const double* __restrict__ a2, const double* __restrict__ b2, double* __restrict__ part)
  {
    __shared__ double sh[3][dot_threads / 32];
    double s0 = 0.0, s1 = 0.0, s2 = 0.0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    {
      s0 += a0[i] * b0[i];
      if (a1) s1 += a1[i] * b1[i];
      if (a2) s2 += a2[i] * b2[i];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
      s0 += __shfl_xor_sync(0xFFFFFFFFu, s0, o);
      s1 += __shfl_xor_sync(0xFFFFFFFFu, s1, o);
      s2 += __shfl_xor_sync(0xFFFFFFFFu, s2, o);
    }
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0)
    {
      sh[0][w] = s0;
      sh[1][w] = s1;
      sh[2][w] = s2;
    }
    __syncthreads();
    if (threadIdx.x < 3)
    {
      double s = 0.0;
      for (int k = 0; k < dot_threads / 32; ++k) s += sh[threadIdx.x][k];
      part[threadIdx.x * dot_blocks + blockIdx.x] = s;
    }
  }

  static __global__ void k_dot3_final(const double* __restrict__ part, int nblocks, double* __restrict__ out)
  {
    __shared__ double sh[32];
    const int q = blockIdx.x;  // which of the 3 dots
    double s = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += blockDim.x) s += part[q * dot_blocks + i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0)
    {
      double t = 0.0;
      for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += sh[k];
      out[q] = t;
    }
  }

  // ---- CG with the scalars kept on the device: no host synchronisation inside an iteration --------------------------
  // sc[0] = r.z, sc[1] = p.Ap, sc[2] = r.r, sc[3] = b.b
  // fused update: alpha = r.z / p.Ap; x += alpha p; r -= alpha Ap; z = Dinv r; block partials of r.z and r.r
  static __global__ void __launch_bounds__(dot_threads) k_cg_update(int n_nodes, const double* __restrict__ sc, const double* __restrict__ p,
      const double* __restrict__ Ap, const double* __restrict__ Dinv, double* __restrict__ x, double* __restrict__ r, double* __restrict__ z,
      double* __restrict__ part)
  {
    __shared__ double sh[2][dot_threads / 32];
    const double alpha = sc[1] > 0.0 ? sc[0] / sc[1] : 0.0;  // p.Ap <= 0: converged to rounding or not SPD, the host decides
    double s0 = 0.0, s1 = 0.0;
    for (long long a = (long long)blockIdx.x * blockDim.x + threadIdx.x; a < n_nodes; a += (long long)gridDim.x * blockDim.x)
    {
      double rv[3], zv[3];
#pragma unroll
      for (int i = 0; i < 3; ++i)
      {
        const long long k = 3 * a + i;
        x[k] += alpha * p[k];
        rv[i] = r[k] - alpha * Ap[k];
        r[k] = rv[i];
      }
      const double* d = Dinv + 9 * a;
#pragma unroll
      for (int i = 0; i < 3; ++i)
      {
        zv[i] = d[3 * i] * rv[0] + d[3 * i + 1] * rv[1] + d[3 * i + 2] * rv[2];
        z[3 * a + i] = zv[i];
        s0 += rv[i] * zv[i];
        s1 += rv[i] * rv[i];
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
      s0 += __shfl_xor_sync(0xFFFFFFFFu, s0, o);
      s1 += __shfl_xor_sync(0xFFFFFFFFu, s1, o);
    }
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0)
    {
      sh[0][w] = s0;
      sh[1][w] = s1;
    }
    __syncthreads();
    if (threadIdx.x < 2)
    {
      double t = 0.0;
      for (int k = 0; k < dot_threads / 32; ++k) t += sh[threadIdx.x][k];
      part[threadIdx.x * dot_blocks + blockIdx.x] = t;
    }
  }

  // one block: sums the partials of r.z (-> beta, new r.z) and r.r; p.Ap partials in mode 1.  host_out (pinned, mapped by a
  // later async copy) is not touched here.
  static __global__ void k_cg_scalars(const double* __restrict__ part, int nblocks, int mode, double* __restrict__ sc)
  {
    __shared__ double sh[2][8];
    double s0 = 0.0, s1 = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += blockDim.x)
    {
      s0 += part[i];
      if (mode == 0) s1 += part[dot_blocks + i];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
      s0 += __shfl_xor_sync(0xFFFFFFFFu, s0, o);
      s1 += __shfl_xor_sync(0xFFFFFFFFu, s1, o);
    }
    if ((threadIdx.x & 31) == 0)
    {
      sh[0][threadIdx.x >> 5] = s0;
      sh[1][threadIdx.x >> 5] = s1;
    }
    __syncthreads();
    if (threadIdx.x == 0)
    {
      double t0 = 0.0, t1 = 0.0;
      for (int k = 0; k < (int)(blockDim.x >> 5); ++k)
      {
        t0 += sh[0][k];
        t1 += sh[1][k];
      }
      if (mode == 1)
        sc[1] = t0;  // p.Ap
      else
      {
        sc[4] = sc[0] > 0.0 ? t0 / sc[0] : 0.0;  // beta
        sc[0] = t0;          // r.z
        sc[2] = t1;          // r.r
      }
    }
  }

  // numproc > 1: the sums arrive all-reduced in t[0..1]; same bookkeeping as k_cg_scalars
  static __global__ void k_cg_derive(int mode, const double* __restrict__ t, double* __restrict__ sc)
  {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (mode == 1)
      sc[1] = t[0];
    else
    {
      sc[4] = sc[0] > 0.0 ? t[0] / sc[0] : 0.0;
      sc[0] = t[0];
      sc[2] = t[1];
    }
  }

  // p = z + beta p, beta = sc[4]
  static __global__ void k_cg_direction(int n, const double* __restrict__ sc, const double* __restrict__ z, double* __restrict__ p)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = z[i] + sc[4] * p[i];
  }

  // y = a x + b y
  static __global__ void k_axpby(int n, double a, const double* __restrict__ x, double b, double* __restrict__ y)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = a * x[i] + (b == 0.0 ? 0.0 : b * y[i]);
  }

  // z = x + a y + b w   (w may be nullptr)
  static __global__ void k_xpaypbw(int n, const double* __restrict__ x, double a, const double* __restrict__ y, double b, const double* __restrict__ w, double* __restrict__ z)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) z[i] = x[i] + a * y[i] + (w ? b * w[i] : 0.0);
  }
}  // namespace fc
