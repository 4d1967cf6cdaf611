// Two ranks, two GPUs, one process: the forward halo import of Discretization::set_state
// (src/core/fem/src/discretization/4C_fem_discretization.cpp:503-558) done by fc_solid_halo_exchange over NCCL,
// followed by struct_calc_nlnstiff on both ranks.  A 4C rank would pass its own ncclComm_t the same way.
//   g++ -std=c++17 examples/halo_example.cpp -I. -I$CUDA_HOME/include -L4c_b200/lib -lfourc_b200 -L$CUDA_HOME/lib64 -lcudart -lnccl -lpthread
#include <cmath>
#include <cstdio>
#include <thread>
#include <vector>

#include <cuda_runtime_api.h>
#include <nccl.h>

#include "include/fourc_b200.h"

static double state(int gid, int j) { return 1e-3 * gid + 0.1 * j; }

struct Rank
{
  fc_ctx* ctx = nullptr;
  std::vector<int> gid, owner, egid, eowner, conn;
  std::vector<double> X;
  int n_row_nodes = 0, n_col_nodes = 0, n_col_eles = 0;
  double* d_u = nullptr;
  int rc = 0;
};

#define CHECK(call)                                                        \
  do                                                                       \
  {                                                                        \
    int rc__ = (call);                                                     \
    if (rc__ != FC_OK)                                                     \
    {                                                                      \
      std::printf("rank %d: %s -> %d (%s)\n", r, #call, rc__, fc_last_error(R.ctx)); \
      return 1;                                                            \
    }                                                                      \
  } while (0)

static int setup(Rank& R, int r, int nranks, const int* n)
{
  fc_grid* g = nullptr;
  const double lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
  if (fc_grid_create(&g, FC_HEX8, n[0], n[1], n[2], lo, hi, nranks, r, 0) != FC_OK) return 1;
  int64_t sizes[6];
  fc_grid_sizes(g, sizes);
  R.n_row_nodes = (int)sizes[0];
  R.n_col_nodes = (int)sizes[1];
  R.n_col_eles = (int)sizes[3];
  R.gid.resize(R.n_col_nodes);
  R.owner.resize(R.n_col_nodes);
  R.X.resize(3 * (size_t)R.n_col_nodes);
  R.egid.resize(R.n_col_eles);
  R.eowner.resize(R.n_col_eles);
  R.conn.resize(8 * (size_t)R.n_col_eles);
  fc_grid_get_nodes(g, R.gid.data(), R.owner.data(), R.X.data());
  fc_grid_get_elements(g, R.egid.data(), R.eowner.data(), R.conn.data());
  std::vector<int> scount(nranks), rcount(nranks);
  fc_grid_halo_counts(g, scount.data(), rcount.data());
  int nsend = 0;
  for (int c : scount) nsend += c;
  std::vector<int> snodes(nsend), slids(3 * (size_t)nsend);
  fc_grid_halo_send_nodes(g, snodes.data());
  fc_grid_destroy(g);
  for (int i = 0; i < nsend; ++i)
    for (int j = 0; j < 3; ++j) slids[3 * i + j] = 3 * snodes[i] + j;
  for (int p = 0; p < nranks; ++p)
  {
    scount[p] *= 3;
    rcount[p] *= 3;
  }
  CHECK(fc_solid_create(&R.ctx, r, r, nranks));
  CHECK(fc_solid_set_mesh(R.ctx, FC_HEX8, FC_FORM_DISPLACEMENT, R.n_col_nodes, R.X.data(), R.gid.data(), R.owner.data(), R.n_col_eles,
      R.conn.data(), R.egid.data()));
  const double par[2] = {10.0, 0.25};
  CHECK(fc_solid_set_material(R.ctx, FC_MAT_ELASTHYPER_COUPNEOHOOKE, par, 2));
  CHECK(fc_solid_build_graph(R.ctx));
  CHECK(fc_solid_halo_set_send(R.ctx, (int)slids.size(), slids.data()));
  CHECK(fc_solid_halo_set_counts(R.ctx, scount.data(), rcount.data()));
  // owned DOFs hold the state, ghosts are poisoned
  std::vector<double> u(3 * (size_t)R.n_col_nodes, std::nan(""));
  for (int i = 0; i < R.n_row_nodes; ++i)
    for (int j = 0; j < 3; ++j) u[3 * i + j] = 1e-2 * state(R.gid[i], j);
  cudaSetDevice(r);
  cudaMalloc((void**)&R.d_u, u.size() * sizeof(double));
  cudaMemcpy(R.d_u, u.data(), u.size() * sizeof(double), cudaMemcpyHostToDevice);
  return 0;
}

int main()
{
  int ndev = 0;
  cudaGetDeviceCount(&ndev);
  if (ndev < 2)
  {
    std::printf("needs 2 GPUs, found %d\n", ndev);
    return 3;
  }
  const int nranks = 2, n[3] = {6, 5, 4};
  Rank R[2];
  for (int r = 0; r < nranks; ++r)
    if (setup(R[r], r, nranks, n)) return 1;
  ncclComm_t comms[2];
  const int devs[2] = {0, 1};
  if (ncclCommInitAll(comms, 2, devs) != ncclSuccess)
  {
    std::printf("ncclCommInitAll failed\n");
    return 1;
  }
  std::vector<std::thread> th;
  for (int r = 0; r < nranks; ++r)
    th.emplace_back(
        [&, r]
        {
          cudaSetDevice(r);
          R[r].rc = fc_solid_halo_exchange(R[r].ctx, comms[r], R[r].d_u);
          cudaDeviceSynchronize();
        });
  for (auto& t : th) t.join();
  int bad = 0;
  double fsum[3] = {0, 0, 0};
  for (int r = 0; r < nranks; ++r)
  {
    if (R[r].rc != FC_OK)
    {
      std::printf("rank %d: halo_exchange -> %d (%s)\n", r, R[r].rc, fc_last_error(R[r].ctx));
      return 1;
    }
    cudaSetDevice(r);
    std::vector<double> u(3 * (size_t)R[r].n_col_nodes);
    cudaMemcpy(u.data(), R[r].d_u, u.size() * sizeof(double), cudaMemcpyDeviceToHost);
    for (int i = 0; i < R[r].n_col_nodes; ++i)
      for (int j = 0; j < 3; ++j)
        if (u[3 * i + j] != 1e-2 * state(R[r].gid[i], j)) ++bad;  // bit-exact copy of the owner's value
    // evaluate with the imported state: host outputs, library stages them
    int n_row_dof = 0, n_col_dof = 0;
    int64_t nnz = 0;
    fc_solid_graph_sizes(R[r].ctx, &n_row_dof, &n_col_dof, &nnz);
    std::vector<double> f(n_row_dof), K(nnz);
    int nbad = 0, gid = -1;
    const int rc = fc_solid_evaluate(R[r].ctx, FC_ACTION_NLNSTIFF, FC_EVAL_ZERO_FIRST, R[r].d_u, f.data(), K.data(), &nbad, &gid);
    if (rc != FC_OK || nbad)
    {
      std::printf("rank %d: evaluate -> %d, %d failing elements (%s)\n", r, rc, nbad, fc_last_error(R[r].ctx));
      return 1;
    }
    for (int i = 0; i < n_row_dof; ++i) fsum[i % 3] += f[i];
    std::printf("rank %d: %d row nodes, %d ghost nodes, nnz %lld\n", r, R[r].n_row_nodes, R[r].n_col_nodes - R[r].n_row_nodes, (long long)nnz);
  }
  // internal forces of the two ranks together are self-equilibrated
  std::printf("ghost mismatches=%d  sum f = %.3e %.3e %.3e\n", bad, fsum[0], fsum[1], fsum[2]);
  const bool ok = bad == 0 && std::fabs(fsum[0]) + std::fabs(fsum[1]) + std::fabs(fsum[2]) < 1e-12;
  std::printf(ok ? "HALO OK\n" : "HALO FAILED\n");
  for (int r = 0; r < nranks; ++r)
  {
    ncclCommDestroy(comms[r]);
    fc_solid_destroy(R[r].ctx);
  }
  return ok ? 0 : 2;
}
