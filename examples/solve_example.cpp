// Two ranks, two GPUs, one process: the Newton linear system K x = b of a clamped hex8 Neo-Hooke block solved by
// fc_solid_linear_solve on the partitioned system (halo import of the search direction before every y = K x, one NCCL
// all-reduce per dot product) and, for comparison, on one rank holding the whole mesh.  The two solutions must agree.
//   g++ -std=c++17 examples/solve_example.cpp -I. -I$CUDA_HOME/include -L4c_b200/lib -lfourc_b200 -L$CUDA_HOME/lib64 -lcudart -lnccl -lpthread
#include <cmath>
#include <cstdio>
#include <map>
#include <thread>
#include <vector>

#include <cuda_runtime_api.h>
#include <nccl.h>

#include "include/fourc_b200.h"

struct Rank
{
  fc_ctx* ctx = nullptr;
  std::vector<int> gid, owner, egid, eowner, conn;
  std::vector<double> X, x;
  int n_row_nodes = 0, n_col_nodes = 0, n_col_eles = 0, n_row_dof = 0, n_col_dof = 0, iters = 0, rc = 0;
  int64_t nnz = 0;
  double relres = 0.0;
  double *d_u = nullptr, *d_K = nullptr, *d_b = nullptr, *d_x = nullptr;
};

#define CHECK(call)                                                                  \
  do                                                                                 \
  {                                                                                  \
    int rc__ = (call);                                                               \
    if (rc__ != FC_OK)                                                               \
    {                                                                                \
      std::printf("rank %d: %s -> %d (%s)\n", r, #call, rc__, fc_last_error(R.ctx)); \
      return 1;                                                                      \
    }                                                                                \
  } while (0)

// mesh, graph, halo plan, Dirichlet toggle (x = 0 face) and right-hand side (transverse load on the x = 1 face)
static int setup(Rank& R, int r, int nranks, int device, const int* n)
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
  CHECK(fc_solid_create(&R.ctx, device, r, nranks));
  CHECK(fc_solid_set_mesh(R.ctx, FC_HEX8, FC_FORM_DISPLACEMENT, R.n_col_nodes, R.X.data(), R.gid.data(), R.owner.data(), R.n_col_eles,
      R.conn.data(), R.egid.data()));
  const double par[2] = {10.0, 0.25};
  CHECK(fc_solid_set_material(R.ctx, FC_MAT_ELASTHYPER_COUPNEOHOOKE, par, 2));
  CHECK(fc_solid_build_graph(R.ctx));
  CHECK(fc_solid_graph_sizes(R.ctx, &R.n_row_dof, &R.n_col_dof, &R.nnz));
  if (nranks > 1)
  {
    CHECK(fc_solid_halo_set_send(R.ctx, (int)slids.size(), slids.data()));
    CHECK(fc_solid_halo_set_counts(R.ctx, scount.data(), rcount.data()));
  }
  std::vector<int> fixed;
  std::vector<double> b((size_t)R.n_row_dof, 0.0);
  for (int i = 0; i < R.n_row_nodes; ++i)
  {
    if (R.X[3 * (size_t)i] < 1e-12)
      for (int j = 0; j < 3; ++j) fixed.push_back(3 * i + j);
    if (R.X[3 * (size_t)i] > 1.0 - 1e-12) b[3 * (size_t)i + 2] = 1e-3;
  }
  CHECK(fc_solid_set_dirichlet_toggle(R.ctx, (int)fixed.size(), fixed.data()));
  cudaSetDevice(device);
  cudaMalloc((void**)&R.d_u, sizeof(double) * (size_t)R.n_col_dof);
  cudaMalloc((void**)&R.d_K, sizeof(double) * (size_t)R.nnz);
  cudaMalloc((void**)&R.d_b, sizeof(double) * (size_t)R.n_row_dof);
  cudaMalloc((void**)&R.d_x, sizeof(double) * (size_t)R.n_row_dof);
  cudaMemset(R.d_u, 0, sizeof(double) * (size_t)R.n_col_dof);
  cudaMemset(R.d_x, 0, sizeof(double) * (size_t)R.n_row_dof);
  cudaMemcpy(R.d_b, b.data(), sizeof(double) * b.size(), cudaMemcpyHostToDevice);
  return 0;
}

// tangent at u = 0, Dirichlet application, solve; every rank runs this (its own thread when nranks > 1)
static int solve(Rank& R, int r, int method)
{
  int nbad = 0, gid = -1;
  CHECK(fc_solid_evaluate(R.ctx, FC_ACTION_NLNSTIFF, FC_EVAL_ZERO_FIRST, R.d_u, nullptr, R.d_K, &nbad, &gid));
  if (nbad) return 1;
  CHECK(fc_solid_apply_dirichlet(R.ctx, R.d_K, 1));
  CHECK(fc_solid_apply_dirichlet_to_system(R.ctx, R.d_x, R.d_b, nullptr));
  CHECK(fc_solid_linear_solve(R.ctx, method, R.d_K, R.d_b, R.d_x, 1e-11, 5000, &R.iters, &R.relres));
  R.x.resize((size_t)R.n_row_dof);
  cudaMemcpy(R.x.data(), R.d_x, sizeof(double) * R.x.size(), cudaMemcpyDeviceToHost);
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
  const int n[3] = {8, 6, 5};
  ncclComm_t comms[2];
  const int devs[2] = {0, 1};
  if (ncclCommInitAll(comms, 2, devs) != ncclSuccess)
  {
    std::printf("ncclCommInitAll failed\n");
    return 1;
  }
  for (int method = 0; method < 2; ++method)
  {
    Rank whole, R[2];
    if (setup(whole, 0, 1, 0, n) || solve(whole, 0, method)) return 1;
    std::map<int, int> where;  // node gid -> row node of the one-rank system
    for (int i = 0; i < whole.n_row_nodes; ++i) where[whole.gid[i]] = i;
    for (int r = 0; r < 2; ++r)
    {
      if (setup(R[r], r, 2, r, n)) return 1;
      if (fc_solid_set_comm(R[r].ctx, comms[r]) != FC_OK) return 1;
    }
    std::vector<std::thread> th;
    for (int r = 0; r < 2; ++r)
      th.emplace_back(
          [&, r]
          {
            cudaSetDevice(r);
            R[r].rc = solve(R[r], r, method);
            cudaDeviceSynchronize();
          });
    for (auto& t : th) t.join();
    double maxdiff = 0.0, maxx = 0.0;
    for (int r = 0; r < 2; ++r)
    {
      if (R[r].rc) return 1;
      for (int i = 0; i < R[r].n_row_nodes; ++i)
        for (int j = 0; j < 3; ++j)
        {
          const double ref = whole.x[3 * (size_t)where[R[r].gid[i]] + j];
          maxdiff = std::fmax(maxdiff, std::fabs(R[r].x[3 * (size_t)i + j] - ref));
          maxx = std::fmax(maxx, std::fabs(ref));
        }
    }
    std::printf("method %d: one rank %d iterations (%.2e), two ranks %d/%d iterations (%.2e); max |x| %.4e, max difference %.2e\n", method,
        whole.iters, whole.relres, R[0].iters, R[1].iters, R[0].relres, maxx, maxdiff);
    if (!(maxx > 0.0) || !(maxdiff <= 1e-8 * maxx) || R[0].iters != R[1].iters) return 1;
    for (Rank* q : {&whole, &R[0], &R[1]})
    {
      fc_solid_destroy(q->ctx);
      cudaFree(q->d_u);
      cudaFree(q->d_K);
      cudaFree(q->d_b);
      cudaFree(q->d_x);
    }
  }
  std::printf("SOLVE OK\n");
  return 0;
}
