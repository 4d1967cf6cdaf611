// Calling sequence of the reference's seam through the C++ host mirror (4c_b200/host/fourc_b200.hpp):
//   discret.set_state(0, "displacement", disnp); discret.evaluate(params, stiff, nullptr, fint, ...); discret.clear_state();
//   (src/structure_new/src/model_evaluator/4C_structure_new_model_evaluator_structure.cpp:420-423, 1293-1312)
// Builds an n^3 hex8 Neo-Hooke cube with the structured generator, evaluates struct_calc_nlnstiff twice
// (zero + evaluate, then a second += evaluate) and prints checksums.  Used by tests/test_host_cpp.py.
//   g++ -std=c++17 examples/host_example.cpp -I. -L4c_b200/lib -lfourc_b200 -Wl,-rpath,$PWD/4c_b200/lib -o /tmp/host_example
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "4c_b200/host/fourc_b200.hpp"
#include <cuda_runtime_api.h>

using namespace FourC::B200;

int main(int argc, char** argv)
{
  const int n = argc > 1 ? std::atoi(argv[1]) : 4;
  try
  {
    // mesh of rank 0 of 1: Core::IO::GridGenerator::create_rectangular_cuboid_discretization
    fc_grid* grid = nullptr;
    const double lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
    if (fc_grid_create(&grid, FC_HEX8, n, n, n, lo, hi, 1, 0, 0) != FC_OK) throw Exception(-1, "fc_grid_create failed");
    int64_t sizes[6];
    fc_grid_sizes(grid, sizes);
    const int n_row_nodes = (int)sizes[0], n_col_nodes = (int)sizes[1], n_col_eles = (int)sizes[3];
    std::vector<int> gid(n_col_nodes), owner(n_col_nodes), egid(n_col_eles), eowner(n_col_eles), conn(8 * (size_t)n_col_eles);
    std::vector<double> X(3 * (size_t)n_col_nodes);
    fc_grid_get_nodes(grid, gid.data(), owner.data(), X.data());
    fc_grid_get_elements(grid, egid.data(), eowner.data(), conn.data());
    fc_grid_destroy(grid);

    Discretization dis(0, 0, 1);
    dis.set_mesh(CellType::hex8, ElementTechnology::none, X, gid, owner, conn, egid);
    dis.set_material(Material::ElastHyperCoupNeoHooke(10.0, 0.25));
    const int64_t nnz = dis.build_graph();

    // device state: u = 0.01 * (x y, y z, z x)
    std::vector<double> u(3 * (size_t)n_col_nodes);
    for (int i = 0; i < n_col_nodes; ++i)
    {
      const double x = X[3 * i], y = X[3 * i + 1], z = X[3 * i + 2];
      u[3 * i] = 0.01 * x * y;
      u[3 * i + 1] = 0.01 * y * z;
      u[3 * i + 2] = 0.01 * z * x;
    }
    double *d_u, *d_f, *d_K;
    cudaMalloc((void**)&d_u, u.size() * sizeof(double));
    cudaMalloc((void**)&d_f, 3 * (size_t)n_row_nodes * sizeof(double));
    cudaMalloc((void**)&d_K, nnz * sizeof(double));
    cudaMemcpy(d_u, u.data(), u.size() * sizeof(double), cudaMemcpyHostToDevice);

    SparseMatrix stiff;
    stiff.values = d_K;
    stiff.nnz = nnz;
    dis.zero(&stiff, d_f);
    dis.set_state(d_u);
    bool ok = dis.evaluate(ActionType::struct_calc_nlnstiff, &stiff, d_f);
    ok = dis.evaluate(ActionType::struct_calc_nlnstiff, &stiff, d_f) && ok;  // += : everything doubles
    dis.clear_state();
    cudaDeviceSynchronize();

    std::vector<double> K(nnz), f(3 * (size_t)n_row_nodes);
    cudaMemcpy(K.data(), d_K, nnz * sizeof(double), cudaMemcpyDeviceToHost);
    cudaMemcpy(f.data(), d_f, f.size() * sizeof(double), cudaMemcpyDeviceToHost);
    double sK = 0, sf = 0;
    for (double v : K) sK += std::fabs(v);
    for (double v : f) sf += std::fabs(v);
    std::printf("ok=%d nnz=%lld sum_abs_K=%.15e sum_abs_f=%.15e\n", (int)ok, (long long)nnz, sK, sf);
    cudaFree(d_u);
    cudaFree(d_f);
    cudaFree(d_K);

    // error convention: evaluate without state throws like FOUR_C_THROW
    try
    {
      dis.evaluate(ActionType::struct_calc_internalforce, nullptr, d_f);
      std::printf("missing-state: no exception\n");
      return 2;
    }
    catch (const Exception& e)
    {
      std::printf("missing-state: exception code %d\n", e.code);
    }
  }
  catch (const Exception& e)
  {
    std::printf("fourc_b200 exception %d: %s\n", e.code, e.what());
    return 1;
  }
  return 0;
}
