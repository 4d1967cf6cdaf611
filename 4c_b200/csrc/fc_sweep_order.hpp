// Element order for L2 locality of the per-element scatter on large meshes (host integer code, no CUDA).
//
// A matrix row receives its element contributions between the first and the last element that touches the node; if
// the elements in between write more K values than L2 holds, the partially summed rows are evicted and fetched again.
// On a z-major sweep that distance is one layer (nx * ny elements, 2 KB of K each).  This order cuts the layers into
// y blocks of at most `layer_elems` elements and sweeps block by block: (y block, z, y, x) on the lattice of the
// mean element extent.  Stands where the reference walks my_col_element_range() in map order
// (src/core/fem/src/discretization/4C_fem_discretization_evaluate.cpp:83).
#pragma once
#include <vector>

namespace fc_host
{
  // Returns false (perm left empty) when a layer already fits, i.e. the caller's order is kept.  nen = nodes per element.
  bool l2_sweep_order(int n_col_nodes, const double* X, int n_ele, int nen, const int* conn, long long layer_elems, std::vector<int>& perm);
}  // namespace fc_host
