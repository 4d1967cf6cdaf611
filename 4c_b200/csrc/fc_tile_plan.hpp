// Tile plan of the hex8 assembly: elements are grouped into spatially compact tiles of <= 16 (4 x 2 x 2 on a
// structured cube) and every (row node, column node) pair that occurs inside a tile becomes ONE "tile block"
// (3 x 3 doubles).  The tile kernel (fc_kernels_hex8nh_tile.cuh) sums the element contributions of a tile in shared
// memory and adds each tile block to the CSR values once: 637 blocks instead of 1024 element blocks on a structured
// 4 x 2 x 2 tile, and the blocks of a row are consecutive in CSR order, so the reductions arrive as long contiguous runs
// (about 110 instead of 216 32-byte reduction sectors per element).
//
// What this replaces in the reference: the per-element row loop of SparseMatrix::assemble
// (src/core/linalg/src/sparse/4C_linalg_sparsematrix.cpp:474-543), whose 24 ExtractMyRowView calls and lower_bound
// searches per element become a table computed once per graph.  Pure host integer code, no CUDA.
#pragma once
#include <cstdint>
#include <vector>

namespace fc_host
{
  struct TilePlanHost
  {
    static constexpr int tile_eles = 16;
    int n_tiles = 0;
    int max_blocks = 0;                   // largest number of blocks of a tile
    std::vector<int> tile_ele;            // [n_tiles][16] element index, -1 = empty slot
    std::vector<unsigned> tile_blk;       // [n_tiles][16][32] accumulator lane (a, q): block of (a, 2q) | block of (a, 2q+1) << 16; 0xFFFF = row not owned
    std::vector<long long> tile_ptr;      // [n_tiles + 1] first block record of a tile
    std::vector<unsigned long long> rec;  // per block: CSR offset of (DOF row 3a, column 3b) | entries per DOF row << 44 (bits 60..62: Dirichlet flags, set on the device)
    std::vector<int> rownode;             // per block: row node a
  };

  // Element order for L2 locality of the per-element scatter on large meshes.  A matrix row receives its element
  // contributions between the first and the last element that touches the node; if the elements in between write more
  // K values than L2 holds, the partially summed rows are evicted and fetched again.  On a z-major sweep that distance is
  // one layer (nx * ny elements, 2 KB of K each).  This order cuts the layers into y blocks of at most `layer_elems`
  // elements and sweeps block by block: (y block, z, y, x) on the lattice of the mean element extent.  Returns false
  // (perm left empty) when a layer already fits, i.e. the caller's order is kept.  nen = nodes per element.
  bool l2_sweep_order(int n_col_nodes, const double* X, int n_ele, int nen, const int* conn, long long layer_elems, std::vector<int>& perm);

  // X [n_col_nodes][3]; conn [n_ele][8] column-node ids; nodes [0, n_row_nodes) own rows; nodeptr/nodeadj = node-block
  // graph of the owned nodes (ascending column-node ids).  Returns false (plan left empty) if an element lists a node
  // twice or couples a pair that is missing from the graph: the caller then stays on the per-element scatter.
  bool build_tile_plan_host(int n_col_nodes, const double* X, int n_ele, const int* conn, int n_row_nodes,
      const int64_t* nodeptr, const int* nodeadj, TilePlanHost& plan);
}  // namespace fc_host
