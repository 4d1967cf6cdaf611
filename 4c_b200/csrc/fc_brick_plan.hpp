// Brick plan of the owner-computed hex8 assembly (k_hex8_nh_brick, fc_kernels_hex8nh_brick.cuh).
//
// What it replaces in the reference: SparseMatrix::assemble adds every element matrix into the rows of its owned
// nodes (src/core/linalg/src/sparse/4C_linalg_sparsematrix.cpp:471-575) on a matrix that SparseMatrix::zero re-created
// just before (:353-378).  On the GPU that is a zero-fill plus 576 FP64 reductions per element, and every K line
// crosses DRAM three times (zero, fetch for the first reduction, write-back).  The brick kernel turns the loop
// around: a CTA OWNS a column of node lines (BY x BZ lines, swept along x), evaluates every element that touches
// them, sums the contributions of a node in shared memory and writes each finished CSR row exactly once with
// plain stores.  No zero-fill, no atomics, and the result does not depend on the order in which CTAs run.
//
// The plan is pure host integer code.  It needs a lattice: every node sits on a point of a rectilinear grid and
// every element spans one cell of it (what Core::IO::GridGenerator::create_rectangular_cuboid_discretization,
// src/core/io/src/4C_io_gridgenerator.cpp:41-323, and any rank's share of it produce).  A mesh without that
// structure has no plan and stays on the per-element reduction kernel.
#pragma once
#include <cstdint>
#include <vector>

namespace fc_host
{
  struct BrickPlanHost
  {
    // unit = BY x BZ node lines x up to `chunk` node planes along the sweep axis (lattice axis 0)
    static constexpr int BY = 7, BZ = 5, NL = BY * BZ;           // owned node lines of a unit
    static constexpr int CY = BY + 1, CZ = BZ + 1, NE = CY * CZ;  // element columns that touch them (48)
    int n_units = 0, n_interior_units = 0;  // units [0, n_interior_units) hold no element with a ghost node
    int chunk = 0;
    std::vector<int> unit_planes;          // [n_units] node planes P of the unit (element layers: P + 1)
    std::vector<long long> unit_node_off;  // [n_units] first entry of the unit in node_tab (P * NL entries)
    std::vector<long long> unit_ele_off;   // [n_units] first entry in ele_tab / rowidx_tab ((P + 1) * NE entries)
    std::vector<int> node_tab;             // node of (plane p, line l) or -1; line l = ly + BY * lz
    std::vector<int> ele_tab;              // element of (layer t, slot k) or -1; layer t lies between planes t-1 and t.
                                           // slot k = 12 * colour + warp, colour = (cy & 1) + 2 (cz & 1),
                                           // warp = (cy >> 1) + 4 (cz >> 1): the 12 elements of a colour share no node
    std::vector<unsigned long long> rowidx_tab;  // per (layer, slot): byte a = accumulator row of local node a =
                                                 // (plane & 1) * NL + line, 0xFF = not a row of this unit
    long long n_element_instances = 0;     // non-empty entries of ele_tab (>= n_ele: elements on unit faces repeat)
  };

  // Lattice index of every node along coordinate axis d (uniform spacing, or the sorted distinct levels of a non-uniform
  // rectilinear grid); false when the coordinates do not sit on such a grid.  conn [n_ele][8].  Also used by the
  // geometric multigrid hierarchy (fc_capi_mg.cu).
  bool lattice_axis(int n_nodes, const double* X, int d, int n_ele, const int* conn, std::vector<int>& idx, int& n_levels);

  // X [n_col_nodes][3]; conn [n_ele][8] column-node ids; nodes [0, n_row_nodes) own rows; max_row_nodes = largest
  // number of coupled nodes of an owned node (the kernel keeps 27 blocks per row).  chunk = node planes per unit
  // along the sweep axis (0 = choose: at least ~8 units per SM-wave of 148).  Returns false (plan left empty) when
  // the mesh is not a lattice of hex cells.
  bool build_brick_plan_host(int n_col_nodes, const double* X, int n_ele, const int* conn, int n_row_nodes, int max_row_nodes,
      int chunk, BrickPlanHost& plan);
}  // namespace fc_host
