// Context of one rank/GPU: mesh, graph and scatter maps resident in HBM.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

#include "fc_material.cuh"

struct fc_ctx;

// one level of the geometric multigrid hierarchy (fc_capi_mg.cu); level 0 is the context's own mesh
struct fc_mg_level
{
  fc_ctx* c = nullptr;               // coarse context (levels >= 1), owned by the hierarchy
  int nn[3] = {0, 0, 0};             // lattice nodes per direction
  int n_nodes = 0;
  std::vector<int> h_lat2lid;        // [nn0 nn1 nn2]
  std::vector<unsigned char> h_free; // [3 n_nodes] 1 = free DOF
  int* d_lat2lid = nullptr;
  int* d_lat = nullptr;              // [n_nodes][3]
  unsigned char* d_free = nullptr;   // [3 n_nodes]
  double* d_K = nullptr;             // tangent of the level (levels >= 1; level 0 uses the caller's)
  double* d_Dinv = nullptr;          // inverse diagonal blocks (levels >= 1; level 0 uses ctx->d_Dinv)
  double omega = 0.6;                // Jacobi damping of the level: 4 / (3 lambda_max(D^-1 K)), estimated by fc_solid_mg_update
  double *d_b = nullptr, *d_x = nullptr, *d_y = nullptr, *d_r = nullptr, *d_u = nullptr;  // [3 n_nodes] each
  double* d_ev = nullptr;            // [3 n_nodes] last iterate of the power iteration: start of the next estimate
  bool have_ev = false;
  float* d_K32 = nullptr;            // [nnz of the level] single-precision copy of the tangent for the cycle (all levels)
  double* d_Ainv = nullptr;          // [(3 n_nodes)^2] dense inverse of the coarsest level's tangent (small levels only)
  bool have_Ainv = false;
};

struct fc_ctx
{
  int device = 0, rank = 0, nranks = 1;
  cudaStream_t stream = nullptr;
  std::string last_error;

  // mesh
  int celltype = -1, formulation = 0, nen = 0;
  int n_col_nodes = 0, n_row_nodes = 0, n_ele = 0;
  std::vector<int> h_conn;        // caller order, column-node LIDs
  std::vector<int> h_node_owner;  // per column node
  std::vector<int> h_ele_gid;
  double* d_X = nullptr;     // [n_col_nodes][3]
  int* d_conn = nullptr;     // [n_ele][nen]
  int* d_ele_gid = nullptr;  // [n_ele]
  double* d_dNtab = nullptr; // [ngp + nen + 1][nen][3]: Gauss points, node positions, centroid
  double* d_gpw = nullptr;   // [ngp] weights
  // nurbs27 (fc_nurbs.cuh): d_dNtab holds one table per element, built by fc_solid_set_nurbs from
  double* d_nurbs_knots = nullptr;    // [n_ele][3][6] knot spans of the elements and
  double* d_nurbs_weights = nullptr;  // [n_col_nodes] control point weights
  bool nurbs_ready = false;
  unsigned char* d_ele_bad = nullptr;  // [n_ele] 1 = det J <= 0 at an element node
  int* d_ele_perm = nullptr; // element order of the tensor-core assembled-path kernels (L2 sweep order, fc_sweep_order.hpp) or nullptr = caller's order
  int n_interior_ele = 0;    // launch slots [0, n_interior_ele) hold the elements without ghost nodes (FC_EVAL_INTERIOR / _BOUNDARY)
  double* d_geom = nullptr;  // hex8 geometry cache [n_ele][72], built by the first evaluation that uses it
  int geom_bad_count = 0, geom_first_bad_gid = -1;

  // material
  fc::MatParams mat{fc::MAT_NEOHOOKE, 10.0, 0.25, 0, {0, 0, 0, 0}, {{0, 0, 0, 0}, {0, 0, 0, 0}, {0, 0, 0, 0}, {0, 0, 0, 0}}};
  double* d_fiber = nullptr;  // unit fibre directions or nullptr: [n_ele][3] (fc_solid_set_fibers) or [n_ele][ngp][3] (fc_solid_set_nodal_fibers)
  int fiber_gp = 0;           // 0: one fibre per element, 3: one per Gauss point
  double* d_eas = nullptr;    // FC_FORM_EAS_*: [n_ele][3 n + n^2 + 24 n + 1] element history (alpha_inc | alpha | s | Kaa^-1 | Kda | step)
  double* d_mulf = nullptr;   // FC_FORM_MULF: [n_ele][ngp][18] prestress history (inverse Jacobian | deformation gradient, column-major)
  bool have_material = false;

  // graph (node-block CSR: every owned node row lists its coupled column nodes, ascending LID)
  bool have_graph = false;
  int n_row_dof = 0, n_col_dof = 0;
  int64_t nnz = 0;
  std::vector<int64_t> h_nodeptr;  // [n_row_nodes+1] in units of nodes
  std::vector<int> h_nodeadj;      // column-node LIDs
  int64_t* d_node_rowbase = nullptr;  // [n_col_nodes] CSR offset of DOF row 3*n, -1 for ghosts
  int* d_node_rowlen = nullptr;       // [n_col_nodes] entries per DOF row (3*neighbours)
  unsigned char* d_slots = nullptr;   // [n_ele][nen][nen] position of node b in the row of node a
  unsigned char* d_dbc = nullptr;     // [n_row_dof] 1 = row skipped by the matrix scatter
  bool have_dbc = false;
  long long* d_nodeptr = nullptr;     // [n_row_nodes+1] node-block graph on the device (SpMV, Dirichlet application)
  int* d_nodeadj = nullptr;
  int* d_node_diagpos = nullptr;      // [n_row_nodes] position of node a in its own neighbour list

  // Dirichlet toggle (Solid::Dbc's dbcmap) and linear-solver workspace
  int n_toggle = 0;
  int* d_toggle_rows = nullptr;       // [n_toggle] DOF row lids
  double* d_solver_ws = nullptr;      // Krylov vectors
  size_t solver_ws_cap = 0;
  double* d_Dinv = nullptr;           // [n_row_nodes][9] inverse diagonal blocks
  size_t Dinv_cap = 0;
  std::vector<int> h_toggle_rows;     // host copy of the toggle (the multigrid hierarchy injects it into the coarse levels)
  long long toggle_version = 0;
  std::vector<fc_mg_level> mg;        // geometric multigrid hierarchy (fc_solid_mg_setup), empty = none
  long long mg_toggle_version = -1;   // toggle the coarse Dirichlet maps were derived from
  bool mg_operators_ready = false;    // fc_solid_mg_update ran for the current state
  int mg_nu = 2;                      // pre- and post-smoothing sweeps of the V-cycle
  bool mg_fine_omega_stale = true;
  double mg_damping = 4.0 / 3.0;      // omega = mg_damping / lambda_max(D^-1 K) estimate
  bool mg_fp32 = true;                // the V-cycle reads single-precision copies of the tangents    // the damping of level 0 is estimated at the first cycle after fc_solid_mg_update
  double* d_dot = nullptr;            // block partials [3][592] + results [3]
  double* h_dot_pinned = nullptr;

  // brick plan of the owner-computed hex8 assembly (fc_brick_plan.hpp): built by the first evaluation that can use it;
  // n_brick_units == 0 = no plan (not a lattice of hex cells), per-element reductions
  bool brick_tried = false;
  int n_brick_units = 0, n_brick_interior = 0;
  int* d_brick_planes = nullptr;
  long long* d_brick_node_off = nullptr;
  long long* d_brick_ele_off = nullptr;
  int* d_brick_nodes = nullptr;
  int* d_brick_eles = nullptr;
  unsigned long long* d_brick_rows = nullptr;
  bool split_store = false;  // an FC_EVAL_INTERIOR evaluation with FC_EVAL_ZERO_FIRST ran on the brick kernel: the
                             // FC_EVAL_BOUNDARY evaluation that completes it stores its rows as well

  // halo send plan
  int n_send = 0;
  int* d_send_lids = nullptr;
  std::vector<int> halo_send_count, halo_recv_count;  // DOFs per peer rank (fc_solid_halo_set_counts)
  double* d_halo_sendbuf = nullptr;
  void* nccl_comm = nullptr;  // caller's ncclComm_t for the collectives of the linear solve (fc_solid_set_comm)
  size_t halo_sendbuf_cap = 0;

  // scratch + error state
  double* d_stage_u = nullptr;  // staging for host callers
  double* d_stage_f = nullptr;
  double* d_stage_K = nullptr;
  size_t stage_u_cap = 0, stage_f_cap = 0, stage_K_cap = 0;
  int* d_err = nullptr;  // [0] = #failing elements, [1] = min failing gid (INT_MAX if none), [2] = flag of the cross-rank reduction
  int* h_err_pinned = nullptr;

  // device timing: every evaluate / spmv records (start, after zero-fill, end) into the next triple of a ring, so a
  // caller can read the kernel times of a whole timed region afterwards without synchronising inside it
  static constexpr int ev_ring = 128;
  cudaEvent_t ev_hist[ev_ring][3] = {};
  cudaEvent_t* ev = ev_hist[0];  // triple of the current / last call
  long long ev_count = 0;
  bool ev_valid = false;
  int64_t launches = 0;
  int eval_elements_flags = 0;
};
