/* fourc_b200.h — C ABI of libfourc_b200.so
 *
 * B200-native (CUDA sm_100a) replacement for ONE hot path of 4C: the per-Newton-iteration
 * evaluation of internal force + tangent stiffness of all 3D solid elements of a rank and the
 * scatter into the rank's CSR system.  Paths below are relative to the reference tree.
 *
 * Reference seam this library is called from:
 *   Solid::ModelEvaluator::Structure::evaluate_internal
 *     src/structure_new/src/model_evaluator/4C_structure_new_model_evaluator_structure.cpp:1293-1312
 *   -> Core::FE::Discretization::evaluate(params, K, 0, f, 0, 0)
 *     src/core/fem/src/discretization/4C_fem_discretization_evaluate.cpp:31-103
 *
 * Conventions
 *   - all entry points return 0 (FC_OK) or a negative fc_status; no exceptions cross the ABI.
 *   - plain pointers + sizes only.  Pointers marked "host|device" are classified with
 *     cudaPointerGetAttributes; host buffers are staged inside the call.
 *   - one fc_ctx per rank/GPU, driven from one host thread.  All device work is enqueued on the
 *     context's stream (fc_solid_set_stream); calls with host output buffers synchronise.
 *   - there is NO CPU fallback: without a CUDA device fc_solid_create fails with
 *     FC_ERR_NO_DEVICE.  The fc_grid_* / fc_graph_* helpers are pure host integer code.
 *   - local node numbering follows the Epetra column-map convention: the n_row_nodes owned nodes
 *     come first (ascending GID), ghosts after them grouped by owner rank.  Local DOF id =
 *     3*node_lid + j; row DOF lid == column DOF lid for owned DOFs.
 */
#ifndef FOURC_B200_H
#define FOURC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C"
{
#endif

  typedef struct fc_ctx fc_ctx;

  typedef enum
  {
    FC_OK = 0,
    FC_ERR_INVALID = -1,           /* bad argument / call order */
    FC_ERR_NO_DEVICE = -2,         /* no CUDA device: there is no CPU fallback */
    FC_ERR_CUDA = -3,              /* CUDA runtime error, see fc_last_error() */
    FC_ERR_NEGATIVE_JACOBIAN = -4, /* det J <= 0 at an element node (4C_solid_3D_ele_calc_lib.hpp:492-494) */
    FC_ERR_GRAPH = -5,             /* element couples DOFs that are not in the CSR graph */
    FC_ERR_UNSUPPORTED = -6,       /* cell type / formulation / material combination not on the path */
    FC_ERR_SOLVER = -7,            /* Krylov breakdown in fc_solid_linear_solve */
    FC_ERR_SOLVER_NOT_CONVERGED = -8 /* max_iter reached; x holds the last iterate, rel_residual its residual */
  } fc_status;

  /* cell types on the path: src/solid_3D_ele/4C_solid_3D_ele_factory.hpp:36-82 */
  enum
  {
    FC_HEX8 = 0,
    FC_HEX27 = 1,
    FC_TET10 = 2,
    /* evaluated by the generic B-operator kernel (no tensor-core variant) */
    FC_HEX20 = 3,   /* 27-point rule */
    FC_TET4 = 4,    /* stiffness tet_1point, mass tet_4point */
    FC_WEDGE6 = 5,  /* wedge_6point */
    FC_PYRAMID5 = 6, /* pyramid_8point; F-bar allowed (4C_solid_3D_ele_factory.hpp:49-50) */
    FC_HEX18 = 7,    /* quad9 x line2 (18 nodes, hex_18point); faces quad9 (bottom, top) and quad6 (sides) */
    FC_NURBS27 = 8   /* 3 x 3 x 3 control points, u fastest (hex_27point); needs fc_solid_set_nurbs after fc_solid_set_mesh */
  };
  /* element technology (TECH none | fbar): src/solid_3D_ele/4C_solid_3D_ele.cpp:71-138 */
  enum
  {
    FC_FORM_DISPLACEMENT = 0, /* 4C_solid_3D_ele_calc_displacement_based.hpp:35-251 */
    FC_FORM_FBAR = 1,         /* 4C_solid_3D_ele_calc_fbar.hpp:43-191 (hex8 and pyramid5 only) */
    FC_FORM_LINEAR = 2,       /* KINEM linear: 4C_solid_3D_ele_calc_displacement_based_linear_kinematics.hpp:36-155 (every cell) */
    FC_FORM_MULF = 3,         /* PRESTRESS_TECH mulf: 4C_solid_3D_ele_calc_mulf.hpp:40-196 (every cell); Gauss-point history in the
                                 context, see fc_solid_update_prestress */
    FC_FORM_EAS_MILD = 4,     /* TECH eas_mild (hex8, 9 enhanced strain modes) and */
    FC_FORM_EAS_FULL = 5,     /* TECH eas_full (hex8, 21 modes): 4C_solid_3D_ele_calc_eas.hpp:44-291; element history in the
                                 context, see fc_solid_recover */
    FC_FORM_MULF_FBAR = 6,    /* TECH fbar with PRESTRESS_TECH mulf (hex8 and pyramid5): 4C_solid_3D_ele_calc_mulf_fbar.hpp:33-300;
                                 the history has one more record per element, for the centroid, behind the Gauss points' */
    FC_FORM_SHELL_ANS = 8,    /* TECH shell_ans (hex8, wedge6): assumed natural strains, t = thickness direction;
                                 4C_solid_3D_ele_calc_shell_ans.hpp:52-196 */
    FC_FORM_SHELL_EAS_ANS = 9, /* TECH shell_eas_ans (hex8: + the 7 modes of eas_sosh8, wedge6: + the mode of eas_sosw6);
                                 4C_solid_3D_ele_calc_shell_eas_ans.hpp:48-345; history and fc_solid_recover like FC_FORM_EAS_*,
                                 with K_da [3 nen][n] */
    FC_FORM_EAS_MILD_LINEAR = 10, /* TECH eas_mild / eas_full with KINEM linear (EASSolidIntegrator<hex8, ..., KinemType::linear>, */
    FC_FORM_EAS_FULL_LINEAR = 11, /* 4C_solid_3D_ele_factory.hpp:58-61): B-operator at F = I, strain = sym grad u + Mtilde alpha */
    FC_FORM_EAS_SHELL = 7     /* TECH shell_eas (hex8, the 7 modes of eas_sosh8 with t as thickness direction, no ANS):
                                 EASSolidIntegrator<hex8, eastype_sh8_7>, 4C_solid_3D_ele_factory.hpp:56-57; like FC_FORM_EAS_MILD */
  };
  /* materials: MAT_Struct_StVenantKirchhoff (src/mat/4C_mat_stvenantkirchhoff.cpp) and
   * MAT_ElastHyper with one ELAST_CoupNeoHooke summand (src/mat/4C_mat_elasthyper.cpp,
   * src/mat/elast/4C_mat_elast_coupneohooke.cpp).  params = {YOUNG, NUE}. */
  enum
  {
    FC_MAT_STVENANTKIRCHHOFF = 0,
    FC_MAT_ELASTHYPER_COUPNEOHOOKE = 1
  };
  /* element actions: src/solid_3D_ele/4C_solid_3D_ele_evaluate.cpp:82-134 */
  enum
  {
    FC_ACTION_NLNSTIFF = 0,     /* struct_calc_nlnstiff: f += fint, K += stiffness */
    FC_ACTION_INTERNALFORCE = 1 /* struct_calc_internalforce: f += fint */
  };
  /* evaluate flags */
  enum
  {
    FC_EVAL_ZERO_FIRST = 1, /* zero K values and f before adding (SparseMatrix::zero +
                               Vector::put_scalar(0), 4C_structure_new_model_evaluator_structure.cpp:136-159) */
    FC_EVAL_FORCE_GENERIC = 2, /* use the generic B-operator kernel even where a specialised
                                  closed-form kernel exists (testing / comparison) */
    FC_EVAL_CLOSED_FORM_FMA = 4, /* hex8 Neo-Hooke closed form on the FP64 FMA pipe instead of the
                                    FP64 tensor cores (comparison) */
    FC_EVAL_OWNER_COMPUTED = 8, /* hex8 displacement-based Neo-Hooke, struct_calc_nlnstiff on a lattice mesh: owner-computed
                             store-only kernel (k_hex8_nh_brick): no zero-fill, no atomics, every CSR row written once, results
                             bit-identical from run to run, 2.8 GB instead of 6.6 GB of DRAM traffic per evaluation at C2 -
                             but 1.9x slower than memset + per-element reductions on B200, so opt-in.
                             Ignored (per-element reductions) where the mesh has no brick plan */
    FC_EVAL_DMMA_TILES = 16, /* tet10 Neo-Hooke: the padded tensor-core tile kernel (k_nh_mma_gen) instead of the default
                                warp-per-element FMA kernel (k_tet10_nh) (comparison) */
    FC_EVAL_INTERIOR = 32, /* numproc > 1: only the elements without ghost nodes (they need no halo import: evaluate them while
                              fc_solid_halo_exchange is in flight on another stream) ... */
    FC_EVAL_BOUNDARY = 64 /* ... then only the elements with ghost nodes, after the import; without FC_EVAL_ZERO_FIRST (the
                             rows it completes are stored if the interior half ran with FC_EVAL_ZERO_FIRST on the store-only
                             kernel, added otherwise).  Tensor-core kernels only (FC_ERR_UNSUPPORTED otherwise); both or
                             neither = all elements */
  };

  /* ---- context ------------------------------------------------------------------------------ */
  int fc_solid_create(fc_ctx** ctx, int device, int rank, int nranks);
  void fc_solid_destroy(fc_ctx* ctx);
  /* cudaStream_t passed as void*; NULL = legacy default stream */
  int fc_solid_set_stream(fc_ctx* ctx, void* cuda_stream);
  const char* fc_last_error(const fc_ctx* ctx);
  const char* fc_version(void);

  /* ---- problem definition ------------------------------------------------------------------- */
  /* Replaces the per-element node/material pointers of Core::Elements::Element
   * (location_vector, 4C_fem_general_element.cpp:474-541; evaluate_element_nodes,
   * 4C_solid_3D_ele_calc_lib.hpp:152-203).
   *   X          [3*n_col_nodes] reference coordinates, node-major (host)
   *   node_gid   [n_col_nodes]   (host)
   *   node_owner [n_col_nodes]   owner rank of each column node (host)
   *   conn       [nen*n_col_ele] column-node LIDs in the reference's node order (host)
   *   ele_gid    [n_col_ele]     (host)
   * Runs the node-Jacobian check of ensure_positive_jacobian_determinant_at_element_nodes once
   * on the device (the result only depends on X); a failure is reported by every evaluate. */
  int fc_solid_set_mesh(fc_ctx* ctx, int celltype, int formulation, int n_col_nodes, const double* X,
      const int* node_gid, const int* node_owner, int n_col_ele, const int* conn, const int* ele_gid);
  /* FC_NURBS27 meshes: what Core::FE::Nurbs::get_my_nurbs_knots_and_weights gives every element evaluation
   * (4C_solid_3D_ele_calc_lib.hpp:193-200): ele_knots [n_col_ele][3][6] = the element's knot span per direction
   * (Knotvector::get_ele_knots, 4C_fem_nurbs_discretization_knotvector.cpp:173-243), node_weights [n_col_nodes] = the control
   * point weights (host arrays).  Builds the per-element shape tables once (the basis depends on the mesh only), then runs the
   * node-Jacobian check and the initial prestress history that set_mesh postpones for this cell; evaluations before this call
   * return FC_ERR_INVALID, an element with a zero-sized knot span is refused like the reference refuses it. */
  int fc_solid_set_nurbs(fc_ctx* ctx, const double* ele_knots, const double* node_weights);

  int fc_solid_set_material(fc_ctx* ctx, int kind, const double* params, int n_params);

  /* MAT_ElastHyper with a list of summands (src/mat/4C_mat_elasthyper.cpp:76-90): kinds [n_summands],
   * params [n_summands][4].  One FC_SUMMAND_COUPNEOHOOKE alone is the same material as
   * FC_MAT_ELASTHYPER_COUPNEOHOOKE (closed-form tensor-core kernels); every other list runs the generic B-operator
   * kernel.  Anisotropic summands take the element fibres of fc_solid_set_fibers. */
  enum
  {
    FC_SUMMAND_COUPNEOHOOKE = 0, /* ELAST_CoupNeoHooke  YOUNG NUE - -   (src/mat/elast/4C_mat_elast_coupneohooke.cpp) */
    FC_SUMMAND_COUPSVK = 1,      /* ELAST_CoupSVK       YOUNG NUE - -   (src/mat/elast/4C_mat_elast_coupSaintVenantKirchhoff.cpp:15-57) */
    FC_SUMMAND_COUPANISOEXPO = 2 /* ELAST_CoupAnisoExpo K1 K2 K1COMP K2COMP, INIT 1, FIBER_ID 1, STRATEGY Standard
                                    (src/mat/elast/4C_mat_elast_coupanisoexpobase.cpp:157-185) */
  };
  int fc_solid_set_material_elasthyper(fc_ctx* ctx, int n_summands, const int* kinds, const double* params);
  /* FIBER1 of every column element ([n_col_ele][3], host; the element line's FIBER1 x y z, src/solid_3D_ele/4C_solid_3D_ele.cpp:60),
   * normalised like Mat::read_anisotropy_fiber (src/mat/4C_mat_anisotropy_utils.cpp:19-44).  NULL removes them.
   * Call after fc_solid_set_mesh. */
  /* FC_FORM_MULF (modified updated Lagrangian prestressing).  The context keeps, per column element and Gauss point, the
   * history of MulfHistoryData (src/solid_3D_ele/4C_solid_3D_ele_calc_lib_mulf.hpp:24-33): 18 doubles = inverse Jacobian
   * (N_xyz = invJ dN) | deformation gradient, both column-major like the Core::LinAlg::Matrix<3, 3> members; fc_solid_set_mesh initialises it (reference Jacobian,
   * identity; FC_FORM_MULF_FBAR: [n_col_ele][ngp + 1][18], the last record of an element is its centroid's).
   * fc_solid_update_prestress is struct_update_prestress (Solid::ModelEvaluator::Structure::update_step_element,
   * 4C_structure_new_model_evaluator_structure.cpp:1555-1592 -> MulfFormulation::update_prestress) with the column
   * displacements (host|device) of the finished step; the caller resets its displacements afterwards like
   * Solid::IMPLICIT::PreStress::post_update.  get/set move the history (host|device, [n_col_ele][ngp][18]) for restart files. */
  int fc_solid_update_prestress(fc_ctx* ctx, const double* u_col);
  int fc_solid_mulf_get_history(fc_ctx* ctx, double* history);
  int fc_solid_mulf_set_history(fc_ctx* ctx, const double* history);
  /* FC_FORM_EAS_MILD / FC_FORM_EAS_FULL (hex8 with enhanced assumed strains, condensed on the element).  The context keeps the
   * history of every column element in the order EASFormulation::pack writes it (4C_solid_3D_ele_calc_eas.hpp:266-274):
   * alpha_inc [n] | alpha [n] | s [n] | Kaa^-1 [n][n] | Kda [24][n] | old_step_length, matrices column-major, n = 9 | 21 | 7, i.e.
   * 3 n + n^2 + 24 n + 1 doubles (fc_solid_eas_history_doubles).  fc_solid_set_mesh zeroes it (old_step_length = 1).
   * Every fc_solid_evaluate / fc_solid_evaluate_elements integrates Kaa, Kda and s with the current alpha, adds the condensed
   * contributions -Kda Kaa^-1 s and -Kda Kaa^-1 Kda^T and stores Kaa^-1, Kda, s (4C_solid_3D_ele_calc.cpp:110-240), also for
   * FC_ACTION_INTERNALFORCE.  fc_solid_recover is struct_calc_recover for a default step (SolidEleCalc::recover,
   * 4C_solid_3D_ele_calc.cpp:244-273; update_alpha, 4C_solid_3D_ele_calc_eas_helpers.hpp:189-228): with the Newton increment
   * du_col (column layout, host|device) alpha_inc = -Kaa^-1 (s + Kda^T du), alpha += step_length alpha_inc.  Call it after every
   * solution update, before the next evaluation.  A rank evaluates and recovers its ghost elements too, so every copy of an
   * element holds the same alpha.  get/set move the history (host|device) for restart or to repeat an evaluation. */
  int fc_solid_recover(fc_ctx* ctx, const double* du_col, double step_length);
  int fc_solid_eas_history_doubles(const fc_ctx* ctx, int* doubles_per_element);
  int fc_solid_eas_get_history(fc_ctx* ctx, double* history);
  int fc_solid_eas_set_history(fc_ctx* ctx, const double* history);
  int fc_solid_set_fibers(fc_ctx* ctx, const double* fiber1);
  /* Fibres given at the nodes (FNODE ... FIBER1; anisotropic summands with INIT 3): node_fiber1 [n_col_nodes][3].  The library
   * interpolates them to the Gauss points of the stiffness rule with the element's shape functions and normalises
   * (Core::Nodes::project_fibers_to_gauss_points, src/core/fem/src/general/node/4C_fem_general_fiber_node_utils.cpp:18-75);
   * replaces element fibres and vice versa; NULL clears. */
  int fc_solid_set_nodal_fibers(fc_ctx* ctx, const double* node_fiber1);

  /* CSR graph of the rank's rows, the state of Epetra_CrsMatrix after FillComplete on the saved
   * graph (4C_linalg_sparsematrix.cpp:353-378,843-865): rows = owned DOFs, column ids local and
   * ascending per row.  set_graph adopts a caller-provided graph (host arrays); build_graph
   * creates the full element-connectivity pattern "zeros included" (4C_linalg_sparsematrix.cpp:
   * 578-611) from the mesh.  Either one precomputes the per-element scatter slot map on the device. */
  int fc_solid_set_graph(fc_ctx* ctx, int n_row_dof, const int64_t* rowptr, const int* colind, int n_col_dof);
  int fc_solid_build_graph(fc_ctx* ctx);
  int fc_solid_graph_sizes(const fc_ctx* ctx, int* n_row_dof, int* n_col_dof, int64_t* nnz);
  /* copy the graph out (host buffers: rowptr [n_row_dof+1], colind [nnz]) */
  int fc_solid_get_graph(const fc_ctx* ctx, int64_t* rowptr, int* colind);

  /* rows skipped by the scatter, like the dbcmaps_ test in SparseMatrix::assemble
   * (4C_linalg_sparsematrix.cpp:480).  Only K rows are skipped; f is assembled for all rows. */
  int fc_solid_set_dirichlet_rows(fc_ctx* ctx, int n, const int* row_lids);

  /* ---- the hot path ---------------------------------------------------------------------------
   * One Discretization::evaluate(params, K, 0, f, 0, 0) over all column elements.
   *   u_col    [n_col_dof]  displacement in column layout (after set_state), host|device
   *   f_row    [n_row_dof]  += internal force of owned rows, host|device, may be NULL
   *   K_values [nnz]        += stiffness in CSR order (NLNSTIFF only), host|device
   *   ele_error_count   (host, may be NULL)  number of elements that failed
   *   first_bad_ele_gid (host, may be NULL)  lowest GID among them, -1 if none
   * Reading the error counters synchronises the stream; pass NULL for both to stay asynchronous
   * (device buffers only) and poll later with fc_solid_error_state. */
  int fc_solid_evaluate(fc_ctx* ctx, int action, int flags, const double* u_col, double* f_row,
      double* K_values, int* ele_error_count, int* first_bad_ele_gid);
  int fc_solid_error_state(fc_ctx* ctx, int* ele_error_count, int* first_bad_ele_gid);
  /* eval_error_check (src/structure_new/src/model_evaluator/4C_structure_new_model_evaluator_generic.cpp:225-258): the
   * reference sums one int error flag over all ranks (Core::Communication::sum_all, :253-257) before it lets NOX go on.
   * any_error (host) = 1 if any rank's last evaluation had a failing element or its mesh failed the node-Jacobian
   * check: one 4-byte ncclAllReduce(max) on the communicator of fc_solid_set_comm.  Synchronises the stream; every
   * rank calls.  One rank: the local flag, no communicator needed. */
  int fc_solid_error_allreduce(fc_ctx* ctx, int* any_error);

  /* SparseMatrix::zero / Vector::put_scalar(0) on device buffers (4C_linalg_sparsematrix.cpp:353-378) */
  int fc_solid_zero(fc_ctx* ctx, double* f_row_dev, double* K_values_dev);

  /* Consistent mass matrix (struct_calc_nlnstiffmass branch, 4C_solid_3D_ele_evaluate.cpp:107-119; add_mass_matrix,
   * 4C_solid_3D_ele_calc_lib.hpp:940-958; tet10 integrates it with tet_11point, 4C_solid_3D_ele_calc.cpp:228-240)
   * added into M_values [nnz] (host|device) in the layout of the stiffness graph, owned rows only.  flags:
   * FC_EVAL_ZERO_FIRST.  Constant for a mesh: call once, not per Newton iteration. */
  int fc_solid_evaluate_mass(fc_ctx* ctx, double density, int flags, double* M_values);

  /* Neumann loads of TYPE Live: the external force that
   * Solid::ModelEvaluator::Structure::apply_force_external assembles through Discretization::evaluate_neumann
   * (src/structure_new/src/model_evaluator/4C_structure_new_model_evaluator_structure.cpp:368-396), added into f_ext
   * [n_row_dof] (host|device, += like the reference's assemble; owned rows only, so pass every face / element that
   * touches an owned node).
   *   surface  neum_live of src/solid_3D_ele/4C_solid_3D_ele_surface_evaluate.cpp:225-330 on faces given by their nodes
   *            (face_nodes [n_faces][4|9|6] column-node LIDs in the reference's quad4 / quad9 / tri6 node order), rules
   *            quad_4point / quad_9point / tri_6point with the reference's 13-digit constants
   *            (4C_solid_3D_ele_surface.cpp:58-77, 4C_fem_general_utils_integration.cpp:6151-6231, 6374-6396)
   *   volume   evaluate_neumann of src/solid_3D_ele/4C_solid_3D_ele_neumann_evaluator.cpp:43-115 on column elements
   *            ele_lids [n_ele], stiffness Gauss rule of the cell
   *   onoff [3], val [3]  ONOFF and VAL of the condition's first three DOFs
   *   gp_factor           [n][ngp][3] (host|device) value of the condition's FUNCT per Gauss point and DOF
   *                       (FunctionOfSpaceTime::evaluate(X_gp, t, dof)); NULL = 1.  The library does not parse
   *                       functions: the caller evaluates them at the reference coordinates of the Gauss points,
   *                       which the *_gp_coords calls return ([n][ngp][3], host|device). */
  enum
  {
    FC_FACE_QUAD4 = 0,
    FC_FACE_QUAD9 = 1,
    FC_FACE_TRI6 = 2,
    FC_FACE_QUAD8 = 3, /* hex20 faces, quad_9point */
    FC_FACE_TRI3 = 4,  /* tet4 / wedge6 / pyramid5 faces, tri_3point */
    FC_FACE_QUAD6 = 5  /* side faces of hex18 (quadratic in r, linear in s), quad_6point */
  };
  int fc_solid_neumann_surface(fc_ctx* ctx, int face_type, int n_faces, const int* face_nodes, const int* onoff, const double* val,
      const double* gp_factor, double* f_ext);
  int fc_solid_neumann_volume(fc_ctx* ctx, int n_ele, const int* ele_lids, const int* onoff, const double* val, const double* gp_factor,
      double* f_ext);
  int fc_solid_neumann_surface_gp_coords(fc_ctx* ctx, int face_type, int n_faces, const int* face_nodes, double* gp_coords);
  int fc_solid_neumann_volume_gp_coords(fc_ctx* ctx, int n_ele, const int* ele_lids, double* gp_coords);
  /* TYPE pseudo_orthopressure / orthopressure (4C_solid_3D_ele_surface_evaluate.cpp:84-96, 139-176, 318-358): the pressure VAL[0]
   * along the normal of the configuration X + u, f[a][dim] += N_a normal[dim] w functfac VAL[0] with the area-weighted normal
   * of surface_integration (:382-394).  u_col [n_col_dof] (host|device): the last converged displacements for
   * pseudo_orthopressure, the current iterate for orthopressure, NULL = material configuration (what the reference uses while
   * MULF prestressing is active).  gp_factor as above, its first component is used.
   * _linearization: the load stiffness of LOADLIN yes (:358-376 with analytical_d_surface_integration :502-549),
   * K[(a, dim)][(b, k)] -= N_a d normal[dim] / d x_(b,k) w functfac VAL[0], added into K_values [nnz] (host|device) in the layout of the
   * stiffness graph: owned rows, Dirichlet rows skipped like the element scatter.  The reference adds it to the tangent with the
   * same sign (the external force enters the residual with a minus). */
  int fc_solid_neumann_surface_orthopressure(fc_ctx* ctx, int face_type, int n_faces, const int* face_nodes, const double* u_col,
      double pressure, const double* gp_factor, double* f_ext);
  int fc_solid_neumann_surface_orthopressure_linearization(fc_ctx* ctx, int face_type, int n_faces, const int* face_nodes, const double* u_col,
      double pressure, const double* gp_factor, double* K_values);
  /* Surface loads on faces of nurbs27 elements (nurbs9, quad_9point): a face is named like the reference's surface element
   * names it - by its parent column element and the parent's surface number 0..5 (eleNodeNumbering_nurbs27_surfaces,
   * 4C_fem_general_utils_local_connectivity_matrices.hpp:250-253) - because its two knot vectors come from the parent's
   * knot spans (Knotvector::get_boundary_ele_and_parent_knots, 4C_fem_nurbs_discretization_knotvector.cpp:251-420;
   * 4C_solid_3D_ele_surface_evaluate.cpp:190-256).  parent_ele [n_faces], surface_id [n_faces]; everything else as above. */
  int fc_solid_neumann_surface_nurbs(fc_ctx* ctx, int n_faces, const int* parent_ele, const int* surface_id, const int* onoff,
      const double* val, const double* gp_factor, double* f_ext);
  int fc_solid_neumann_surface_nurbs_gp_coords(fc_ctx* ctx, int n_faces, const int* parent_ele, const int* surface_id, double* gp_coords);

  /* struct_calc_stress post-processing: Gauss-point stresses (calculate_stress,
   * 4C_solid_3D_ele_calc.cpp:347-390; Cauchy = F S F^T / det F, 4C_solid_3D_ele_utils.cpp:20-33) and their extrapolation
   * to the nodes with nodal averaging (4C_fem_general_utils_gauss_point_extrapolation.cpp:260-320, 418-427).
   * gp_stress [n_col_ele][ngp][6], nodal_stress [n_row_nodes][6] (host|device, either may be NULL); stress-Voigt
   * xx,yy,zz,xy,yz,zx.  With several ranks the nodal values of owned nodes are complete (every element at an owned
   * node is a column element). */
  enum fc_stress_type
  {
    FC_STRESS_2PK = 0,
    FC_STRESS_CAUCHY = 1
  };
  int fc_solid_evaluate_stress(fc_ctx* ctx, const double* u_col, int stress_type, double* gp_stress, double* nodal_stress);
  /* The strains the same action writes next to the stresses (STRUCT_STRAIN gl | ea | log; assemble_strain_type_to_matrix_row,
   * 4C_solid_3D_ele_calc_lib_io.hpp:113-153): Green-Lagrange E, Euler-Almansi F^-T E F^-1 or logarithmic strains
   * (4C_solid_3D_ele_utils.cpp:35-84) of the formulation's strain measure (F-bar: of the modified F; EAS: the enhanced strains;
   * KINEM linear: sym grad u), in the stress-like Voigt order of the output xx,yy,zz,xy,yz,zx (shear components not doubled).
   * Same layouts and extrapolation as fc_solid_evaluate_stress. */
  enum fc_strain_type
  {
    FC_STRAIN_GL = 0,
    FC_STRAIN_EA = 1,
    FC_STRAIN_LOG = 2
  };
  int fc_solid_evaluate_strain(fc_ctx* ctx, const double* u_col, int strain_type, double* gp_strain, double* nodal_strain);

  /* Nodal stresses as fc_solid_evaluate_stress computes them (Gauss-point stresses extrapolated to the nodes and averaged),
   * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
   * NO COUNTERPART IN THE REFERENCE.
   * The call allocates nothing (no cudaMalloc / cudaFree / pinned or library allocation) and does not synchronise: its
   * scratch (Gauss-point stresses, extrapolation matrix, elements-per-node counts) is one fixed device array of about
   * 1.7 MB inside the library, held by every GPU that loads it until exit.  Limits: n_col_ele * ngp <=
   * FC_NOALLOC_MAX_GAUSS_POINTS and n_col_nodes <= FC_NOALLOC_MAX_COL_NODES; beyond them the call returns FC_ERR_INVALID
   * with a message naming the limit (use fc_solid_evaluate_stress then).  u_col_dev [n_col_dof] and nodal_stress_dev
   * [n_row_nodes][6] must be DEVICE buffers (a host buffer would need staging, i.e. an allocation).  All calls on one GPU
   * share the array, which keeps the previous call's contents (the call clears what must start at zero): calls that
   * overlap on the GPU (different streams or contexts of one device) race; serialise them. */
#define FC_NOALLOC_MAX_GAUSS_POINTS 32768
#define FC_NOALLOC_MAX_COL_NODES 32768
  int fc_solid_evaluate_stress_noalloc(fc_ctx* ctx, const double* u_col_dev, int stress_type, double* nodal_stress_dev);

  /* Element-local evaluation without scatter (Solid::evaluate seam, 4C_solid_3D_ele.hpp:120-124):
   * fe [n_col_ele][3*nen], Ke [n_col_ele][3*nen][3*nen] column-major per element, in the caller's
   * element order.  host|device.  Ke may be NULL. */
  int fc_solid_evaluate_elements(fc_ctx* ctx, const double* u_col, double* fe, double* Ke);
  /* FC_EVAL_* kernel-selection flags used by fc_solid_evaluate_elements (default 0) */
  int fc_solid_set_element_kernel_flags(fc_ctx* ctx, int flags);

  /* ---- halo plan for Discretization::set_state (4C_fem_discretization.cpp:503-558) -----------
   * The forward import row layout -> column layout.  Ghost DOFs sit behind the owned ones grouped
   * by owner rank, so the receive side needs no unpack: rank p's message lands at
   * u_col + recv_offset[p].  The send side is a gather of owned DOFs (fc_solid_halo_pack).
   * The exchange itself (NCCL grouped send/recv) is issued by the caller between pack and evaluate. */
  int fc_solid_halo_set_send(fc_ctx* ctx, int n_send, const int* send_row_lids);
  int fc_solid_halo_pack(fc_ctx* ctx, const double* u_dev, double* sendbuf_dev);
  /* The whole forward import inside the library (what Epetra_Import does in set_state, 4C_fem_discretization.cpp:546-551):
   * send_dof_count[p] DOFs of the send list go to rank p (the list is ordered by peer), recv_dof_count[p] ghost DOFs
   * arrive from rank p.  fc_solid_halo_exchange packs, then issues one grouped ncclSend/ncclRecv per peer on the
   * context's stream; rank p's message lands at u_col_dev + n_row_dof + sum_{r<p} recv_dof_count[r].
   * nccl_comm is the caller's ncclComm_t whose rank ids equal the ranks given to fc_solid_create.  The library
   * does not link NCCL: the symbols are resolved from the process (or libnccl.so.2) at the first call. */
  int fc_solid_halo_set_counts(fc_ctx* ctx, const int* send_dof_count, const int* recv_dof_count);
  int fc_solid_halo_exchange(fc_ctx* ctx, void* nccl_comm, double* u_col_dev);
  /* communicator (ncclComm_t, rank ids = the ranks of fc_solid_create) used by fc_solid_linear_solve on numproc > 1: halo
   * import of the search direction before every y = K x, one all-reduce per dot product */
  int fc_solid_set_comm(fc_ctx* ctx, void* nccl_comm);

  /* ---- timing helpers (device time of the last evaluate, CUDA events on the ctx stream) ----- */
  int fc_solid_last_kernel_ms(fc_ctx* ctx, float* zero_ms, float* element_ms);
  /* the same for the last n <= 128 evaluate / spmv calls, oldest first (zero_ms, element_ms: host arrays of n, either may
   * be NULL): the device times of a whole timed region, read after it.  Synchronises on the last call only. */
  int fc_solid_kernel_ms_history(fc_ctx* ctx, int n, float* zero_ms, float* element_ms);
  int fc_solid_kernel_launch_count(const fc_ctx* ctx, int64_t* launches);

  /* FP64 peak probes (TFLOP/s, best of 5) for the roofline denominators: DFMA loop and
   * mma.sync.m8n8k4.f64 loop on all SMs */
  int fc_probe_fp64_fma(int device, double* tflops);
  int fc_probe_fp64_mma(int device, double* tflops);

  /* ---- structured cuboid generator + reference decomposition (host, integer arithmetic) ------
   * Core::IO::GridGenerator::create_rectangular_cuboid_discretization with PARTITION structured
   * (src/core/io/src/4C_io_gridgenerator.cpp:41-323,328-393), unique node ownership
   * (src/core/rebalance/src/4C_rebalance_graph_based.cpp:159-214), element owner / ghosting rule
   * (src/core/fem/src/discretization/4C_fem_discretization_partition.cpp:510-587) and DOF numbering
   * (src/core/fem/src/dofset/4C_fem_dofset.cpp:344-352), evaluated for ONE rank in closed form. */
  typedef struct fc_grid fc_grid;
  int fc_grid_create(fc_grid** g, int celltype, int nx, int ny, int nz, const double* lo,
      const double* hi, int nranks, int rank, int node_offset);
  void fc_grid_destroy(fc_grid* g);
  /* sizes[0..5] = n_row_nodes, n_col_nodes, n_row_eles, n_col_eles, min_node_gid (global), nen */
  int fc_grid_sizes(const fc_grid* g, int64_t* sizes);
  int fc_grid_subdivisions(const fc_grid* g, int* sub3);
  /* column nodes (owned first, ascending GID; ghosts grouped by owner, ascending GID) */
  int fc_grid_get_nodes(const fc_grid* g, int* node_gid, int* node_owner, double* X);
  /* column elements (row elements first); conn = column-node LIDs */
  int fc_grid_get_elements(const fc_grid* g, int* ele_gid, int* ele_owner, int* conn);
  /* halo plan: counts per peer, then the lists.  send lists are row-node LIDs in ascending GID
   * per peer; recv lists are implicit (ghost nodes grouped by owner). */
  int fc_grid_halo_counts(const fc_grid* g, int* send_node_count /*[nranks]*/, int* recv_node_count /*[nranks]*/);
  int fc_grid_halo_send_nodes(const fc_grid* g, int* send_node_lids /*[sum send counts]*/);

  /* ---- Dirichlet application and linear-solve hand-off, K resident in HBM ---------------------
   * Stand where Solid::Dbc::apply_dirichlet_to_local_jacobian / apply_dirichlet_to_rhs
   * (src/structure_new/src/utils/4C_structure_new_dbc.cpp:244-290) and the Newton linear system hand K and the
   * residual to core/linear_solver.  All vectors and K_values are DEVICE buffers.
   *   set_dirichlet_toggle   the dbcmap (owned DOF row lids, host array); independent of set_dirichlet_rows, which
   *                          is the dbcmaps_ of the matrix that makes the scatter skip rows
   *   apply_dirichlet        SparseMatrix::apply_dirichlet (4C_linalg_sparsematrix.cpp:978-1100, non-explicit
   *                          branch :1066-1092): zero the rows, 1.0 on the diagonal if diagonalblock
   *   apply_dirichlet_to_system  x[i] = b[i] = dbcval[i] on the toggled rows
   *                          (4C_linalg_utils_sparse_algebra_assemble.cpp:126-140); x or b may be NULL, dbcval NULL = zeros
   *   spmv                   y_row = K x_col on the rank's rows
   *   linear_solve           K x = b, preconditioned (node-block Jacobi) CG or BiCGStab; x holds the initial guess on
   *                          entry.  CG needs a tangent that is symmetric positive definite on the free DOFs and x = b
   *                          on the Dirichlet rows on entry (apply_dirichlet_to_system sets that); BiCGStab has no such
   *                          requirement (F-bar tangents are not symmetric).  numproc > 1: b and x are the rank's rows, the halo plan
   *                          (fc_solid_halo_set_send/_set_counts) and fc_solid_set_comm must be in place; every rank calls. */
  enum
  {
    FC_SOLVER_CG = 0,
    FC_SOLVER_BICGSTAB = 1,
    FC_SOLVER_CG_MG = 2 /* CG preconditioned by a geometric multigrid V-cycle (fc_solid_mg_setup / _update) */
  };
  /* Geometric multigrid hierarchy for FC_SOLVER_CG_MG, in place of the MueLu preconditioner the reference configures
   * for large structural problems (src/core/linear_solver).  hex8 lattice mesh on one rank; isotropic materials.
   *   mg_setup   once per mesh + graph: coarse lattices (every second node, plus the last node of a direction with an odd
   *              number of elements) down to 2 elements per direction or max_levels levels; n_levels (host) = levels
   *              built, FC_ERR_UNSUPPORTED when there is no coarser one
   *   mg_update  once per Newton iteration, after fc_solid_set_dirichlet_toggle: coarse tangents = the same elements and
   *              material re-discretised on the coarse lattices at the injected state u_col_dev (device), coarse
   *              Dirichlet rows = the toggled DOFs of the coincident nodes, node-block Jacobi smoother blocks
   * The V(2,2) cycle uses damped node-block Jacobi (omega = 4 / (3 lambda_max(D^-1 K)) per level, lambda_max from a few
   * power iterations in mg_update), trilinear transfers, 60 sweeps on the coarsest level. */
  int fc_solid_mg_setup(fc_ctx* ctx, int max_levels, int* n_levels);
  int fc_solid_mg_update(fc_ctx* ctx, const double* u_col_dev);
  int fc_solid_mg_set_sweeps(fc_ctx* ctx, int nu); /* pre- and post-smoothing sweeps of the cycle, default 2 */
  /* The V-cycle reads single-precision copies of the level tangents (default; vectors, sums and the CG around it stay FP64:
   * the cycle is then the exact V-cycle of a matrix that differs from K by 6e-8 relative, still a symmetric positive
   * definite preconditioner) or the FP64 values (0).  Call fc_solid_mg_update afterwards. */
  int fc_solid_mg_set_precision(fc_ctx* ctx, int single_precision_cycle);
  /* Jacobi damping of every level: omega = factor / lambda_max(D^-1 K), 0 < factor < 2 (default 4/3).  lambda_max is a
   * power-iteration estimate, i.e. a lower bound: should a level's damping come out too large (the cycle is then indefinite and
   * CG sees r.z <= 0), the solver multiplies every level's damping by 0.7 and restarts from its current iterate, up to four
   * times, before it gives up with FC_ERR_SOLVER.  Call fc_solid_mg_update afterwards. */
  int fc_solid_mg_set_damping(fc_ctx* ctx, double factor);
  /* nodes and Jacobi damping of one level (0 = finest; its damping is known after the first solve that follows mg_update) */
  int fc_solid_mg_level_info(fc_ctx* ctx, int level, int* n_nodes, double* omega);
  int fc_solid_set_dirichlet_toggle(fc_ctx* ctx, int n, const int* row_lids);
  int fc_solid_apply_dirichlet(fc_ctx* ctx, double* K_values, int diagonalblock);
  int fc_solid_apply_dirichlet_to_system(fc_ctx* ctx, double* x, double* b, const double* dbcval);
  int fc_solid_spmv(fc_ctx* ctx, const double* K_values, const double* x_col, double* y_row);
  int fc_solid_linear_solve(fc_ctx* ctx, int method, const double* K_values, const double* b, double* x, double rel_tol, int max_iter,
      int* iters, double* rel_residual);

  /* ---- brick plan of the owner-computed hex8 assembly (host integer code; exposed for inspection and tests) ------
   * Replaces the per-element row loop of SparseMatrix::assemble (4C_linalg_sparsematrix.cpp:471-575) and the zero-fill
   * of SparseMatrix::zero (:353-378): the owned nodes of a lattice mesh are cut into units of 7 x 5 node lines x
   * `chunk` node planes; the CTA that owns a unit evaluates every element touching it, sums in shared memory and
   * stores each CSR row once.  The first evaluation that can use it creates the context's plan; these entry points
   * build the same plan from plain arrays.
   *   unit_planes [n_units], unit_node_off [n_units], unit_ele_off [n_units]
   *   node_tab  [sum planes * 35]        node of (plane, line) or -1
   *   ele_tab   [sum (planes + 1) * 48]  element of (layer, slot) or -1; slot = 12 * colour + warp, the 12 elements of a
   *                                      colour share no node
   *   rowidx_tab  same shape             byte a = accumulator row of local node a = (plane & 1) * 35 + line, 0xFF = the
   *                                      node is not a row of this unit
   * units [0, n_interior_units) contain no element with a ghost node.  max_row_nodes = largest number of nodes a row
   * couples to (27 on a lattice); chunk = 0 lets the library choose.  create returns FC_ERR_GRAPH when the mesh is
   * not a lattice of hex cells; the library then scatters element by element. */
  typedef struct fc_brick_plan fc_brick_plan;
  int fc_brick_plan_create(int n_col_nodes, const double* X, int n_ele, const int* conn, int n_row_nodes, int max_row_nodes, int chunk,
      fc_brick_plan** out);
  int fc_brick_plan_sizes(const fc_brick_plan* p, int* n_units, int* n_interior_units, int* chunk, int64_t* n_node_entries,
      int64_t* n_ele_entries, int64_t* n_element_instances);
  int fc_brick_plan_get(const fc_brick_plan* p, int* unit_planes, int64_t* unit_node_off, int64_t* unit_ele_off, int* node_tab, int* ele_tab,
      uint64_t* rowidx_tab);
  void fc_brick_plan_destroy(fc_brick_plan* p);
  /* element order the hex8 assembled-path kernels visit on large meshes (host integer code): y blocks of z layers of at
   * most layer_elems elements, so that a matrix row's element contributions arrive while the row is still in L2.
   * perm [n_ele] = element visited by launch slot k; returns 1 if the order differs from the caller's, 0 if it is kept. */
  int fc_l2_sweep_order(int n_col_nodes, const double* X, int n_ele, int nen, const int* conn, int64_t layer_elems, int* perm);

#ifdef __cplusplus
}
#endif
#endif /* FOURC_B200_H */
