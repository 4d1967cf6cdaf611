// ORACLE — TEST INFRASTRUCTURE ONLY. Never linked into or called by the product library.
// C ABI over the CPU restatement so that tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs can drive it through ctypes.
#include <chrono>
#include <cstring>

#include "fc_oracle_assemble.hpp"
#include "fc_oracle_neumann.hpp"

using namespace fc_oracle;

namespace
{
  struct JobState
  {
    Job job;
    std::vector<std::vector<double>> u_row, f_row, K_values;
  };
  void finish_job(JobState* js)
  {
    const int nr = js->job.part.nranks;
    js->job.sys.resize(nr);
    for (int r = 0; r < nr; ++r) build_rank_system(js->job.mesh, js->job.part, r, js->job.sys[r]);
    js->u_row.resize(nr);
    js->f_row.resize(nr);
    js->K_values.resize(nr);
    for (int r = 0; r < nr; ++r)
    {
      js->u_row[r].assign(js->job.sys[r].row_dof_gid.size(), 0.0);
      js->f_row[r].assign(js->job.sys[r].row_dof_gid.size(), 0.0);
    }
  }
  template <class T>
  void copy_out(const std::vector<T>& v, T* out)
  {
    if (out && !v.empty()) std::memcpy(out, v.data(), v.size() * sizeof(T));
  }
}  // namespace

extern "C"
{
  // ---- unit-level entry points (golden vector checks) -----------------------------------------
  void oracle_calculate_gamma_delta(
      const double* prinv, const double* dPI, const double* ddPII, double* gamma, double* delta)
  {
    calculate_gamma_delta(gamma, delta, prinv, dPI, ddPII);
  }
  void oracle_material_evaluate(int kind, double young, double nue, const double* F,
      const double* glstrain, double* S, double* cmat)
  {
    MaterialParams mp{kind, young, nue};
    material_evaluate(mp, F, glstrain, S, cmat);
  }
  double oracle_svk_strain_energy(double young, double nue, const double* glstrain)
  {
    return svk_strain_energy(young, nue, glstrain);
  }
  void oracle_rcg_strain_like(const double* E, double* C)
  {
    for (int i = 0; i < 6; ++i) C[i] = 2.0 * E[i];
    for (int i = 0; i < 3; ++i) C[i] += 1.0;
  }
  double oracle_voigt_determinant(int strain, const double* t)
  {
    return strain ? voigt_determinant_strain(t) : voigt_determinant_stress(t);
  }
  void oracle_voigt_invariants(int strain, const double* t, double* prinv)
  {
    if (strain)
      voigt_invariants_principal_strain(prinv, t);
    else
      voigt_invariants_principal_stress(prinv, t);
  }
  void oracle_voigt_inverse(int strain, const double* t, double* ti)
  {
    if (strain)
      voigt_inverse_tensor<true>(t, ti);
    else
      voigt_inverse_tensor<false>(t, ti);
  }
  void oracle_voigt_to_stress_like(const double* in, double* out) { voigt_strain_to_stress_like(in, out); }
  void oracle_voigt_matrix_to_vector_strain(const double* m, double* v) { voigt_matrix_to_vector_strain(m, v); }
  void oracle_add_holzapfel_product(double* cmat, const double* invc, double scalar)
  {
    add_holzapfel_product(cmat, invc, scalar);
  }
  double oracle_invert3x3(const double* in, double* out) { return invert3x3(out, in, nullptr); }
  double oracle_determinant3x3(const double* in) { return determinant3x3(in); }
  // generic small products for the bit-exact check against the reference header (oracle/_ref)
  void oracle_multiply_nn_6_6_24(double* out, const double* l, const double* r) { multiply_nn<6, 6, 24>(out, l, r); }
  void oracle_multiply_tn_24_6_24(double fac, double* out, const double* l, const double* r)
  {
    multiply_tn_scaled<24, 6, 24>(fac, out, l, r, 1.0);
  }
  void oracle_multiply_nt_3_8_3(double* out, const double* l, const double* r) { multiply_nt<3, 8, 3>(out, l, r); }
  void oracle_multiply_tn_3_3_3(double* out, const double* l, const double* r) { multiply_tn<3, 3, 3>(out, l, r); }

  int oracle_num_nodes(int ct) { return num_nodes(ct); }
  int oracle_num_gp(int ct) { return num_gp(ct); }
  void oracle_shape(int ct, const double* xi, double* N, double* dN)
  {
    if (N) shape_function(ct, xi, N);
    if (dN) shape_function_deriv1(ct, xi, dN);
  }
  void oracle_gauss_rule(int ct, double* xi, double* w)
  {
    double x[27][3];
    gauss_rule(ct, x, w);
    for (int g = 0; g < num_gp(ct); ++g)
      for (int d = 0; d < 3; ++d) xi[3 * g + d] = x[g][d];
  }
  void oracle_node_param_coords(int ct, double* xi)
  {
    for (int a = 0; a < num_nodes(ct); ++a) node_param_coords(ct, a, xi + 3 * a);
  }
  int oracle_element_evaluate(int ct, int form, int kind, double young, double nue, const double* X,
      const double* u, double* fe, double* Ke, double* Me, double rho)
  {
    MaterialParams mp{kind, young, nue};
    return element_evaluate(ct, form, mp, X, u, fe, Ke, Me, rho);
  }
  // Gauss-point stresses of one element ([ngp][6], stress-Voigt xx,yy,zz,xy,yz,zx); stress_type 0 = 2PK, 1 = Cauchy
  int oracle_element_gp_stress(int ct, int form, int kind, double young, double nue, const double* X, const double* u,
      int stress_type, double* gp_stress)
  {
    MaterialParams mp{kind, young, nue};
    return element_evaluate(ct, form, mp, X, u, nullptr, nullptr, nullptr, 0.0, gp_stress, stress_type);
  }
  // Neumann loads TYPE Live on n independent faces / elements (fc_oracle_neumann.hpp): x [n][nen][3] reference coordinates,
  // functfac [n][ngp][3] or NULL, fe [n][nen][3] += (may be NULL), gp_coord [n][ngp][3] (may be NULL)
  void oracle_surface_neumann(int ft, int n, const double* x, const int* onoff, const double* val, const double* functfac, double* fe,
      double* gp_coord)
  {
    const int nn = face_num_nodes(ft), ng = face_num_gp(ft);
    for (int k = 0; k < n; ++k)
      surface_neumann_live(ft, x + (size_t)k * nn * 3, onoff, val, functfac ? functfac + (size_t)k * ng * 3 : nullptr,
          fe ? fe + (size_t)k * nn * 3 : nullptr, gp_coord ? gp_coord + (size_t)k * ng * 3 : nullptr);
  }
  void oracle_surface_neumann_orthopressure(int ft, int n, const double* xc, double value, const double* functfac, double* fe)
  {
    const int nn = face_num_nodes(ft), ng = face_num_gp(ft);
    for (int k = 0; k < n; ++k)
      surface_neumann_orthopressure(ft, xc + (size_t)k * nn * 3, value, functfac ? functfac + (size_t)k * ng * 3 : nullptr, fe + (size_t)k * nn * 3);
  }
  void oracle_surface_neumann_orthopressure_linearization(int ft, int n, const double* xc, double value, const double* functfac, double* Ke)
  {
    const int nn = face_num_nodes(ft), ng = face_num_gp(ft);
    for (int k = 0; k < n; ++k)
      surface_neumann_orthopressure_linearization(ft, xc + (size_t)k * nn * 3, value, functfac ? functfac + (size_t)k * ng * 3 : nullptr,
          Ke + (size_t)k * 9 * nn * nn);
  }
  void oracle_volume_neumann(int ct, int n, const double* X, const int* onoff, const double* val, const double* functfac, double* fe,
      double* gp_coord)
  {
    const int nn = num_nodes(ct), ng = num_gp(ct);
    for (int k = 0; k < n; ++k)
      volume_neumann_live(ct, X + (size_t)k * nn * 3, onoff, val, functfac ? functfac + (size_t)k * ng * 3 : nullptr,
          fe ? fe + (size_t)k * nn * 3 : nullptr, gp_coord ? gp_coord + (size_t)k * ng * 3 : nullptr);
  }
  // batch of independent elements (X,u: [n][3*nen]); used as the checker for the CUDA
  // element-level entry point.  returns the number of failing elements.
  int oracle_element_evaluate_batch(int ct, int form, int kind, double young, double nue, int n,
      const double* X, const double* u, double* fe, double* Ke)
  {
    MaterialParams mp{kind, young, nue};
    const int nd = 3 * num_nodes(ct);
    int bad = 0;
    for (int e = 0; e < n; ++e)
    {
      std::memset(fe + (size_t)e * nd, 0, sizeof(double) * nd);
      if (Ke) std::memset(Ke + (size_t)e * nd * nd, 0, sizeof(double) * nd * nd);
      bad += 0 != element_evaluate(ct, form, mp, X + (size_t)e * nd, u + (size_t)e * nd,
                      fe + (size_t)e * nd, Ke ? Ke + (size_t)e * nd * nd : nullptr);
    }
    return bad;
  }

  // ---- job-level entry points -------------------------------------------------------------------
  void* oracle_job_create_cuboid(int ct, int nx, int ny, int nz, const double* lo, const double* hi,
      int nranks, int node_offset)
  {
    auto* js = new JobState;
    const int interval[3] = {nx, ny, nz};
    create_cuboid(ct, interval, lo, hi, nranks, node_offset, js->job.mesh, js->job.part);
    finish_job(js);
    return js;
  }
  void* oracle_job_create_mesh(int ct, int n_nodes, const int* node_gid, const double* X, int n_ele,
      const int* ele_gid, const int* conn)
  {
    auto* js = new JobState;
    Mesh& m = js->job.mesh;
    m.ct = ct;
    m.nen = num_nodes(ct);
    std::vector<int> order(n_nodes);
    for (int i = 0; i < n_nodes; ++i) order[i] = i;
    std::sort(order.begin(), order.end(), [&](int a, int b) { return node_gid[a] < node_gid[b]; });
    m.node_gid.resize(n_nodes);
    m.node_X.resize(3 * (size_t)n_nodes);
    for (int i = 0; i < n_nodes; ++i)
    {
      m.node_gid[i] = node_gid[order[i]];
      for (int d = 0; d < 3; ++d) m.node_X[3 * (size_t)i + d] = X[3 * (size_t)order[i] + d];
    }
    m.ele_gid.assign(ele_gid, ele_gid + n_ele);
    m.ele_nodes.assign(conn, conn + (size_t)n_ele * m.nen);
    m.finalize();
    trivial_partition(m, js->job.part);
    finish_job(js);
    return js;
  }
  void oracle_job_destroy(void* h) { delete (JobState*)h; }
  // nurbs27 meshes: knots [n_ele][3][6] in the mesh's element order, weights [n_nodes] parallel to node_gid [n_nodes]
  void oracle_job_set_nurbs(void* h, const double* knots, int n_nodes, const int* node_gid, const double* weights)
  {
    auto* js = (JobState*)h;
    Mesh& m = js->job.mesh;
    m.nurbs_knots.assign(knots, knots + 18 * m.ele_gid.size());
    m.nurbs_weights.assign(m.node_gid.size(), 1.0);
    for (int i = 0; i < n_nodes; ++i) m.nurbs_weights[(size_t)m.node_index.at(node_gid[i])] = weights[i];
  }
  // unit-level calls on one nurbs27 element / nurbs9 face (oracle_shape, oracle_element_evaluate, oracle_neumann_*): the knots
  // ([3][6] or [2][6]) and weights ([27] or [9]) that the following calls of this thread use
  void oracle_set_nurbs_element(const double* knots, int n_knots, const double* weights, int n_weights)
  {
    static thread_local double kn[18], w[27];
    std::copy(knots, knots + n_knots, kn);
    std::copy(weights, weights + n_weights, w);
    nurbs_context().knots = kn;
    nurbs_context().weights = w;
  }
  // FORM_MULF: history of every element as MulfFormulation holds it before the first prestress update
  static void mulf_init(JobState* js)
  {
    js->job.mulf_hist.clear();
    js->job.eas_hist.clear();
    if (form_is_eas(js->job.form))
    {
      // EASHistoryData of a new element: everything zero, old_step_length = 1
      const int neas = eas_modes(js->job.form), nh = eas_history_doubles(neas);
      js->job.eas_hist.assign((size_t)nh * js->job.mesh.ele_gid.size(), 0.0);
      for (size_t e = 0; e < js->job.mesh.ele_gid.size(); ++e) eas_history_init(neas, &js->job.eas_hist[(size_t)nh * e]);
    }
    if (js->job.form == FORM_SHELL_EAS_ANS)
    {
      const int nd = 3 * js->job.mesh.nen, neas = shell_eas_modes(js->job.mesh.ct, js->job.form), nh = shell_history_doubles(neas, nd);
      js->job.eas_hist.assign((size_t)nh * js->job.mesh.ele_gid.size(), 0.0);
      for (size_t e = 0; e < js->job.mesh.ele_gid.size(); ++e) js->job.eas_hist[(size_t)nh * (e + 1) - 1] = 1.0;
    }
    const Mesh& m = js->job.mesh;
    const int ngp = mulf_records(m.ct, js->job.form);  // records per element (MULF with F-bar: + the centroid)
    if (ngp == 0) return;
    js->job.mulf_hist.assign((size_t)18 * ngp * m.ele_gid.size(), 0.0);
    std::vector<double> X((size_t)3 * m.nen);
    for (size_t e = 0; e < m.ele_gid.size(); ++e)
    {
      for (int a = 0; a < m.nen; ++a)
        for (int i = 0; i < 3; ++i) X[3 * a + i] = m.node_X[3 * (size_t)m.node_index.at(m.ele_nodes[e * m.nen + a]) + i];
      const NurbsScope nurbs(m, e);
      mulf_history(m.ct, X.data(), nullptr, &js->job.mulf_hist[(size_t)18 * ngp * e], js->job.form == FORM_MULF_FBAR);
    }
  }
  void oracle_job_set_problem(void* h, int form, int kind, double young, double nue)
  {
    auto* js = (JobState*)h;
    js->job.form = form;
    js->job.mp = MaterialParams{kind, young, nue};
    mulf_init(js);
  }
  // struct_update_prestress with the displacements of the finished step, u_nodes [n_nodes][3] in the mesh's node order
  void oracle_job_update_prestress(void* h, const double* u_nodes)
  {
    auto* js = (JobState*)h;
    const Mesh& m = js->job.mesh;
    if (js->job.mulf_hist.empty()) return;
    const int ngp = mulf_records(m.ct, js->job.form);
    std::vector<double> X((size_t)3 * m.nen), u((size_t)3 * m.nen);
    for (size_t e = 0; e < m.ele_gid.size(); ++e)
    {
      for (int a = 0; a < m.nen; ++a)
      {
        const size_t ni = (size_t)m.node_index.at(m.ele_nodes[e * m.nen + a]);
        for (int i = 0; i < 3; ++i)
        {
          X[3 * a + i] = m.node_X[3 * ni + i];
          u[3 * a + i] = u_nodes[3 * ni + i];
        }
      }
      const NurbsScope nurbs(m, e);
      mulf_history(m.ct, X.data(), u.data(), &js->job.mulf_hist[(size_t)18 * ngp * e], js->job.form == FORM_MULF_FBAR);
    }
  }
  // struct_calc_recover of the EAS elements (default step): du_nodes [n_nodes][3] = the Newton increment in the mesh's node order
  void oracle_job_recover(void* h, const double* du_nodes, double step_length)
  {
    auto* js = (JobState*)h;
    const Mesh& m = js->job.mesh;
    if (js->job.eas_hist.empty()) return;
    const bool shell = js->job.form == FORM_SHELL_EAS_ANS;
    const int nd = 3 * m.nen, neas = shell ? shell_eas_modes(m.ct, js->job.form) : eas_modes(js->job.form);
    const int nh = shell ? shell_history_doubles(neas, nd) : eas_history_doubles(neas);
    double du[24];
    for (size_t e = 0; e < m.ele_gid.size(); ++e)
    {
      for (int a = 0; a < m.nen; ++a)
      {
        const size_t ni = (size_t)m.node_index.at(m.ele_nodes[e * m.nen + a]);
        for (int i = 0; i < 3; ++i) du[3 * a + i] = du_nodes[3 * ni + i];
      }
      if (shell)
        shell_recover(neas, nd, &js->job.eas_hist[(size_t)nh * e], du, step_length);
      else
        eas_recover(neas, &js->job.eas_hist[(size_t)nh * e], du, step_length);
    }
  }
  // the job's EAS history [n_ele][eas_history_doubles]; get returns its length (hist may be NULL)
  long long oracle_job_get_eas_history(void* h, double* hist)
  {
    auto* js = (JobState*)h;
    if (hist) std::copy(js->job.eas_hist.begin(), js->job.eas_hist.end(), hist);
    return (long long)js->job.eas_hist.size();
  }
  void oracle_job_set_eas_history(void* h, const double* hist)
  {
    auto* js = (JobState*)h;
    std::copy(hist, hist + js->job.eas_hist.size(), js->job.eas_hist.begin());
  }
  // one hex8 EAS element (neas = 9 | 21) with the history hist [eas_history_doubles(neas)]: fe [24] and Ke [24 x 24] are
  // overwritten (Ke may be NULL); invKaa, Kda, s of the history are written.  recover: alpha += step * alpha_inc(du).
  int oracle_eas_history_doubles(int neas) { return eas_history_doubles(neas); }
  int oracle_element_evaluate_eas(int neas, int kind, double young, double nue, const double* X, const double* u, double* hist,
      double* fe, double* Ke, int linear)
  {
    MaterialParams mp{kind, young, nue};
    std::memset(fe, 0, sizeof(double) * 24);
    if (Ke) std::memset(Ke, 0, sizeof(double) * 576);
    return element_evaluate_eas(neas, mp, X, u, hist, fe, Ke, nullptr, 0, nullptr, 0, 0, linear != 0);
  }
  void oracle_element_recover_eas(int neas, double* hist, const double* du, double step_length) { eas_recover(neas, hist, du, step_length); }
  // the job's prestress history [n_ele][ngp][18] (inverse Jacobian | deformation gradient, both column-major); returns its length
  long long oracle_job_get_mulf_history(void* h, double* hist)
  {
    auto* js = (JobState*)h;
    if (hist) std::copy(js->job.mulf_hist.begin(), js->job.mulf_hist.end(), hist);
    return (long long)js->job.mulf_hist.size();
  }
  // Gauss-point stresses of element e (index in the mesh's element order) with the job's formulation, material, fibres and
  // history: gp_stress [ngp][6]
  int oracle_job_element_gp_stress(void* h, int e, const double* u_nodes, int stress_type, double* gp_stress)
  {
    auto* js = (JobState*)h;
    const Mesh& m = js->job.mesh;
    const int ngp = num_gp(m.ct);
    std::vector<double> X((size_t)3 * m.nen), u((size_t)3 * m.nen);
    for (int a = 0; a < m.nen; ++a)
    {
      const size_t ni = (size_t)m.node_index.at(m.ele_nodes[(size_t)e * m.nen + a]);
      for (int i = 0; i < 3; ++i)
      {
        X[3 * a + i] = m.node_X[3 * ni + i];
        u[3 * a + i] = u_nodes[3 * ni + i];
      }
    }
    const Job& job = js->job;
    const NurbsScope nurbs(m, (size_t)e);
    const double* fib = job.ele_fiber.empty() ? nullptr : job.ele_fiber.data() + (job.fiber_gp_stride ? (size_t)job.fiber_gp_stride * ngp : 3) * (size_t)e;
    if (job.form == FORM_SHELL_ANS || job.form == FORM_SHELL_EAS_ANS)
    {
      const int neas = shell_eas_modes(m.ct, job.form);
      const bool strain = stress_type >= OUTPUT_STRAIN;
      return element_evaluate_shell(m.ct, neas, job.mp, X.data(), u.data(),
          neas ? job.eas_hist.data() + (size_t)shell_history_doubles(neas, 3 * m.nen) * e : nullptr, nullptr, nullptr, fib, job.fiber_gp_stride,
          gp_stress, strain ? 1 : 0, strain ? stress_type - OUTPUT_STRAIN : stress_type);
    }
    if (!job.eas_hist.empty())
    {
      const int neas = eas_modes(job.form);
      const bool strain = stress_type >= OUTPUT_STRAIN;
      return element_evaluate_eas(neas, job.mp, X.data(), u.data(), job.eas_hist.data() + (size_t)eas_history_doubles(neas) * e, nullptr,
          nullptr, fib, job.fiber_gp_stride, gp_stress, strain ? 1 : 0, strain ? stress_type - OUTPUT_STRAIN : stress_type, eas_linear(job.form));
    }
    return element_evaluate(m.ct, job.form, job.mp, X.data(), u.data(), nullptr, nullptr, nullptr, 0.0, gp_stress, stress_type, fib,
        job.fiber_gp_stride, job.mulf_hist.empty() ? nullptr : job.mulf_hist.data() + (size_t)18 * mulf_records(m.ct, job.form) * e);
  }
  // MAT_ElastHyper with a summand list (kinds [n], par [n][4]) and optional element fibres FIBER1 [n_ele][3] in the job's
  // element order (NULL = none)
  void oracle_job_set_elasthyper(void* h, int form, int n, const int* kinds, const double* par, const double* ele_fiber)
  {
    auto* js = (JobState*)h;
    js->job.form = form;
    MaterialParams mp{MAT_ELASTHYPER, 0.0, 0.0};
    mp.n_summands = n;
    for (int k = 0; k < n && k < 4; ++k)
    {
      mp.sum_kind[k] = kinds[k];
      for (int j = 0; j < 4; ++j) mp.sum_par[k][j] = par[4 * k + j];
    }
    js->job.mp = mp;
    js->job.ele_fiber.clear();
    js->job.fiber_gp_stride = 0;
    if (ele_fiber) js->job.ele_fiber.assign(ele_fiber, ele_fiber + 3 * js->job.mesh.ele_gid.size());
    mulf_init(js);
  }
  // one element with a summand list: fe, Ke (col-major), optional Gauss-point stresses
  // nodal fibres (FNODE ... FIBER1, materials with INIT 3): node_fiber [n_nodes][3] in the order of the mesh's nodes;
  // projected to the Gauss points of every element
  void oracle_job_set_nodal_fibers(void* h, const double* node_fiber)
  {
    auto* js = static_cast<JobState*>(h);
    const Mesh& m = js->job.mesh;
    const int ngp = num_gp(m.ct);
    js->job.ele_fiber.assign((size_t)3 * ngp * m.ele_gid.size(), 0.0);
    js->job.fiber_gp_stride = 3;
    std::vector<double> nodal((size_t)3 * m.nen);
    for (size_t e = 0; e < m.ele_gid.size(); ++e)
    {
      for (int a = 0; a < m.nen; ++a)
      {
        const int ni = m.node_index.at(m.ele_nodes[e * m.nen + a]);
        for (int i = 0; i < 3; ++i) nodal[3 * a + i] = node_fiber[3 * (size_t)ni + i];
      }
      project_fibers_to_gauss_points(m.ct, nodal.data(), &js->job.ele_fiber[(size_t)3 * ngp * e]);
    }
  }
  void oracle_project_fibers_to_gauss_points(int ct, const double* nodal, double* gp_fiber) { project_fibers_to_gauss_points(ct, nodal, gp_fiber); }

  // fiber: [3] FIBER1 of the element, or [ngp][3] fibres at the Gauss points when fiber_gp_stride = 3
  int oracle_element_evaluate_elasthyper(int ct, int form, int n, const int* kinds, const double* par, const double* fiber, const double* X,
      const double* u, double* fe, double* Ke, double* gp_stress, int stress_type, int fiber_gp_stride)
  {
    MaterialParams mp{MAT_ELASTHYPER, 0.0, 0.0};
    mp.n_summands = n;
    for (int k = 0; k < n && k < 4; ++k)
    {
      mp.sum_kind[k] = kinds[k];
      for (int j = 0; j < 4; ++j) mp.sum_par[k][j] = par[4 * k + j];
    }
    return element_evaluate(ct, form, mp, X, u, fe, Ke, nullptr, 0.0, gp_stress, stress_type, fiber, fiber_gp_stride);
  }
  int oracle_job_nranks(void* h) { return ((JobState*)h)->job.part.nranks; }
  void oracle_job_subdivisions(void* h, int* s)
  {
    for (int d = 0; d < 3; ++d) s[d] = ((JobState*)h)->job.part.subdiv[d];
  }
  int oracle_job_n_nodes(void* h) { return (int)((JobState*)h)->job.mesh.node_gid.size(); }
  int oracle_job_n_eles(void* h) { return (int)((JobState*)h)->job.mesh.ele_gid.size(); }
  void oracle_job_get_nodes(void* h, int* gid, double* X, int* owner)
  {
    auto* js = (JobState*)h;
    copy_out(js->job.mesh.node_gid, gid);
    copy_out(js->job.mesh.node_X, X);
    copy_out(js->job.part.node_owner, owner);
  }
  void oracle_job_get_eles(void* h, int* gid, int* conn, int* owner)
  {
    auto* js = (JobState*)h;
    copy_out(js->job.mesh.ele_gid, gid);
    copy_out(js->job.mesh.ele_nodes, conn);
    copy_out(js->job.part.ele_owner, owner);
  }
  // sizes[0..7] = n_row_nodes, n_col_nodes, n_row_eles, n_col_eles, n_row_dof, n_col_dof, nnz_lo, nnz_hi
  void oracle_job_rank_sizes(void* h, int rank, int64_t* sizes)
  {
    auto* js = (JobState*)h;
    const RankSystem& rs = js->job.sys[rank];
    sizes[0] = rs.n_row_nodes;
    sizes[1] = (int64_t)rs.col_node_gid.size();
    sizes[2] = js->job.part.n_row_eles[rank];
    sizes[3] = (int64_t)js->job.part.col_eles[rank].size();
    sizes[4] = (int64_t)rs.row_dof_gid.size();
    sizes[5] = (int64_t)rs.col_dof_gid.size();
    sizes[6] = (int64_t)rs.colind.size();
  }
  void oracle_job_rank_maps(void* h, int rank, int* col_node_gid, int* col_eles, int* row_dof_gid,
      int* col_dof_gid, int64_t* rowptr, int* colind)
  {
    auto* js = (JobState*)h;
    const RankSystem& rs = js->job.sys[rank];
    copy_out(rs.col_node_gid, col_node_gid);
    copy_out(js->job.part.col_eles[rank], col_eles);
    copy_out(rs.row_dof_gid, row_dof_gid);
    copy_out(rs.col_dof_gid, col_dof_gid);
    copy_out(rs.rowptr, rowptr);
    copy_out(rs.colind, colind);
  }
  void oracle_job_set_dirichlet_rows(void* h, int rank, int n, const int* row_lids)
  {
    auto* js = (JobState*)h;
    auto& d = js->job.sys[rank].dbc_row;
    std::fill(d.begin(), d.end(), 0);
    for (int i = 0; i < n; ++i) d[row_lids[i]] = 1;
  }
  void oracle_job_set_u(void* h, int rank, const double* u_row)
  {
    auto* js = (JobState*)h;
    std::memcpy(js->u_row[rank].data(), u_row, js->u_row[rank].size() * sizeof(double));
  }
  // u given per global node (mesh node order, 3 per node): scattered to every rank's row layout
  void oracle_job_set_u_nodal(void* h, const double* u_nodes)
  {
    auto* js = (JobState*)h;
    const Mesh& m = js->job.mesh;
    for (int r = 0; r < js->job.part.nranks; ++r)
    {
      const auto& rn = js->job.part.row_nodes[r];
      for (size_t i = 0; i < rn.size(); ++i)
      {
        const int ni = m.node_index.at(rn[i]);
        for (int j = 0; j < 3; ++j) js->u_row[r][3 * i + j] = u_nodes[3 * (size_t)ni + j];
      }
    }
  }
  // zero f and K (reset(), sparsematrix.cpp:353-378), then evaluate all ranks.
  // returns #failing elements; seconds = wall time of the evaluate+assemble region only.
  int oracle_job_evaluate(void* h, int action, int max_threads, int* first_bad, double* seconds)
  {
    auto* js = (JobState*)h;
    const int nr = js->job.part.nranks;
    for (int r = 0; r < nr; ++r)
    {
      std::fill(js->f_row[r].begin(), js->f_row[r].end(), 0.0);
      if (action == 0) js->K_values[r].assign(js->job.sys[r].colind.size(), 0.0);
    }
    const auto t0 = std::chrono::steady_clock::now();
    const int bad =
        evaluate_job(js->job, action, js->u_row, js->f_row, js->K_values, max_threads, first_bad);
    const auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    return bad;
  }
  void oracle_job_get_f(void* h, int rank, double* f) { copy_out(((JobState*)h)->f_row[rank], f); }
  void oracle_job_get_K(void* h, int rank, double* K) { copy_out(((JobState*)h)->K_values[rank], K); }
}
