"""SURVEY 8f F4 cell types (hex20, tet4, wedge6, pyramid5; pyramid5 also with F-bar) on the generic kernel: CUDA against the CPU
oracle on the meshes of the reference's own element tests (tests/golden), with the nodes moved off their regular places so
that no Jacobian is constant.  Element level (fe, Ke), assembled system, consistent mass and Neumann loads."""
import numpy as np
import pytest
import structure_driver as sd

pytestmark = pytest.mark.gpu
TOL = 1e-12
DISP, FBAR, LINEAR = 0, 1, 2
SVK, NH = 0, 1

CASES = [
    ("solid_ele_hex20_Standard_dynamic", DISP, NH, 0.25),
    ("solid_ele_hex20_Standard_dynamic", DISP, SVK, 0.3),
    ("solid_ele_tet4_Standard_dynamic", DISP, NH, 0.25),
    ("solid_ele_tet4_Standard_dynamic", DISP, SVK, 0.3),
    ("solid_ele_wedge6_Standard_dynamic", DISP, NH, 0.25),
    ("solid_ele_pyramid5_Standard_dynamic", DISP, NH, 0.25),
    ("solid_ele_pyramid5_Standard_dynamic", FBAR, NH, 0.45),
    ("solid_ele_pyramid5_Standard_dynamic", FBAR, SVK, 0.3),
    # KINEM linear on every cell type (hex8 / hex27 / tet10 meshes of the reference's element tests as well)
    ("solid_ele_hex8_Standard_eas_none_dynamic", LINEAR, SVK, 0.3),
    ("solid_ele_hex27_Standard_dynamic", LINEAR, SVK, 0.3),
    ("solid_ele_tet10_Standard_dynamic", LINEAR, NH, 0.25),
    ("solid_ele_hex20_Standard_dynamic", LINEAR, SVK, 0.3),
    ("solid_ele_tet4_Standard_dynamic", LINEAR, SVK, 0.3),
    ("solid_ele_wedge6_Standard_dynamic", LINEAR, SVK, 0.3),
    ("solid_ele_pyramid5_Standard_dynamic", LINEAR, NH, 0.25),
    # hex18 (quad9 x line2)
    ("solid_ele_hex18_Standard_dynamic", DISP, NH, 0.25),
    ("solid_ele_hex18_Standard_dynamic", DISP, SVK, 0.3),
    ("solid_ele_hex18_Standard_dynamic", LINEAR, SVK, 0.3),
]


def distorted(oracle, name):
    p = sd.Problem(sd.load(name), oracle.node_param_coords)
    rng = np.random.default_rng(11)
    X = p.X + 0.04 * (rng.random(p.X.shape) - 0.5)
    u = 0.05 * (rng.random(p.X.shape) - 0.5) + 0.03 * X * X[:, ::-1]
    return p, X, u


@pytest.mark.parametrize("name,form,mat,nue", CASES)
def test_element_and_assembly_parity(fc, oracle, name, form, mat, nue):
    p, X, u = distorted(oracle, name)
    ct, conn = p.ct, p.conn
    s = fc.Solid(0)
    s.set_mesh(ct, form, X, p.node_gid, np.zeros(len(X), np.int32), conn, p.ele_gid)
    s.set_material(mat, 10.0, nue)
    n_row, n_col, nnz = s.build_graph()
    bad, fe_ref, Ke_ref = oracle.element_evaluate_batch(ct, form, mat, 10.0, nue, X[conn], u[conn])
    assert bad == 0
    fe, Ke = s.evaluate_elements(np.ascontiguousarray(u.reshape(-1)), want_K=True)
    for e in range(len(conn)):
        kr, fr = Ke_ref[e], fe_ref[e]
        assert np.max(np.abs(Ke[e] - kr)) <= TOL * np.max(np.abs(kr)), (e, np.max(np.abs(Ke[e] - kr)) / np.max(np.abs(kr)))
        assert np.max(np.abs(fe[e] - fr)) <= TOL * max(np.max(np.abs(fr)), np.max(np.abs(kr))), e
    # assembled system against the oracle's owner-row assembly, and the residual-only action
    job = oracle.Job.from_mesh(ct, p.node_gid, X, p.ele_gid, p.node_gid[conn])
    job.set_problem(form, mat, 10.0, nue)
    o = job.rank(0)
    job.set_u_nodal(u)
    assert job.evaluate(oracle.ACTION_NLNSTIFF)[0] == 0
    f_ref, K_ref = job.get_f(0, n_row), job.get_K(0, o["nnz"])
    rowptr, colind = s.get_graph()
    assert nnz == o["nnz"] and np.array_equal(rowptr, o["rowptr"]) and np.array_equal(colind, o["colind"])
    K, f = np.zeros(nnz), np.zeros(n_row)
    assert s.evaluate(fc.FC_ACTION_NLNSTIFF, u.reshape(-1), f, K, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
    starts = np.asarray(rowptr[:-1], np.int64)
    rmax = np.maximum.reduceat(np.abs(K_ref), starts)
    emax = np.maximum.reduceat(np.abs(K - K_ref), starts)
    assert np.max(emax / rmax) <= TOL
    assert np.max(np.abs(f - f_ref)) <= TOL * max(np.max(np.abs(f_ref)), np.max(np.abs(K_ref)))
    f2 = np.zeros(n_row)
    assert s.evaluate(fc.FC_ACTION_INTERNALFORCE, u.reshape(-1), f2, None, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
    assert np.max(np.abs(f2 - f)) <= 1e-14 * np.max(np.abs(f))
    if (ct, form, mat) == (0, LINEAR, SVK):  # default: the tensor-core closed form (k_hex8_mma_ext<VAR_SVK_LINEAR>); also the generic kernel
        K3, f3 = np.zeros(nnz), np.zeros(n_row)
        assert s.evaluate(fc.FC_ACTION_NLNSTIFF, u.reshape(-1), f3, K3, flags=fc.FC_EVAL_ZERO_FIRST | fc.FC_EVAL_FORCE_GENERIC) == (0, -1)
        assert np.max(np.maximum.reduceat(np.abs(K3 - K_ref), starts) / rmax) <= TOL
        assert np.max(np.abs(f3 - f_ref)) <= TOL * max(np.max(np.abs(f_ref)), np.max(np.abs(K_ref)))
    s.close()


@pytest.mark.parametrize("name", ["solid_ele_hex20_Standard_dynamic", "solid_ele_tet4_Standard_dynamic", "solid_ele_wedge6_Standard_dynamic",
                                  "solid_ele_pyramid5_Standard_dynamic", "solid_ele_hex18_Standard_dynamic"])
def test_mass_and_neumann_parity(fc, oracle, name):
    p, X, _ = distorted(oracle, name)
    ct, conn = p.ct, p.conn
    s = fc.Solid(0)
    s.set_mesh(ct, DISP, X, p.node_gid, np.zeros(len(X), np.int32), conn, p.ele_gid)
    s.set_material(NH, 10.0, 0.25)
    n_row, n_col, nnz = s.build_graph()
    rowptr, colind = s.get_graph()
    # consistent mass (tet4: tet_4point, the others their stiffness rule)
    rho = 0.37
    M = np.zeros(nnz)
    s.evaluate_mass(rho, M)
    Mo = np.zeros((n_row, n_row))
    for en in conn:
        err, _, _, Me = oracle.element_evaluate(ct, 0, 1, 10.0, 0.25, X[en], np.zeros((len(en), 3)), want_K=False, want_M=True, rho=rho)
        assert err == 0
        dofs = (3 * en[:, None] + np.arange(3)).reshape(-1)
        Mo[np.ix_(dofs, dofs)] += Me
    assert np.max(np.abs(sd.csr_to_dense(rowptr, colind, M, n_row) - Mo)) <= 1e-13 * np.max(np.abs(Mo))
    # volume loads on every second element, surface loads on the loaded faces of the reference's input (quad8 / tri3 / quad4)
    rng = np.random.default_rng(3)
    els = np.arange(len(conn), dtype=np.int32)[::2]
    xg = s.neumann_volume_gp_coords(els)
    fac = 1.0 + rng.random(xg.shape)
    fe_ref, xg_ref = oracle.volume_neumann(ct, X[conn[els]], [1, 0, 1], [2.0, 9.0, -0.5], fac, want_coords=True)
    assert np.max(np.abs(xg - xg_ref)) <= 1e-14
    f_ref = np.full(n_row, 0.125)
    np.add.at(f_ref.reshape(-1, 3), conn[els], fe_ref)
    f = np.full(n_row, 0.125)
    s.neumann_volume(els, [1, 0, 1], [2.0, 9.0, -0.5], f, gp_factor=fac)
    assert np.max(np.abs(f - f_ref)) <= 1e-14 * np.max(np.abs(f_ref))
    assert p.face_lists
    for c, ft, faces in p.face_lists:
        xs = s.neumann_surface_gp_coords(ft, faces)
        fac = 1.0 + rng.random(xs.shape)
        fs_ref, xs_ref = oracle.surface_neumann(ft, X[faces], [1, 1, 1], [1.0, -2.0, 0.3], fac, want_coords=True)
        assert np.max(np.abs(xs - xs_ref)) <= 1e-14
        f_ref = np.zeros(n_row)
        np.add.at(f_ref.reshape(-1, 3), faces, fs_ref)
        f = np.zeros(n_row)
        s.neumann_surface(ft, faces, [1, 1, 1], [1.0, -2.0, 0.3], f, gp_factor=fac)
        assert np.max(np.abs(f - f_ref)) <= 1e-14 * np.max(np.abs(f_ref))
    s.close()


@pytest.mark.parametrize("name", ["solid_ele_hex8_Standard_eas_none_dynamic", "solid_ele_hex20_Standard_dynamic", "solid_ele_tet10_Standard_dynamic",
                                  "solid_ele_hex18_Standard_dynamic"])
def test_orthopressure_parity(fc, oracle, name):
    """TYPE (pseudo_)orthopressure on the loaded faces of the reference's inputs: pressure along the normal of X + u, and of X"""
    p, X, u = distorted(oracle, name)
    s = fc.Solid(0)
    s.set_mesh(p.ct, DISP, X, p.node_gid, np.zeros(len(X), np.int32), p.conn, p.ele_gid)
    n_row = 3 * len(X)
    rng = np.random.default_rng(4)
    assert p.face_lists
    for c, ft, faces in p.face_lists:
        fac = 1.0 + rng.random((len(faces), fc.FACE_NGP[ft], 3))
        for uu in (u, None):
            xc = X[faces] + (0.0 if uu is None else uu[faces])
            f_ref = np.full(n_row, 0.5)
            np.add.at(f_ref.reshape(-1, 3), faces, oracle.surface_neumann_orthopressure(ft, xc, -1.7, fac))
            f = np.full(n_row, 0.5)
            s.neumann_surface_orthopressure(ft, faces, None if uu is None else uu.reshape(-1), -1.7, f, gp_factor=fac)
            assert np.max(np.abs(f - f_ref)) <= 1e-14 * np.max(np.abs(f_ref))
        # LOADLIN yes: the load stiffness in the CSR layout of the graph against the oracle's face matrices
        n_dof, n_col, nnz = s.build_graph() if not getattr(s, "_g", None) else s._g
        s._g = (n_dof, n_col, nnz)
        rowptr, colind = s.get_graph()
        K = np.zeros(nnz)
        s.neumann_surface_orthopressure_linearization(ft, faces, u.reshape(-1), -1.7, K, gp_factor=fac)
        K_ref = np.zeros((n_dof, n_dof))
        for nodes, ke in zip(faces, oracle.surface_neumann_orthopressure_linearization(ft, X[faces] + u[faces], -1.7, fac)):
            dofs = (3 * nodes[:, None] + np.arange(3)).reshape(-1)
            K_ref[np.ix_(dofs, dofs)] += ke
        assert np.max(np.abs(sd.csr_to_dense(rowptr, colind, K, n_dof) - K_ref)) <= 1e-14 * np.max(np.abs(K_ref))
    s.close()


@pytest.mark.parametrize("name,form", [("solid_ele_hex20_Standard_dynamic", DISP), ("solid_ele_tet4_Standard_dynamic", DISP),
                                       ("solid_ele_wedge6_Standard_dynamic", DISP), ("solid_ele_pyramid5_Standard_dynamic", DISP),
                                       ("solid_ele_pyramid5_Standard_dynamic", FBAR), ("solid_ele_hex18_Standard_dynamic", DISP)])
def test_stress_output(fc, oracle, name, form):
    """fc_solid_evaluate_stress on the F4 cells: Gauss-point stresses against the oracle, nodal values against the driver's
    restatement of the extrapolation (base cell per Gauss-point count; pyramid5: the reference's fixed matrix)"""
    import torch
    p, X, u = distorted(oracle, name)
    p.X = X
    s = fc.Solid(0)
    s.set_mesh(p.ct, form, X, p.node_gid, np.zeros(len(X), np.int32), p.conn, p.ele_gid)
    s.set_material(NH, 10.0, 0.25)
    s.build_graph()
    E = sd.extrapolation_matrix(p.ct, oracle.gauss_rule(p.ct)[0], oracle.shape, oracle.node_param_coords(p.ct))
    for st in (fc.FC_STRESS_2PK, fc.FC_STRESS_CAUCHY):
        gp, nodal = s.evaluate_stress(u.reshape(-1), st)
        ref = [oracle.element_gp_stress(p.ct, form, NH, 10.0, 0.25, X[en], u[en], st)[1] for en in p.conn]
        for e in range(len(p.conn)):
            assert np.max(np.abs(gp[e] - ref[e])) <= 1e-12 * np.max(np.abs(ref[e]))
        nodal_ref = sd.nodal_stresses(p, lambda e: ref[e], E)
        assert np.max(np.abs(nodal - nodal_ref)) <= 1e-12 * np.max(np.abs(nodal_ref))
    # the allocation-free variant keeps to hex8 / hex27 / tet10
    with pytest.raises(fc.FourCError) as ei:
        s.evaluate_stress_noalloc(torch.from_numpy(u.reshape(-1)).cuda(), torch.zeros((len(X), 6), dtype=torch.float64, device="cuda"))
    assert ei.value.code == fc.FC_ERR_UNSUPPORTED
    s.close()


@pytest.mark.parametrize("name,mat,MULF", [("solid_ele_hex8_Standard_eas_none_dynamic", NH, 3), ("solid_ele_hex27_Standard_dynamic", NH, 3),
                                           ("solid_ele_tet10_Standard_dynamic", SVK, 3), ("solid_ele_hex20_Standard_dynamic", NH, 3),
                                           ("solid_ele_tet4_Standard_dynamic", NH, 3), ("solid_ele_wedge6_Standard_dynamic", SVK, 3),
                                           ("solid_ele_pyramid5_Standard_dynamic", NH, 3), ("solid_ele_hex18_Standard_dynamic", NH, 3),
                                           ("solid_ele_hex8_Standard_eas_none_dynamic", NH, 6), ("solid_ele_pyramid5_Standard_dynamic", SVK, 6)])
def test_mulf_prestress_parity(fc, oracle, name, mat, MULF):
    """PRESTRESS_TECH mulf on every cell type (MULF = 6: with TECH fbar, hex8 and pyramid5): two prestress updates with different
    displacements, then the assembled system, the Gauss-point stresses and the history itself against an oracle job that saw the
    same updates; history round trip."""
    assert (fc.FC_FORM_MULF, fc.FC_FORM_MULF_FBAR) == (3, 6)
    p, X, u = distorted(oracle, name)
    ct, conn = p.ct, p.conn
    s = fc.Solid(0)
    s.set_mesh(ct, MULF, X, p.node_gid, np.zeros(len(X), np.int32), conn, p.ele_gid)
    s.set_material(mat, 10.0, 0.3)
    n_row, n_col, nnz = s.build_graph()
    job = oracle.Job.from_mesh(ct, p.node_gid, X, p.ele_gid, p.node_gid[conn])
    job.set_problem(MULF, mat, 10.0, 0.3)
    o = job.rank(0)
    assert np.max(np.abs(s.mulf_history() - job.mulf_history())) <= 1e-14 * np.max(np.abs(job.mulf_history()))
    for k, uk in enumerate((0.6 * u, -0.3 * u[::-1])):
        s.update_prestress(uk.reshape(-1) if k == 0 else __import__("torch").from_numpy(np.ascontiguousarray(uk.reshape(-1))).cuda())
        job.update_prestress(uk)
        h_ref = job.mulf_history()
        assert np.max(np.abs(s.mulf_history() - h_ref)) <= 1e-14 * np.max(np.abs(h_ref))
    job.set_u_nodal(u)
    assert job.evaluate(oracle.ACTION_NLNSTIFF)[0] == 0
    f_ref, K_ref = job.get_f(0, n_row), job.get_K(0, o["nnz"])
    rowptr, _ = s.get_graph()
    K, f = np.zeros(nnz), np.zeros(n_row)
    assert s.evaluate(fc.FC_ACTION_NLNSTIFF, u.reshape(-1), f, K, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
    starts = np.asarray(rowptr[:-1], np.int64)
    assert np.max(np.maximum.reduceat(np.abs(K - K_ref), starts) / np.maximum.reduceat(np.abs(K_ref), starts)) <= TOL
    assert np.max(np.abs(f - f_ref)) <= TOL * max(np.max(np.abs(f_ref)), np.max(np.abs(K_ref)))
    f2 = np.zeros(n_row)
    assert s.evaluate(fc.FC_ACTION_INTERNALFORCE, u.reshape(-1), f2, None, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
    assert np.max(np.abs(f2 - f)) <= 1e-14 * np.max(np.abs(f))
    for st in (fc.FC_STRESS_2PK, fc.FC_STRESS_CAUCHY):
        gp, _ = s.evaluate_stress(u.reshape(-1), st, want_nodal=False)
        for e in range(len(conn)):
            ref = job.element_gp_stress(e, u, st)[1]
            assert np.max(np.abs(gp[e] - ref)) <= 1e-12 * np.max(np.abs(ref))
    # restart: the history moved to a fresh context reproduces the system; the unsupported operators say so
    s2 = fc.Solid(0)
    s2.set_mesh(ct, MULF, X, p.node_gid, np.zeros(len(X), np.int32), conn, p.ele_gid)
    s2.set_material(mat, 10.0, 0.3)
    s2.build_graph()
    s2.set_mulf_history(s.mulf_history())
    K2, f3 = np.zeros(nnz), np.zeros(n_row)
    assert s2.evaluate(fc.FC_ACTION_NLNSTIFF, u.reshape(-1), f3, K2, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
    assert np.max(np.abs(K2 - K)) <= 1e-14 * np.max(np.abs(K)) and np.max(np.abs(f3 - f)) <= 1e-14 * np.max(np.abs(f))
    s2.close()
    s3 = fc.Solid(0)
    s3.set_mesh(ct, DISP, X, p.node_gid, np.zeros(len(X), np.int32), conn, p.ele_gid)
    with pytest.raises(fc.FourCError) as ei:
        s3.update_prestress(u.reshape(-1))
    assert ei.value.code == fc.FC_ERR_INVALID
    s3.close()
    s.close()


@pytest.mark.parametrize("form,mat", [(4, NH), (4, SVK), (5, NH), (5, SVK), (7, NH), (10, SVK), (11, NH)])
def test_eas_parity(fc, oracle, form, mat):
    """hex8 TECH eas_mild / eas_full on a distorted 3x2x2 block: three rounds of evaluate + recover with the library and with an
    oracle job; element fe / Ke, assembled system, history (alpha, s, Kaa^-1, Kda) and Gauss-point stresses / strains."""
    neas = oracle.NEAS[form]
    m = fc.Grid(fc.FC_HEX8, (3, 2, 2), (0.0, 0.0, 0.0), (1.5, 1.0, 1.0))
    rng = np.random.default_rng(5)
    X = m.X + 0.06 * (rng.random(m.X.shape) - 0.5)
    conn = m.conn
    u = 0.05 * (rng.random(X.shape) - 0.5) + 0.04 * X * X[:, ::-1]
    s = fc.Solid(0)
    s.set_mesh(fc.FC_HEX8, form, X, m.node_gid, np.zeros(len(X), np.int32), conn, m.ele_gid)
    s.set_material(mat, 10.0, 0.3)
    n_row, n_col, nnz = s.build_graph()
    rowptr, _ = s.get_graph()
    starts = np.asarray(rowptr[:-1], np.int64)
    job = oracle.Job.from_mesh(fc.FC_HEX8, m.node_gid, X, m.ele_gid, m.node_gid[conn])
    job.set_problem(form, mat, 10.0, 0.3)
    o = job.rank(0)
    assert s.eas_history().shape == (len(conn), oracle.eas_history_doubles(neas))
    assert np.array_equal(s.eas_history(), job.eas_history())
    for it in range(3):
        uk = (0.4 + 0.3 * it) * u
        # element level: the oracle element with the library's history before the evaluation
        h0 = s.eas_history()
        fe, Ke = s.evaluate_elements(np.ascontiguousarray(uk.reshape(-1)), want_K=True)
        for e in range(len(conn)):
            he = np.ascontiguousarray(h0[e])
            err, fr, kr = oracle.element_evaluate_eas(neas, mat, 10.0, 0.3, X[conn[e]], uk[conn[e]], he, linear=form in (10, 11))
            assert err == 0
            assert np.max(np.abs(Ke[e] - kr)) <= 1e-11 * np.max(np.abs(kr)), (it, e, np.max(np.abs(Ke[e] - kr)) / np.max(np.abs(kr)))
            assert np.max(np.abs(fe[e].reshape(8, 3) - fr)) <= 1e-11 * max(np.max(np.abs(fr)), np.max(np.abs(kr)))
            assert np.max(np.abs(s.eas_history()[e] - he)) <= 1e-10 * np.max(np.abs(he))
        # assembled system: both sides continue from their own history
        job.set_u_nodal(uk)
        assert job.evaluate(oracle.ACTION_NLNSTIFF)[0] == 0
        f_ref, K_ref = job.get_f(0, n_row), job.get_K(0, o["nnz"])
        K, f = np.zeros(nnz), np.zeros(n_row)
        assert s.evaluate(fc.FC_ACTION_NLNSTIFF, uk.reshape(-1), f, K, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
        assert np.max(np.maximum.reduceat(np.abs(K - K_ref), starts) / np.maximum.reduceat(np.abs(K_ref), starts)) <= 1e-10
        assert np.max(np.abs(f - f_ref)) <= 1e-10 * max(np.max(np.abs(f_ref)), np.max(np.abs(K_ref)))
        f2 = np.zeros(n_row)
        assert s.evaluate(fc.FC_ACTION_INTERNALFORCE, uk.reshape(-1), f2, None, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
        assert np.max(np.abs(f2 - f)) <= 1e-14 * np.max(np.abs(f))
        h1, h2 = s.eas_history(), job.eas_history()
        assert np.max(np.abs(h1 - h2)) <= 1e-9 * np.max(np.abs(h2))
        # recover with a made-up increment (host array, then a device tensor)
        du = 0.02 * (rng.random(X.shape) - 0.5)
        if it == 1:
            s.recover(__import__("torch").from_numpy(np.ascontiguousarray(du.reshape(-1))).cuda(), 0.75)
            __import__("torch").cuda.synchronize()
        else:
            s.recover(du.reshape(-1), 0.75)
        job.recover(du, 0.75)
        h1, h2 = s.eas_history(), job.eas_history()
        assert np.max(np.abs(h1[:, :2 * neas] - h2[:, :2 * neas])) <= 1e-9 * np.max(np.abs(h2[:, :2 * neas]))
        assert np.all(h1[:, -1] == 0.75)
    # Gauss-point output with the enhanced strains (the oracle job takes the library's history)
    job.set_eas_history(s.eas_history())
    for st in (fc.FC_STRESS_2PK, fc.FC_STRESS_CAUCHY):
        gp, _ = s.evaluate_stress(u.reshape(-1), st, want_nodal=False)
        for e in range(len(conn)):
            ref = job.element_gp_stress(e, u, st)[1]
            assert np.max(np.abs(gp[e] - ref)) <= 1e-12 * np.max(np.abs(ref))
    for st in (fc.FC_STRAIN_GL, fc.FC_STRAIN_EA, fc.FC_STRAIN_LOG):
        gp, _ = s.evaluate_strain(u.reshape(-1), st, want_nodal=False)
        for e in range(len(conn)):
            ref = job.element_gp_stress(e, u, oracle.OUTPUT_STRAIN + st)[1]
            assert np.max(np.abs(gp[e] - ref)) <= 1e-12 * np.max(np.abs(ref))
    assert np.array_equal(s.eas_history(), job.eas_history())  # the output leaves the history alone
    # EAS exists for hex8 only
    s2 = fc.Solid(0)
    p = sd.Problem(sd.load("solid_ele_tet4_Standard_dynamic"), oracle.node_param_coords)
    with pytest.raises(fc.FourCError) as ei:
        s2.set_mesh(p.ct, form, p.X, p.node_gid, np.zeros(len(p.X), np.int32), p.conn, p.ele_gid)
    assert ei.value.code == fc.FC_ERR_UNSUPPORTED
    s2.close()
    s.close()


@pytest.mark.parametrize("form", [3, 6, 4, 5, 7, 8, 9])
def test_history_formulations_on_several_ranks(fc, oracle, form):
    """MULF, MULF + F-bar, eas_mild, eas_full on a 3-rank split of a hex8 block: every rank evaluates and updates its column
    elements, ghosts included; assembled rows against the oracle's before and after a history update, and the history of an
    element held by two ranks is the same on both, bit for bit."""
    from test_gpu_parity import setup_rank, rel_per_row
    n, nranks = (6, 4, 3), 3
    job = oracle.Job.cuboid(fc.FC_HEX8, n, nranks=nranks)
    job.set_problem(form, NH, 10.0, 0.3)
    u1 = oracle.smooth_displacement(job.X, 1.0 / max(n), amp=0.01)
    rng = np.random.default_rng(2)
    u2 = 1.3 * u1 + 0.002 * (rng.random(u1.shape) - 0.5)
    ranks = [setup_rank(fc, job, r, fc.FC_HEX8, form, NH, 10.0, 0.3) for r in range(nranks)]
    tol = 1e-12 if form in (3, 6) else 1e-10

    def compare(u):
        job.set_u_nodal(u)
        assert job.evaluate(oracle.ACTION_NLNSTIFF)[0] == 0
        for r, (s, o, col) in enumerate(ranks):
            K, f = np.zeros(o["nnz"]), np.zeros(o["n_row_dof"])
            assert s.evaluate(fc.FC_ACTION_NLNSTIFF, np.ascontiguousarray(u[col].reshape(-1)), f, K, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
            K_ref, f_ref = job.get_K(r, o["nnz"]), job.get_f(r, o["n_row_dof"])
            assert rel_per_row(K, K_ref, o["rowptr"]) <= tol, (r, rel_per_row(K, K_ref, o["rowptr"]))
            assert np.max(np.abs(f - f_ref)) <= tol * max(np.max(np.abs(f_ref)), np.max(np.abs(K_ref)) / max(n))

    compare(u1)
    for s, o, col in ranks:
        if form in (3, 6):
            s.update_prestress(np.ascontiguousarray(u1[col].reshape(-1)))
        else:
            s.recover(np.ascontiguousarray((u2 - u1)[col].reshape(-1)))  # a no-op without condensed variables (form 8)
    if form in (3, 6):
        job.update_prestress(u1)
    else:
        job.recover(u2 - u1)
    compare(u2)
    # copies of an element on different ranks
    seen = {}
    shared = 0
    for s, o, col in ranks:
        if form == 8:  # ANS alone keeps no history
            shared = 1
            break
        h = s.mulf_history() if form in (3, 6) else s.eas_history()
        for k, g in enumerate(o["col_eles"]):
            if int(g) in seen:
                assert np.array_equal(seen[int(g)], h[k])
                shared += 1
            else:
                seen[int(g)] = h[k].copy()
    assert shared > 0
    for s, o, col in ranks:
        s.close()


@pytest.mark.parametrize("form,mat", [(8, NH), (9, NH), (9, SVK)])
def test_wedge6_solid_shell_parity(fc, oracle, form, mat):
    """TECH shell_ans / shell_eas_ans on wedge6 (5 sampling points, one EAS mode): two rounds of evaluate (+ recover) against an oracle
    job on a distorted copy of the reference's wedge mesh; Gauss-point stresses and strains"""
    p, X, u = distorted(oracle, "solid_ele_wedge6_Standard_dynamic")
    ct, conn = p.ct, p.conn
    s = fc.Solid(0)
    s.set_mesh(ct, form, X, p.node_gid, np.zeros(len(X), np.int32), conn, p.ele_gid)
    s.set_material(mat, 10.0, 0.3)
    n_row, n_col, nnz = s.build_graph()
    rowptr, _ = s.get_graph()
    starts = np.asarray(rowptr[:-1], np.int64)
    job = oracle.Job.from_mesh(ct, p.node_gid, X, p.ele_gid, p.node_gid[conn])
    job.set_problem(form, mat, 10.0, 0.3)
    o = job.rank(0)
    rng = np.random.default_rng(9)
    for it in range(2):
        uk = (0.5 + 0.4 * it) * u
        job.set_u_nodal(uk)
        assert job.evaluate(oracle.ACTION_NLNSTIFF)[0] == 0
        f_ref, K_ref = job.get_f(0, n_row), job.get_K(0, o["nnz"])
        K, f = np.zeros(nnz), np.zeros(n_row)
        assert s.evaluate(fc.FC_ACTION_NLNSTIFF, uk.reshape(-1), f, K, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
        assert np.max(np.maximum.reduceat(np.abs(K - K_ref), starts) / np.maximum.reduceat(np.abs(K_ref), starts)) <= 1e-10
        assert np.max(np.abs(f - f_ref)) <= 1e-10 * max(np.max(np.abs(f_ref)), np.max(np.abs(K_ref)))
        fe, Ke = s.evaluate_elements(np.ascontiguousarray(uk.reshape(-1)), want_K=True)
        assert np.all(np.isfinite(Ke)) and np.max(np.abs(Ke - Ke.transpose(0, 2, 1))) <= 1e-10 * np.max(np.abs(Ke))
        du = 0.02 * (rng.random(X.shape) - 0.5)
        s.recover(du.reshape(-1))
        job.recover(du)
        if form == 9:
            h1, h2 = s.eas_history(), job.eas_history()
            assert h1.shape == h2.shape == (len(conn), 3 + 1 + 18 + 1)
            assert np.max(np.abs(h1 - h2)) <= 1e-9 * np.max(np.abs(h2))
    if form == 9:
        job.set_eas_history(s.eas_history())
    for st in (fc.FC_STRESS_2PK, fc.FC_STRESS_CAUCHY):
        gp, _ = s.evaluate_stress(u.reshape(-1), st, want_nodal=False)
        for e in range(len(conn)):
            ref = job.element_gp_stress(e, u, st)[1]
            assert np.max(np.abs(gp[e] - ref)) <= 1e-12 * np.max(np.abs(ref))
    for st in (fc.FC_STRAIN_GL, fc.FC_STRAIN_EA, fc.FC_STRAIN_LOG):
        gp, _ = s.evaluate_strain(u.reshape(-1), st, want_nodal=False)
        for e in range(len(conn)):
            ref = job.element_gp_stress(e, u, oracle.OUTPUT_STRAIN + st)[1]
            assert np.max(np.abs(gp[e] - ref)) <= 1e-12 * np.max(np.abs(ref))
    s.close()
