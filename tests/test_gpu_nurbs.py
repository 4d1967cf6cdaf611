"""nurbs27 solid elements (the last cell of ImplementedSolidCellTypes, 4C_solid_3D_ele_factory.hpp:36-38) through the C ABI against
the CPU oracle: the reference's 2 x 3 x 4 patch (tests/golden/solid_ele_nurbs27_Standard_Static_multielement.json, non-uniform
knot spans) with moved control points and random weights.  Assembled system for the three formulations of the cell, prestress
history, mass, body and surface loads, Gauss-point stresses, a two-rank split, and what the library refuses."""
import numpy as np
import pytest
import structure_driver as sd

pytestmark = pytest.mark.gpu
TOL = 1e-12
DISP, LINEAR, MULF = 0, 2, 3
SVK, NH = 0, 1
NAME = "solid_ele_nurbs27_Standard_Static_multielement"


def patch(oracle):
    p = sd.Problem(sd.load(NAME), oracle.node_param_coords)
    assert p.ct == 8 and p.conn.shape == (24, 27) and p.ele_knots.shape == (24, 3, 6)
    rng = np.random.default_rng(5)
    X = p.X + 0.02 * (rng.random(p.X.shape) - 0.5)
    u = 0.03 * (rng.random(p.X.shape) - 0.5) + 0.02 * X * X[:, ::-1]
    p.X = X
    p.weights = 0.7 + 0.6 * rng.random(len(X))
    return p, X, u


def contexts(fc, oracle, p, form, mat, nue=0.3):
    s = fc.Solid(0)
    s.set_mesh(p.ct, form, p.X, p.node_gid, np.zeros(len(p.X), np.int32), p.conn, p.ele_gid)
    s.set_nurbs(p.ele_knots, p.weights)
    s.set_material(mat, 10.0, nue)
    job = oracle.Job.from_mesh(p.ct, p.node_gid, p.X, p.ele_gid, p.node_gid[p.conn])
    job.set_nurbs(p.ele_knots, p.node_gid, p.weights)
    job.set_problem(form, mat, 10.0, nue)
    return s, job


def compare_system(fc, oracle, s, job, u, tol=TOL, flags=0):
    n_row, n_col, nnz = s.build_graph() if not getattr(s, "_graph", None) else s._graph
    s._graph = (n_row, n_col, nnz)
    o = job.rank(0)
    job.set_u_nodal(u)
    assert job.evaluate(oracle.ACTION_NLNSTIFF)[0] == 0
    f_ref, K_ref = job.get_f(0, n_row), job.get_K(0, o["nnz"])
    rowptr, colind = s.get_graph()
    assert nnz == o["nnz"] and np.array_equal(rowptr, o["rowptr"]) and np.array_equal(colind, o["colind"])
    K, f = np.zeros(nnz), np.zeros(n_row)
    assert s.evaluate(fc.FC_ACTION_NLNSTIFF, u.reshape(-1), f, K, flags=fc.FC_EVAL_ZERO_FIRST | flags) == (0, -1)
    starts = np.asarray(rowptr[:-1], np.int64)
    assert np.max(np.maximum.reduceat(np.abs(K - K_ref), starts) / np.maximum.reduceat(np.abs(K_ref), starts)) <= tol
    assert np.max(np.abs(f - f_ref)) <= tol * max(np.max(np.abs(f_ref)), np.max(np.abs(K_ref)))
    f2 = np.zeros(n_row)
    assert s.evaluate(fc.FC_ACTION_INTERNALFORCE, u.reshape(-1), f2, None, flags=fc.FC_EVAL_ZERO_FIRST | flags) == (0, -1)
    assert np.max(np.abs(f2 - f)) <= 1e-13 * np.max(np.abs(f))
    return K, f, rowptr, colind


@pytest.mark.parametrize("form,mat", [(DISP, NH), (DISP, SVK), (LINEAR, SVK), (MULF, NH)])
def test_assembled_system(fc, oracle, form, mat):
    p, X, u = patch(oracle)
    s, job = contexts(fc, oracle, p, form, mat)
    # element level first: fe, Ke of every element
    fe, Ke = s.evaluate_elements(np.ascontiguousarray(u.reshape(-1)), want_K=True)
    for e, en in enumerate(p.conn):
        oracle.set_nurbs_element(p.ele_knots[e], p.weights[en])
        if form == MULF:
            break  # the element-level oracle call has no history argument; the assembled comparison below covers it
        err, fr, kr, _ = oracle.element_evaluate(p.ct, form, mat, 10.0, 0.3, X[en], u[en])
        assert err == 0
        assert np.max(np.abs(Ke[e] - kr)) <= TOL * np.max(np.abs(kr)), e
        assert np.max(np.abs(fe[e] - fr)) <= TOL * max(np.max(np.abs(fr)), np.max(np.abs(kr))), e
    if form == MULF:
        assert np.max(np.abs(s.mulf_history() - job.mulf_history())) <= 1e-14 * np.max(np.abs(job.mulf_history()))
        s.update_prestress(np.ascontiguousarray((0.5 * u).reshape(-1)))
        job.update_prestress(0.5 * u)
        h_ref = job.mulf_history()
        assert np.max(np.abs(s.mulf_history() - h_ref)) <= 1e-13 * np.max(np.abs(h_ref))
    compare_system(fc, oracle, s, job, u)
    if form == DISP and mat == NH:  # default: the tensor-core closed form (k_nh_mma_gen<NURBS27>); also the generic kernel
        compare_system(fc, oracle, s, job, u, flags=fc.FC_EVAL_FORCE_GENERIC)
    for st in (fc.FC_STRESS_2PK, fc.FC_STRESS_CAUCHY):
        gp, _ = s.evaluate_stress(u.reshape(-1), st, want_nodal=False)
        for e in range(len(p.conn)):
            ref = job.element_gp_stress(e, u, st)[1]
            assert np.max(np.abs(gp[e] - ref)) <= 1e-12 * np.max(np.abs(ref))
    # values at the control points: per element the inverse of its own basis at the Gauss points
    # (evaluate_gauss_points_to_nurbs_knots_extrapolation_matrix), averaged over the elements of a control point
    gp, nodal = s.evaluate_stress(u.reshape(-1), fc.FC_STRESS_CAUCHY)
    gxi = oracle.gauss_rule(p.ct)[0]
    acc, cnt = np.zeros((len(X), 6)), np.zeros(len(X))
    for e, en in enumerate(p.conn):
        oracle.set_nurbs_element(p.ele_knots[e], p.weights[en])
        M = np.array([oracle.shape(p.ct, xi) for xi in gxi])  # [gp][a]
        acc[en] += np.linalg.solve(M, job.element_gp_stress(e, u, fc.FC_STRESS_CAUCHY)[1])
        cnt[en] += 1
    ref = acc / cnt[:, None]
    assert np.max(np.abs(nodal - ref)) <= 1e-10 * np.max(np.abs(ref))
    s.close()


def test_mass_and_loads(fc, oracle):
    p, X, _ = patch(oracle)
    s, job = contexts(fc, oracle, p, DISP, NH)
    n_row, n_col, nnz = s.build_graph()
    rowptr, colind = s.get_graph()
    rho = 0.37
    M = np.zeros(nnz)
    s.evaluate_mass(rho, M)
    Mo = np.zeros((n_row, n_row))
    for e, en in enumerate(p.conn):
        oracle.set_nurbs_element(p.ele_knots[e], p.weights[en])
        err, _, _, Me = oracle.element_evaluate(p.ct, 0, 1, 10.0, 0.25, X[en], np.zeros((27, 3)), want_K=False, want_M=True, rho=rho)
        assert err == 0
        dofs = (3 * en[:, None] + np.arange(3)).reshape(-1)
        Mo[np.ix_(dofs, dofs)] += Me
    assert np.max(np.abs(sd.csr_to_dense(rowptr, colind, M, n_row) - Mo)) <= 1e-13 * np.max(np.abs(Mo))
    # body loads on every second element
    rng = np.random.default_rng(3)
    els = np.arange(len(p.conn), dtype=np.int32)[::2]
    xg = s.neumann_volume_gp_coords(els)
    fac = 1.0 + rng.random(xg.shape)
    f_ref = np.full(n_row, 0.125)
    for k, e in enumerate(els):
        en = p.conn[e]
        oracle.set_nurbs_element(p.ele_knots[e], p.weights[en])
        fe_ref, xg_ref = oracle.volume_neumann(p.ct, X[en][None], [1, 0, 1], [2.0, 9.0, -0.5], fac[k][None], want_coords=True)
        assert np.max(np.abs(xg[k] - xg_ref[0])) <= 1e-14
        np.add.at(f_ref.reshape(-1, 3), en, fe_ref[0])
    f = np.full(n_row, 0.125)
    s.neumann_volume(els, [1, 0, 1], [2.0, 9.0, -0.5], f, gp_factor=fac)
    assert np.max(np.abs(f - f_ref)) <= 1e-14 * np.max(np.abs(f_ref))
    # surface loads on all six faces of every element (the control points of neighbouring elements overlap, so these are the
    # parameter-space faces of the knot spans, inside the patch as well as on its boundary)
    faces = np.array([(e, sid) for e in range(len(p.conn)) for sid in range(6)], np.int32)
    nb = sd.OracleNeumann(oracle, p)
    xs = s.neumann_surface_nurbs_gp_coords(faces[:, 0], faces[:, 1])
    assert np.max(np.abs(xs - nb.surface_coords(6, faces))) <= 1e-14
    fac = 1.0 + rng.random(xs.shape)
    f_ref, f = np.zeros(n_row), np.zeros(n_row)
    nb.surface(6, faces, [1, 1, 1], [1.0, -2.0, 0.3], fac, f_ref)
    s.neumann_surface_nurbs(faces[:, 0], faces[:, 1], [1, 1, 1], [1.0, -2.0, 0.3], f, gp_factor=fac)
    assert np.max(np.abs(f - f_ref)) <= 1e-14 * np.max(np.abs(f_ref))
    # a constant traction sums to area times value in every direction (partition of unity on every face)
    f1 = np.zeros(n_row)
    s.neumann_surface_nurbs(faces[:, 0], faces[:, 1], [1, 1, 1], [1.0, -2.0, 0.3], f1)
    area = f1.reshape(-1, 3).sum(0) / np.array([1.0, -2.0, 0.3])
    assert np.max(np.abs(area - area[0])) <= 1e-13 * area[0] and area[0] > 0.0
    with pytest.raises(fc.FourCError):
        s.neumann_surface(fc.FC_FACE_QUAD9, p.conn[:1, :9], [1, 1, 1], [1.0, 0.0, 0.0], f1)
    with pytest.raises(fc.FourCError):
        s.neumann_surface_nurbs([0], [6], [1, 1, 1], [1.0, 0.0, 0.0], f1)
    s.close()


def test_two_ranks(fc, oracle):
    """owner by height: each rank holds its column elements (ghost nodes behind the owned ones) with their knot spans and the
    weights of its column nodes; owned rows against the one-rank oracle system"""
    p, X, u = patch(oracle)
    s1, job = contexts(fc, oracle, p, DISP, NH)
    K, f, rowptr, colind = compare_system(fc, oracle, s1, job, u)
    Kd = sd.csr_to_dense(rowptr, colind, K, len(f))
    s1.close()
    owner = (X[:, 2] > np.median(X[:, 2])).astype(np.int32)
    for r in range(2):
        els = np.array([e for e, en in enumerate(p.conn) if np.any(owner[en] == r)])
        nodes = np.unique(p.conn[els])
        col = np.concatenate([nodes[owner[nodes] == r], nodes[owner[nodes] != r]])
        lid = {int(n): i for i, n in enumerate(col)}
        conn = np.vectorize(lid.get)(p.conn[els]).astype(np.int32)
        s = fc.Solid(0, r, 2)
        s.set_mesh(p.ct, DISP, X[col], p.node_gid[col], owner[col], conn, p.ele_gid[els])
        with pytest.raises(fc.FourCError) as ei:  # the tables come with set_nurbs
            s.set_material(NH, 10.0, 0.3)
            s.evaluate(fc.FC_ACTION_INTERNALFORCE, np.zeros(3 * len(col)), np.zeros(3 * int(np.sum(owner[col] == r))), None)
        assert ei.value.code == fc.FC_ERR_INVALID
        s.set_nurbs(p.ele_knots[els], p.weights[col])
        n_row, n_col, nnz = s.build_graph()
        rp, ci = s.get_graph()
        Kr, fr = np.zeros(nnz), np.zeros(n_row)
        assert s.evaluate(fc.FC_ACTION_NLNSTIFF, np.ascontiguousarray(u[col].reshape(-1)), fr, Kr, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
        dofs = (3 * col[:, None] + np.arange(3)).reshape(-1)
        ref = Kd[np.ix_(dofs[:n_row], dofs)]
        got = np.zeros((n_row, 3 * len(col)))
        for i in range(n_row):
            got[i, ci[rp[i]:rp[i + 1]]] = Kr[rp[i]:rp[i + 1]]
        assert np.max(np.abs(got - ref)) <= TOL * np.max(np.abs(ref))
        assert np.max(np.abs(fr - f[dofs[:n_row]])) <= TOL * np.max(np.abs(K))
        s.close()


def test_refusals(fc, oracle):
    p, X, u = patch(oracle)
    s = fc.Solid(0)
    for form in (1, 4, 8):  # F-bar, EAS and the solid shells do not exist for this cell
        with pytest.raises(fc.FourCError) as ei:
            s.set_mesh(p.ct, form, X, p.node_gid, np.zeros(len(X), np.int32), p.conn, p.ele_gid)
        assert ei.value.code == fc.FC_ERR_UNSUPPORTED
    s.set_mesh(p.ct, DISP, X, p.node_gid, np.zeros(len(X), np.int32), p.conn, p.ele_gid)
    kn = p.ele_knots.copy()
    kn[3, 1, 3] = kn[3, 1, 2]  # zero-sized knot span
    with pytest.raises(fc.FourCError) as ei:
        s.set_nurbs(kn, p.weights)
    assert ei.value.code == fc.FC_ERR_INVALID
    w = p.weights.copy()
    w[5] = 0.0
    with pytest.raises(fc.FourCError):
        s.set_nurbs(p.ele_knots, w)
    # a folded element: the node-Jacobian check runs in set_nurbs and every evaluation reports it
    Xb = X.copy()
    en = p.conn[7]
    Xb[en[26]] = Xb[en[0]] - 0.5
    s.set_mesh(p.ct, DISP, Xb, p.node_gid, np.zeros(len(X), np.int32), p.conn, p.ele_gid)
    s.set_nurbs(p.ele_knots, p.weights)
    s.set_material(NH, 10.0, 0.3)
    cnt, gid = s.evaluate(fc.FC_ACTION_INTERNALFORCE, u.reshape(-1), np.zeros(3 * len(X)), None)
    assert cnt >= 1 and gid in p.ele_gid
    assert s.error_state()[0] >= 1
    s2 = fc.Solid(0)
    s2.set_mesh(fc.FC_HEX27, DISP, X[:27], p.node_gid[:27], np.zeros(27, np.int32), np.arange(27, dtype=np.int32)[None], p.ele_gid[:1])
    with pytest.raises(fc.FourCError):
        s2.set_nurbs(p.ele_knots[:1], p.weights[:27])
    s2.close()
    s.close()
