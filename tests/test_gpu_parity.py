"""Parity of the CUDA path (through the C ABI) against the CPU oracle.  FP64 tolerance of
BASELINE.json's north_star: 1e-12 relative, applied as |a-b| <= 1e-12 * max|.| per element and
per ROW of the assembled system (entry-wise relative error is meaningless on cancelling entries)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-12

HEX8, HEX27, TET10 = 0, 1, 2
DISP, FBAR = 0, 1
SVK, NH = 0, 1


def tet10_mesh(n, rng=None):
    """n^3 cubes split into 6 tets each on the (2n+1)^3 lattice; node order of
    4C_fem_general_utils_fem_shapefunctions.hpp:244-259 (4 corners, then mids 01 12 02 03 13 23)."""
    N = 2 * n + 1
    nid = lambda i, j, k: (k * N + j) * N + i
    tets = []
    # Kuhn split of the cube along the (0,0,0)-(1,1,1) diagonal
    perms = [(0, 1, 2), (0, 2, 1), (1, 0, 2), (1, 2, 0), (2, 0, 1), (2, 1, 0)]
    for k in range(n):
        for j in range(n):
            for i in range(n):
                o = np.array([2 * i, 2 * j, 2 * k])
                for p in perms:
                    c = [o.copy()]
                    for ax in p:
                        v = c[-1].copy()
                        v[ax] += 2
                        c.append(v)
                    c = np.array(c)
                    # positive orientation
                    if np.linalg.det(c[1:] - c[0]) < 0:
                        c[[2, 3]] = c[[3, 2]]
                    mids = [(0, 1), (1, 2), (0, 2), (0, 3), (1, 3), (2, 3)]
                    nodes = [nid(*v) for v in c] + [nid(*((c[a] + c[b]) // 2)) for a, b in mids]
                    tets.append(nodes)
    tets = np.array(tets, np.int32)
    used = np.unique(tets)
    X = np.zeros((len(used), 3))
    X[:, 0] = (used % N) / (2 * n)
    X[:, 1] = ((used // N) % N) / (2 * n)
    X[:, 2] = (used // (N * N)) / (2 * n)
    return used.astype(np.int32), X, tets


def make_case(oracle, ct, n, nranks=1):
    if ct == TET10:
        gid, X, conn = tet10_mesh(n[0])
        job = oracle.Job.from_mesh(TET10, gid, X, np.arange(len(conn), dtype=np.int32), conn)
    else:
        job = oracle.Job.cuboid(ct, n, nranks=nranks)
    return job


def setup_rank(fc, job, r, ct, form, mat, young, nue, use_set_graph=False):
    o = job.rank(r)
    idx = {int(g): i for i, g in enumerate(job.node_gid)}
    col = np.array([idx[int(g)] for g in o["col_node_gid"]])
    eidx = {int(g): i for i, g in enumerate(job.ele_gid)}
    erows = np.array([eidx[int(g)] for g in o["col_eles"]], dtype=np.int64)
    lid = {int(g): i for i, g in enumerate(o["col_node_gid"])}
    conn = np.vectorize(lid.get)(job.conn[erows]).astype(np.int32) if len(erows) else np.zeros((0, fc.NEN[ct]), np.int32)
    s = fc.Solid(0, r, job.nranks)
    s.set_mesh(ct, form, job.X[col], o["col_node_gid"], job.node_owner[col], conn, o["col_eles"])
    s.set_material(mat, young, nue)
    if use_set_graph:
        s.set_graph(o["rowptr"], o["colind"])
    else:
        s.build_graph()
    return s, o, col


def kernel_variants(fc, ct, form, mat):
    """kernel-selection flags to cover: the default kernel of the combination plus its alternatives"""
    if ct == HEX8 and form == DISP and mat == NH:
        # DMMA closed form: per-element reductions (default) and owner-computed rows (lattice meshes, assembled K only);
        # FMA closed form; generic
        return [0, fc.FC_EVAL_OWNER_COMPUTED, fc.FC_EVAL_CLOSED_FORM_FMA, fc.FC_EVAL_FORCE_GENERIC]
    if ct == HEX8 and ((form == FBAR and mat == NH) or (form == DISP and mat == SVK)):
        return [0, fc.FC_EVAL_FORCE_GENERIC]  # tensor-core closed form (k_hex8_mma_ext), generic
    if ct == TET10 and form == DISP and mat == NH:
        return [0, fc.FC_EVAL_DMMA_TILES, fc.FC_EVAL_FORCE_GENERIC]  # warp-per-element closed form (k_tet10_nh), tensor-core tiles, generic
    if ct == HEX27 and form == DISP and mat == NH:
        return [0, fc.FC_EVAL_FORCE_GENERIC]  # tensor-core closed form (k_nh_mma_gen), generic
    return [0]


def rel(a, b):
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)


def rel_per_row(K, K_ref, rowptr):
    """max over the rows of |K - K_ref|_inf / |K_ref|_inf of that row"""
    starts = np.asarray(rowptr[:-1], np.int64)
    keep = starts < len(K_ref)
    rmax = np.maximum.reduceat(np.abs(K_ref), starts[keep])
    emax = np.maximum.reduceat(np.abs(K - K_ref), starts[keep])
    return float(np.max(emax / np.maximum(rmax, 1e-300)))


@pytest.mark.parametrize("ct,form,mat,nue,n", [
    (HEX8, DISP, NH, 0.25, (6, 5, 4)),
    (HEX8, DISP, SVK, 0.3, (5, 4, 3)),
    (HEX8, FBAR, NH, 0.25, (4, 4, 3)),
    (HEX8, FBAR, NH, 0.499, (3, 3, 3)),
    (HEX8, FBAR, SVK, 0.3, (3, 3, 3)),
    (HEX27, DISP, NH, 0.25, (3, 2, 2)),
    (HEX27, DISP, SVK, 0.3, (2, 2, 2)),
    (TET10, DISP, NH, 0.25, (2, 2, 2)),
    (TET10, DISP, SVK, 0.3, (2, 2, 2)),
])
def test_element_level_parity(fc, oracle, ct, form, mat, nue, n):
    """fe, Ke of every element vs oracle.element_evaluate (Solid::evaluate seam)."""
    job = make_case(oracle, ct, n)
    s, o, col = setup_rank(fc, job, 0, ct, form, mat, 10.0, nue)
    h = 1.0 / max(n)
    u = oracle.smooth_displacement(job.X, h, amp=0.01)
    nidx = {int(g): i for i, g in enumerate(job.node_gid)}
    conn = np.vectorize(nidx.get)(job.conn)
    bad, fe_ref, Ke_ref = oracle.element_evaluate_batch(ct, form, mat, 10.0, nue, job.X[conn], u[conn])
    assert bad == 0
    ucol = np.ascontiguousarray(u[col].reshape(-1))
    variants = kernel_variants(fc, ct, form, mat)
    eidx = {int(g): i for i, g in enumerate(job.ele_gid)}
    order = np.array([eidx[int(g)] for g in o["col_eles"]])
    for flags in variants:
        fe, Ke = s.evaluate_elements(ucol, want_K=True, flags=flags)
        for e in range(len(order)):
            kr, fr = Ke_ref[order[e]], fe_ref[order[e]]
            assert np.max(np.abs(Ke[e] - kr)) <= TOL * np.max(np.abs(kr)), (flags, e, rel(Ke[e], kr))
            assert np.max(np.abs(fe[e] - fr)) <= TOL * max(np.max(np.abs(fr)), np.max(np.abs(kr)) * h), (flags, e)
    s.close()


@pytest.mark.parametrize("ct,form,n,summands,with_fibers", [
    (HEX8, DISP, (4, 3, 3), [(1, [10.0, 0.3, 0, 0])], False),                                  # ELAST_CoupSVK alone
    (HEX8, DISP, (4, 3, 3), [(0, [10.0, 0.25, 0, 0]), (1, [4.0, 0.1, 0, 0])], False),           # CoupNeoHooke + CoupSVK
    (HEX8, FBAR, (3, 3, 3), [(0, [10.0, 0.25, 0, 0]), (2, [100.0, 1.0, 0.0, 1.0])], True),      # the material of solid_ele_hex8_Standard_fbar.dat
    (HEX8, DISP, (3, 3, 2), [(0, [10.0, 0.3, 0, 0]), (2, [50.0, 2.0, 5.0, 0.5])], True),        # fibres active in compression too
    (HEX27, DISP, (2, 2, 2), [(1, [10.0, 0.3, 0, 0]), (2, [20.0, 1.0, 0.0, 1.0])], True),
    (TET10, DISP, (2, 2, 2), [(0, [10.0, 0.25, 0, 0]), (2, [20.0, 1.0, 0.0, 1.0])], True),
    # fibres given at the nodes (INIT 3), interpolated to the Gauss points by the library
    (HEX8, DISP, (3, 3, 2), [(0, [10.0, 0.25, 0, 0]), (2, [100.0, 1.0, 0.0, 1.0])], "nodal"),
    (HEX27, DISP, (2, 2, 2), [(0, [10.0, 0.25, 0, 0]), (2, [20.0, 1.0, 5.0, 0.5])], "nodal"),
    (TET10, DISP, (2, 2, 2), [(0, [10.0, 0.25, 0, 0]), (2, [100.0, 1.0, 0.0, 1.0])], "nodal"),
])
def test_elasthyper_summand_list_parity(fc, oracle, ct, form, n, summands, with_fibers):
    """MAT_ElastHyper with a summand list (CoupNeoHooke, CoupSVK, CoupAnisoExpo with element fibres): fe, Ke of every element
    and the Gauss-point Cauchy stresses against the oracle's restatement of the summand loop."""
    job = make_case(oracle, ct, n)
    s, o, col = setup_rank(fc, job, 0, ct, form, NH, 10.0, 0.25)
    rng = np.random.default_rng(11)
    ne = len(job.conn)
    nodal = with_fibers == "nodal"
    fibers = rng.standard_normal((ne, 3)) * 3.0 if with_fibers and not nodal else None  # not normalised: the library normalises like the reference
    s.set_material_elasthyper(summands, fibers)
    node_fibers = None
    if nodal:
        node_fibers = rng.standard_normal((len(job.X), 3)) + np.array([2.5, 0.0, 0.0])  # no near-zero interpolants
        s.set_nodal_fibers(node_fibers[col])
    h = 1.0 / max(n)
    u = oracle.smooth_displacement(job.X, h, amp=0.03)
    nidx = {int(g): i for i, g in enumerate(job.node_gid)}
    conn = np.vectorize(nidx.get)(job.conn)
    ucol = np.ascontiguousarray(u[col].reshape(-1))
    fe, Ke = s.evaluate_elements(ucol, want_K=True)
    gp, _ = s.evaluate_stress(ucol, fc.FC_STRESS_CAUCHY, want_nodal=False)
    eidx = {int(g): i for i, g in enumerate(job.ele_gid)}
    order = np.array([eidx[int(g)] for g in o["col_eles"]])
    for e in range(ne):
        j = order[e]
        fib = oracle.project_fibers_to_gauss_points(ct, node_fibers[conn[j]]) if nodal else (None if fibers is None else fibers[e])
        err, fr, kr, gr = oracle.element_evaluate_elasthyper(ct, form, summands, fib, job.X[conn[j]], u[conn[j]], 1)
        assert err == 0
        assert np.max(np.abs(Ke[e] - kr)) <= TOL * np.max(np.abs(kr)), (e, rel(Ke[e], kr))
        assert np.max(np.abs(fe[e] - fr)) <= TOL * max(np.max(np.abs(fr)), np.max(np.abs(kr)) * h), e
        assert np.max(np.abs(gp[e] - gr)) <= TOL * np.max(np.abs(gr)), e
    s.close()


@pytest.mark.parametrize("ct,form,mat,nue,n,nranks", [
    (HEX8, DISP, NH, 0.25, (10, 10, 10), 1),
    (HEX8, DISP, NH, 0.25, (7, 6, 5), 4),
    (HEX8, DISP, SVK, 0.3, (10, 10, 10), 2),
    (HEX8, FBAR, NH, 0.45, (6, 6, 6), 3),
    (HEX27, DISP, NH, 0.25, (4, 3, 3), 2),
    (TET10, DISP, NH, 0.25, (3, 3, 3), 1),
])
def test_assembled_parity(fc, oracle, ct, form, mat, nue, n, nranks):
    """assembled f and K of every rank vs the oracle's owner-row assembly, same CSR positions."""
    job = make_case(oracle, ct, n, nranks)
    job.set_problem(form, mat, 10.0, nue)
    h = 1.0 / max(n)
    u = oracle.smooth_displacement(job.X, h, amp=0.01)
    job.set_u_nodal(u)
    bad, _ = job.evaluate(oracle.ACTION_NLNSTIFF, max_threads=4)
    assert bad == 0
    for r in range(job.nranks):
        s, o, col = setup_rank(fc, job, r, ct, form, mat, 10.0, nue, use_set_graph=(r % 2 == 1))
        rp, ci = s.get_graph()
        assert np.array_equal(rp, o["rowptr"]) and np.array_equal(ci, o["colind"])  # sparsity pattern bit-exact
        ucol = np.ascontiguousarray(u[col].reshape(-1))
        K_ref, f_ref = job.get_K(r, o["nnz"]), job.get_f(r, o["n_row_dof"])
        variants = kernel_variants(fc, ct, form, mat)
        for flags in variants:
            K, f = np.full(o["nnz"], 7.0), np.full(o["n_row_dof"], 7.0)
            cnt, gid = s.evaluate(fc.FC_ACTION_NLNSTIFF, ucol, f, K, flags=flags | fc.FC_EVAL_ZERO_FIRST)
            assert (cnt, gid) == (0, -1)
            if o["nnz"]:
                assert rel_per_row(K, K_ref, o["rowptr"]) <= TOL, (r, flags, rel_per_row(K, K_ref, o["rowptr"]))
                assert np.max(np.abs(f - f_ref)) <= TOL * max(np.max(np.abs(f_ref)), np.max(np.abs(K_ref)) * h)
            # += semantics (caller owns and pre-zeroes, 4C_structure_new_model_evaluator_structure.cpp:136-152)
            cnt, gid = s.evaluate(fc.FC_ACTION_NLNSTIFF, ucol, f, K, flags=flags)
            if o["nnz"]:
                assert rel_per_row(K, 2 * K_ref, o["rowptr"]) <= 2 * TOL
            # residual only (struct_calc_internalforce)
            f2 = np.zeros(o["n_row_dof"])
            s.evaluate(fc.FC_ACTION_INTERNALFORCE, ucol, f2, None, flags=flags)
            if o["nnz"]:
                assert np.max(np.abs(f2 - f_ref)) <= TOL * max(np.max(np.abs(f_ref)), np.max(np.abs(K_ref)) * h)
        s.close()
