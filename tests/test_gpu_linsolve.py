"""SURVEY 8f F1 on the device: Dirichlet application (SparseMatrix::apply_dirichlet, apply_dirichlet_to_system), y = K x
and the Krylov hand-off (CG / BiCGStab with node-block Jacobi), K resident in HBM.  Checked against scipy on the same
CSR values, and by replaying the reference's solve-level golden vectors with every linear solve done on the device."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla
import structure_driver as sd

pytestmark = pytest.mark.gpu


def setup(fc, n=(6, 5, 4), form=0, mat=1, nue=0.25):
    import torch
    g = fc.Grid(fc.FC_HEX8, n)
    s = fc.Solid(0)
    s.set_mesh(fc.FC_HEX8, form, g.X, g.node_gid, g.node_owner, g.conn, g.ele_gid)
    s.set_material(mat, 10.0, nue)
    n_row, n_col, nnz = s.build_graph()
    rowptr, colind = s.get_graph()
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(3)
    u = torch.from_numpy(1e-3 * rng.standard_normal(n_col)).to(dev)
    K = torch.empty(nnz, dtype=torch.float64, device=dev)
    f = torch.empty(n_row, dtype=torch.float64, device=dev)
    assert s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
    fixed = np.flatnonzero(np.repeat(g.X[:, 0] < 1e-12, 3)).astype(np.int32)  # x- face clamped
    return s, g, rowptr, colind, K, f, fixed, dev, rng


def test_spmv_and_dirichlet_application_match_host(fc):
    import torch
    s, g, rowptr, colind, K, f, fixed, dev, rng = setup(fc)
    n = len(rowptr) - 1
    A = sp.csr_matrix((K.cpu().numpy(), colind, rowptr), shape=(n, n))
    x = rng.standard_normal(n)
    y = torch.empty(n, dtype=torch.float64, device=dev)
    s.spmv(K, torch.from_numpy(x).to(dev), y)
    ref = A @ x
    assert np.max(np.abs(y.cpu().numpy() - ref)) <= 1e-13 * np.max(np.abs(ref))
    # apply_dirichlet: rows zeroed, 1.0 on the diagonal (4C_linalg_sparsematrix.cpp:1066-1092); without diagonalblock: 0
    s.set_dirichlet_toggle(fixed)
    for diag in (True, False):
        K2 = K.clone()
        s.apply_dirichlet(K2, diagonalblock=diag)
        B = A.tolil(copy=True)
        for r in fixed:
            B.data[r] = [0.0] * len(B.rows[r])
            if diag:
                B[r, r] = 1.0
        B = sp.csr_matrix(B)
        got = sp.csr_matrix((K2.cpu().numpy(), colind, rowptr), shape=(n, n))
        assert abs(got - B).max() == 0.0
    # apply_dirichlet_to_system: x = b = dbcval on the rows, untouched elsewhere (4C_linalg_utils_sparse_algebra_assemble.cpp:126-140)
    xv, bv, dv = (torch.from_numpy(rng.standard_normal(n)).to(dev) for _ in range(3))
    x0, b0 = xv.clone(), bv.clone()
    s.apply_dirichlet_to_system(xv, bv, dv)
    free = np.setdiff1d(np.arange(n), fixed)
    assert torch.equal(xv[fixed], dv[fixed]) and torch.equal(bv[fixed], dv[fixed])
    assert torch.equal(xv[free], x0[free]) and torch.equal(bv[free], b0[free])
    s.apply_dirichlet_to_system(None, bv, None)
    assert float(bv[fixed].abs().max()) == 0.0
    s.close()


@pytest.mark.parametrize("method,form,nue", [(0, 0, 0.25), (1, 0, 0.25), (1, 1, 0.45)])
def test_linear_solve_matches_direct_solve(fc, method, form, nue):
    """Newton-type system: Dirichlet rows applied, right-hand side zero there (CG) / prescribed values (BiCGStab)."""
    import torch
    s, g, rowptr, colind, K, f, fixed, dev, rng = setup(fc, form=form, nue=nue)
    n = len(rowptr) - 1
    s.set_dirichlet_toggle(fixed)
    s.apply_dirichlet(K)
    b = torch.from_numpy(rng.standard_normal(n)).to(dev)
    dbc = torch.zeros(n, dtype=torch.float64, device=dev)
    if method == 1:
        dbc[fixed] = torch.from_numpy(0.1 * rng.standard_normal(len(fixed))).to(dev)
    x = torch.zeros(n, dtype=torch.float64, device=dev)
    s.apply_dirichlet_to_system(x, b, dbc)
    it, rr = s.linear_solve(K, b, x, method=method, rel_tol=1e-12, max_iter=5000)
    assert 0 < it < 5000 and rr <= 1e-12
    A = sp.csc_matrix(sp.csr_matrix((K.cpu().numpy(), colind, rowptr), shape=(n, n)))
    ref = spla.spsolve(A, b.cpu().numpy())
    assert np.max(np.abs(x.cpu().numpy() - ref)) <= 1e-8 * np.max(np.abs(ref))
    assert torch.equal(x[fixed], dbc[fixed]) or float((x[fixed] - dbc[fixed]).abs().max()) <= 1e-14
    # zero right-hand side: x = 0 in zero iterations
    it, rr = s.linear_solve(K, torch.zeros_like(b), x, method=method)
    assert it == 0 and float(x.abs().max()) == 0.0
    s.close()


class DevCsr:
    """CSR values on the device in the layout of the context's graph: a*A + b*B and A @ x for the mock time integrator"""

    def __init__(self, s, vals):
        self.s, self.v = s, vals

    def __mul__(self, a):
        return DevCsr(self.s, self.v * a)

    __rmul__ = __mul__

    def __add__(self, o):
        return DevCsr(self.s, self.v + o.v)

    def __matmul__(self, x):
        import torch
        y = torch.empty(len(x), dtype=torch.float64, device=self.v.device)
        self.s.spmv(self.v, torch.from_numpy(np.ascontiguousarray(x)).to(self.v.device), y)
        return y.cpu().numpy()


CASES = [
    ("solid_ele_hex8_Standard_eas_none_stressout", 0),
    ("solid_ele_hex8_Standard_eas_none_dynamic", 0),
    ("solid_ele_tet10_Standard_stressout", 0),
    ("solid_ele_tet10_Standard_dynamic", 0),
    ("solid_runtime_ghosting_output", 0),
    ("solid_ele_hex8_Standard_eas_none_stressout", 1),
    ("solid_ele_tet10_Standard_stressout", 1),
]
# sohex8fbar_cooks_nl_new_struc (F-bar, nu = 0.4999, condition ~1e8) is out of reach of a node-block-Jacobi Krylov method:
# BiCGStab stagnates; the reference solves it with a direct solver (UMFPACK).  F-bar tangents at moderate nu are covered by
# test_linear_solve_matches_direct_solve.  BiCGStab with this preconditioner is also erratic on the 640-element large-deformation
# SVK case (solid_runtime_ghosting_output: converges or stalls at 1e-7 depending on the last bits of K), so that case runs CG.


@pytest.mark.parametrize("case,method", CASES)
def test_reference_golden_solves_with_device_linear_solver(fc, oracle, case, method):
    """the reference's RESULT DESCRIPTION values reached with K (and M) never leaving the device: element evaluation,
    Dirichlet application and the Krylov solve are all library calls on device buffers"""
    import torch
    p = sd.Problem(sd.load(case), oracle.node_param_coords)
    dev = torch.device("cuda", 0)
    s = fc.Solid(0)
    s.set_mesh(p.ct, p.form, p.X, p.node_gid, np.zeros(len(p.node_gid), np.int32), p.conn, p.ele_gid)
    s.set_material(p.mat, p.young, p.nue)
    n_row, n_col, nnz = s.build_graph()
    s.set_dirichlet_toggle(np.asarray(p.fixed, np.int32))
    iters = []

    def assemble(d):
        K = torch.empty(nnz, dtype=torch.float64, device=dev)
        f = torch.empty(n_row, dtype=torch.float64, device=dev)
        assert s.evaluate(fc.FC_ACTION_NLNSTIFF, torch.from_numpy(np.ascontiguousarray(d)).to(dev), f, K, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
        return f.cpu().numpy(), DevCsr(s, K)

    def mass():
        M = torch.empty(nnz, dtype=torch.float64, device=dev)
        s.evaluate_mass(p.dens, M)
        return DevCsr(s, M)

    def solve(A, rhs, fixed):
        J = A.v.clone()
        s.apply_dirichlet(J)
        b = torch.from_numpy(np.ascontiguousarray(rhs)).to(dev)
        x = torch.zeros_like(b)
        s.apply_dirichlet_to_system(x, b, None)
        # BiCGStab's recursive residual drifts from the true one below ~1e-12: ask it for 1e-11, CG for 1e-13
        it, rr = s.linear_solve(J, b, x, method=method, rel_tol=1e-13 if method == 0 else 1e-11, max_iter=20000, check=False)
        assert rr <= 1e-8, (case, it, rr)
        iters.append(it)
        return x.cpu().numpy()

    d, hist = sd.run(p, assemble, mass, solve=solve)
    err, n = p.result_error(d)
    assert n >= 1 and err <= 1.0, (case, err, hist, iters)
    s.close()


@pytest.mark.parametrize("n,levels", [((16, 16, 16), 4), ((24, 12, 9), 4), ((25, 25, 25), 5), ((32, 32, 32), 5)])
def test_multigrid_preconditioned_cg(fc, n, levels):
    """FC_SOLVER_CG_MG: geometric multigrid V-cycle (coarse tangents re-discretised by the library's own element kernels at
    the injected state) against node-block-Jacobi CG: same solution, an iteration count that does not grow with the mesh."""
    import torch
    dev = torch.device("cuda", 0)
    g = fc.Grid(fc.FC_HEX8, n)
    s = fc.Solid(0)
    s.set_mesh(fc.FC_HEX8, 0, g.X, g.node_gid, g.node_owner, g.conn, g.ele_gid)
    s.set_material(1, 10.0, 0.3)
    n_row, n_col, nnz = s.build_graph()
    assert s.mg_setup() == levels
    # a deformed state, so that the coarse operators are evaluated at an injected displacement
    u = torch.from_numpy((0.05 * np.sin(2.0 * g.X[:, [1, 2, 0]]) + 0.02 * g.X).reshape(-1)).to(dev)
    K = torch.empty(nnz, dtype=torch.float64, device=dev)
    f = torch.empty(n_row, dtype=torch.float64, device=dev)
    assert s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
    fixed = np.flatnonzero(np.repeat(g.X[:, 0] < 1e-12, 3)).astype(np.int32)
    s.set_dirichlet_toggle(fixed)
    s.apply_dirichlet(K)
    s.mg_update(u)
    b = np.zeros(n_row)
    loaded = np.flatnonzero(g.X[:, 0] > 1.0 - 1e-12)
    b[3 * loaded + 2] = 0.02 / len(loaded)
    b[3 * loaded + 1] = -0.01 / len(loaded)
    b = torch.from_numpy(b).to(dev)
    x_mg = torch.zeros(n_row, dtype=torch.float64, device=dev)
    s.apply_dirichlet_to_system(x_mg, b, None)
    it_mg, rr_mg = s.linear_solve(K, b, x_mg, method=fc.FC_SOLVER_CG_MG, rel_tol=1e-10, max_iter=200)  # default: FP32 cycle matrices
    s.mg_set_precision(False)
    with pytest.raises(fc.FourCError):  # the copies / estimates belong to an update
        s.linear_solve(K, b, x_mg.clone(), method=fc.FC_SOLVER_CG_MG)
    s.mg_update(u)
    x_64 = torch.zeros(n_row, dtype=torch.float64, device=dev)
    it_64, rr_64 = s.linear_solve(K, b, x_64, method=fc.FC_SOLVER_CG_MG, rel_tol=1e-10, max_iter=200)
    assert rr_64 <= 1e-10 and abs(it_64 - it_mg) <= 2 and float((x_64 - x_mg).abs().max()) <= 1e-7 * float(x_mg.abs().max())
    s.mg_set_precision(True)
    s.mg_update(u)  # second update of the level: the eigenvalue estimate starts from the stored iterate
    x_w = torch.zeros(n_row, dtype=torch.float64, device=dev)
    it_w, _ = s.linear_solve(K, b, x_w, method=fc.FC_SOLVER_CG_MG, rel_tol=1e-10, max_iter=200)
    assert abs(it_w - it_mg) <= 2
    # a damping beyond 2 / lambda_max makes the cycle indefinite: the solver cuts it and restarts instead of giving up
    s.mg_set_damping(1.95)
    s.mg_update(u)
    x_d = torch.zeros(n_row, dtype=torch.float64, device=dev)
    it_d, rr_d = s.linear_solve(K, b, x_d, method=fc.FC_SOLVER_CG_MG, rel_tol=1e-10, max_iter=300)
    assert rr_d <= 1e-10 and float((x_d - x_mg).abs().max()) <= 1e-7 * float(x_mg.abs().max())
    s.mg_set_damping(4.0 / 3.0)
    s.mg_update(u)
    x_j = torch.zeros(n_row, dtype=torch.float64, device=dev)
    it_j, rr_j = s.linear_solve(K, b, x_j, method=fc.FC_SOLVER_CG, rel_tol=1e-10, max_iter=20000)
    print(f"n={n}: CG-MG {it_mg} iterations (rel. residual {rr_mg:.1e}), node-block Jacobi CG {it_j}")
    assert rr_mg <= 1e-10 and it_mg <= 40 and it_mg < it_j / 3
    assert float((x_mg - x_j).abs().max()) <= 1e-7 * float(x_j.abs().max())
    # the solve needs a fresh fc_solid_mg_update after the toggle changed
    s.set_dirichlet_toggle(fixed[:-3])
    with pytest.raises(fc.FourCError):
        s.linear_solve(K, b, x_mg, method=fc.FC_SOLVER_CG_MG)
    s.close()
