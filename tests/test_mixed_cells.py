"""Discretizations with several cell types and element technologies - the reference's soweg6_disp_patchtest (4 hex8 + 6 wedge6,
St.Venant-Kirchhoff) and prestress_new (tet4, tet10, hex8, hex8 with F-bar, hex20, hex27, all PRESTRESS_TECH mulf, seven load
steps of which the first three end with a prestress update): one library context per (cell type, technology), all on the union
graph of the whole mesh (fc_solid_set_graph) and all adding into the same CSR value array and residual - the reference's Discretization::evaluate loops over elements of any type and assembles into one
matrix (4C_fem_discretization_evaluate.cpp:83, 4C_linalg_sparsematrix.cpp:426-576).  Checked against the reference's listed
result, oracle-driven and CUDA-driven."""
import copy

import numpy as np
import pytest
import structure_driver as sd

CASES = {"soweg6_disp_patchtest": 2, "prestress_new": 6}  # case -> number of (cell type, technology) groups


def split(oracle, case):
    """one Problem per (cell type, element options): all nodes, the elements of that group"""
    fx = sd.load(case)
    keys = []
    for e in fx["elements"]:
        if (e["type"], e["opts"]) not in keys:
            keys.append((e["type"], e["opts"]))
    parts = []
    for typ, opts in keys:
        f = copy.deepcopy(fx)
        f["elements"] = [e for e in fx["elements"] if (e["type"], e["opts"]) == (typ, opts)]
        f["part_of_mixed_mesh"] = True
        parts.append(sd.Problem(f, oracle.node_param_coords))
    assert len(parts) == CASES[case] and len({p.n_dof for p in parts}) == 1
    return parts


def union_graph(parts):
    n = len(parts[0].ids)
    adj = [set() for _ in range(n)]
    for p in parts:
        for en in p.conn:
            for a in en:
                adj[int(a)].update(int(b) for b in en)
    rowptr, colind = [0], []
    for i in range(n):
        cols = sorted(adj[i])
        for d in range(3):
            colind += [3 * c + k for c in cols for k in range(3)]
            rowptr.append(len(colind))
    return np.array(rowptr, np.int64), np.array(colind, np.int32)


def run_and_check(parts, assemble, update_prestress=None):
    p0 = parts[0]
    f0 = p0.f_ext
    p0.f_ext = lambda t: f0(t) + sum(p.f_ext(t) for p in parts[1:])  # a loaded face belongs to one element, of one group
    d, hist = sd.run(p0, assemble, update_prestress=update_prestress)
    err, n = p0.result_error(d)
    assert n >= 3 and err <= 1.0, (err, hist)
    return d, hist


def oracle_group(oracle, p):
    """oracle job of one group; its rows are the DOFs of the nodes its elements touch, placed into the mesh's numbering"""
    assert p.summands is None
    job = oracle.Job.from_mesh(p.ct, p.node_gid, p.X, p.ele_gid, p.node_gid[p.conn])
    job.set_problem(p.form, p.mat, p.young, p.nue)
    o = job.rank(0)
    rows, cols = o["row_dof_gid"], o["col_dof_gid"]

    def assemble(d):
        job.set_u_nodal(d.reshape(-1, 3))
        assert job.evaluate(oracle.ACTION_NLNSTIFF)[0] == 0
        f, K = np.zeros(p.n_dof), np.zeros((p.n_dof, p.n_dof))
        f[rows] = job.get_f(0, len(rows))
        vals = job.get_K(0, o["nnz"])
        for r in range(len(rows)):
            a, b = o["rowptr"][r], o["rowptr"][r + 1]
            K[rows[r], cols[o["colind"][a:b]]] = vals[a:b]
        return f, K
    return assemble, (lambda d: job.update_prestress(d.reshape(-1, 3)))


def oracle_solve(oracle, case):
    parts = split(oracle, case)
    for p in parts:
        p.neumann = sd.OracleNeumann(oracle, p)
    backends = [oracle_group(oracle, p) for p in parts]

    def assemble(d):
        out = [a(d) for a, _ in backends]
        return sum(o[0] for o in out), sum(o[1] for o in out)
    return run_and_check(parts, assemble, lambda d: [u(d) for _, u in backends])


@pytest.mark.parametrize("case", list(CASES))
def test_oracle(oracle, case):
    oracle_solve(oracle, case)


@pytest.mark.gpu
@pytest.mark.parametrize("case", list(CASES))
def test_cuda_contexts_share_one_matrix(fc, oracle, case):
    parts = split(oracle, case)
    rowptr, colind = union_graph(parts)
    n_dof = parts[0].n_dof
    ctxs = []
    for p in parts:
        s = fc.Solid(0)
        s.set_mesh(p.ct, p.form, p.X, p.node_gid, np.zeros(len(p.node_gid), np.int32), p.conn, p.ele_gid)
        s.set_material(p.mat, p.young, p.nue)
        n_row, n_col, nnz = s.set_graph(rowptr, colind)
        assert (n_row, nnz) == (n_dof, len(colind))
        p.neumann = sd.CudaNeumann(s)
        ctxs.append(s)

    def assemble(d):
        K, f = np.full(len(colind), np.nan), np.full(n_dof, np.nan)
        for k, s in enumerate(ctxs):  # the first call zeroes, the others add into the same arrays
            assert s.evaluate(fc.FC_ACTION_NLNSTIFF, d, f, K, flags=fc.FC_EVAL_ZERO_FIRST if k == 0 else 0) == (0, -1)
        return f, sd.csr_to_dense(rowptr, colind, K, n_dof)
    d, hist = run_and_check(parts, assemble, lambda dd: [s.update_prestress(dd) for s in ctxs if s.formulation in (3, 6)])
    # the same Newton history as the oracle-driven solve
    d2, hist2 = oracle_solve(oracle, case)
    assert [len(h) for h in hist] == [len(h) for h in hist2] and np.max(np.abs(d - d2)) <= 1e-10 * np.max(np.abs(d2))
    for s in ctxs:
        s.close()
