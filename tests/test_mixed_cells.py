"""A discretization with two cell types (the reference's soweg6_disp_patchtest: 4 hex8 + 6 wedge6, St.Venant-Kirchhoff): one
library context per cell type, both on the union graph of the whole mesh (fc_solid_set_graph) and both adding into the same
CSR value array and residual - the reference's Discretization::evaluate loops over elements of any type and assembles into one
matrix (4C_fem_discretization_evaluate.cpp:83, 4C_linalg_sparsematrix.cpp:426-576).  Checked against the reference's listed
result, oracle-driven and CUDA-driven."""
import copy

import numpy as np
import pytest
import structure_driver as sd

CASE = "soweg6_disp_patchtest"


def split(oracle):
    """one Problem per cell type: all nodes, the elements of that type; the first one carries loads summed over both"""
    fx = sd.load(CASE)
    parts = []
    for typ in ("HEX8", "WEDGE6"):
        f = copy.deepcopy(fx)
        f["elements"] = [e for e in fx["elements"] if e["type"] == typ]
        f["part_of_mixed_mesh"] = True
        parts.append(sd.Problem(f, oracle.node_param_coords))
    assert (len(parts[0].conn), len(parts[1].conn)) == (4, 6) and parts[0].n_dof == parts[1].n_dof == 48
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


def run_and_check(parts, assemble):
    p0, p1 = parts
    f0 = p0.f_ext
    p0.f_ext = lambda t: f0(t) + p1.f_ext(t)  # a loaded face belongs to one element, of either type
    d, hist = sd.run(p0, assemble)
    err, n = p0.result_error(d)
    assert n == 3 and err <= 1.0, (err, hist)
    return d, hist


def test_oracle(oracle):
    import test_golden_reference_solves as T
    parts = split(oracle)
    for p in parts:
        p.neumann = sd.OracleNeumann(oracle, p)
    backends = [T.oracle_backend(oracle, p)[0] for p in parts]

    def assemble(d):
        out = [a(d) for a in backends]
        return out[0][0] + out[1][0], out[0][1] + out[1][1]
    run_and_check(parts, assemble)


@pytest.mark.gpu
def test_cuda_contexts_share_one_matrix(fc, oracle):
    parts = split(oracle)
    rowptr, colind = union_graph(parts)
    ctxs = []
    for p in parts:
        s = fc.Solid(0)
        s.set_mesh(p.ct, p.form, p.X, p.node_gid, np.zeros(len(p.node_gid), np.int32), p.conn, p.ele_gid)
        s.set_material(p.mat, p.young, p.nue)
        n_row, n_col, nnz = s.set_graph(rowptr, colind)
        assert (n_row, nnz) == (48, len(colind))
        p.neumann = sd.CudaNeumann(s)
        ctxs.append(s)

    def assemble(d):
        K, f = np.full(len(colind), np.nan), np.full(48, np.nan)
        assert ctxs[0].evaluate(fc.FC_ACTION_NLNSTIFF, d, f, K, flags=fc.FC_EVAL_ZERO_FIRST) == (0, -1)
        assert ctxs[1].evaluate(fc.FC_ACTION_NLNSTIFF, d, f, K) == (0, -1)  # += into the same arrays
        return f, sd.csr_to_dense(rowptr, colind, K, 48)
    d, hist = run_and_check(parts, assemble)
    # the same Newton history as the oracle-driven solve
    import test_golden_reference_solves as T
    parts2 = split(oracle)
    for p in parts2:
        p.neumann = sd.OracleNeumann(oracle, p)
    backends = [T.oracle_backend(oracle, p)[0] for p in parts2]
    d2, hist2 = run_and_check(parts2, lambda dd: tuple(x + y for x, y in zip(backends[0](dd), backends[1](dd))))
    assert len(hist[0]) == len(hist2[0]) and np.max(np.abs(d - d2)) <= 1e-11 * np.max(np.abs(d2))
    for s in ctxs:
        s.close()
