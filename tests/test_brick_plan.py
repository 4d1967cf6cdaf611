"""Brick plan of the owner-computed hex8 assembly (fc_brick_plan_*, host integer code) and the L2 sweep order.

Every owned node is a row of exactly one unit, every (element, owned local node) pair is accumulated exactly once, the
12 elements of a colour round share no node, and replaying the kernel's data flow (accumulator window of two node planes,
rows flushed when their plane is complete) reproduces the reference's per-element assembly (SparseMatrix::assemble,
4C_linalg_sparsematrix.cpp:471-575) entry by entry."""
import numpy as np
import pytest


def node_graph(conn, n_row):
    """owned node -> ascending column nodes coupled through any element"""
    adj = [set() for _ in range(n_row)]
    for en in conn:
        for a in en:
            if a < n_row:
                adj[a].update(int(b) for b in en)
    nodeptr = np.zeros(n_row + 1, np.int64)
    nodeadj = []
    for n in range(n_row):
        s = sorted(adj[n])
        nodeadj += s
        nodeptr[n + 1] = len(nodeadj)
    return nodeptr, np.array(nodeadj, np.int32)


def direct_scatter(conn, n_row, nodeptr, nodeadj, Ke):
    """owned rows only, node-blocked CSR (row 3a+i starts at 9*nodeptr[a] + i*3*L)"""
    K = np.zeros(9 * nodeptr[n_row])
    for e, en in enumerate(conn):
        for a in range(8):
            na = en[a]
            if na >= n_row:
                continue
            row = nodeadj[nodeptr[na]:nodeptr[na + 1]]
            L = len(row)
            for b in range(8):
                pos = int(np.searchsorted(row, en[b]))
                assert row[pos] == en[b]
                for i in range(3):
                    o = 9 * nodeptr[na] + i * 3 * L + 3 * pos
                    K[o:o + 3] += Ke[e, a, i, b, :]
    return K


def plan_scatter(plan, conn, n_row, nodeptr, nodeadj, Ke):
    """the brick kernel's data flow on the host: window of 2 x 35 accumulator rows, flush of plane t-1 after layer t"""
    NL, NE = 35, 48
    K = np.full(9 * nodeptr[n_row], np.nan)
    for u in range(plan["n_units"]):
        P = int(plan["unit_planes"][u])
        nodes = plan["node_tab"][plan["unit_node_off"][u]:plan["unit_node_off"][u] + P * NL].reshape(P, NL)
        eo = plan["unit_ele_off"][u]
        eles = plan["ele_tab"][eo:eo + (P + 1) * NE].reshape(P + 1, NE)
        rows = plan["rowidx_tab"][eo:eo + (P + 1) * NE].reshape(P + 1, NE, 8)
        win = [None] * (2 * NL)       # per window row: dict column node -> 3x3 block
        win_node = [-1] * (2 * NL)
        for t in range(P + 1):
            if t < P:
                for l in range(NL):
                    w = (t & 1) * NL + l
                    assert win[w] is None  # the previous occupant was flushed
                    win_node[w] = int(nodes[t, l])
                    win[w] = {} if nodes[t, l] >= 0 else None
            for colour in range(4):
                touched = set()
                for wi in range(12):
                    k = 12 * colour + wi
                    e = int(eles[t, k])
                    if e < 0:
                        assert np.all(rows[t, k] == 0xFF)
                        continue
                    en = conn[e]
                    assert not (touched & set(en.tolist())), "elements of one colour round share a node"
                    touched |= set(en.tolist())
                    assert np.any(rows[t, k] != 0xFF)
                    for a in range(8):
                        w = int(rows[t, k, a])
                        if w == 0xFF:
                            continue
                        assert win_node[w] == en[a] and en[a] < n_row
                        for b in range(8):
                            win[w].setdefault(int(en[b]), np.zeros((3, 3)))
                            win[w][int(en[b])] += Ke[e, a, :, b, :]
            if t >= 1:
                for l in range(NL):
                    w = ((t - 1) & 1) * NL + l
                    na = win_node[w]
                    if na < 0:
                        continue
                    row = nodeadj[nodeptr[na]:nodeptr[na + 1]]
                    L = len(row)
                    base = 9 * nodeptr[na]
                    assert np.all(np.isnan(K[base:base + 9 * L])), "a row is written twice"
                    K[base:base + 9 * L] = 0.0
                    for nb, blk in win[w].items():
                        pos = int(np.searchsorted(row, nb))
                        for i in range(3):
                            K[base + i * 3 * L + 3 * pos:base + i * 3 * L + 3 * pos + 3] = blk[i]
                    win[w] = None
        assert all(x is None for x in win)
    return K


@pytest.mark.parametrize("n,nranks,rank,chunk", [((8, 4, 4), 1, 0, 0), ((5, 3, 2), 1, 0, 2), ((6, 5, 4), 2, 1, 0), ((7, 6, 5), 4, 2, 3), ((1, 1, 1), 1, 0, 0),
                                                 ((4, 4, 4), 8, 7, 0), ((9, 2, 3), 3, 0, 4), ((16, 9, 7), 1, 0, 5), ((3, 12, 11), 2, 0, 0)])
def test_plan_reproduces_per_element_assembly(fc, n, nranks, rank, chunk):
    g = fc.Grid(fc.FC_HEX8, n, nranks=nranks, rank=rank)
    conn, n_row = g.conn, g.n_row_nodes
    nodeptr, nodeadj = node_graph(conn, n_row)
    plan = fc.capi.brick_plan(g.X, conn, n_row, chunk=chunk)
    assert plan is not None
    # every owned node is a row of exactly one unit
    nt = plan["node_tab"]
    assert sorted(nt[nt >= 0].tolist()) == list(range(n_row))
    # every (element, owned local node) pair exactly once
    seen = np.zeros((len(conn), 8), np.int32)
    valid = plan["ele_tab"] >= 0
    assert plan["n_element_instances"] == int(valid.sum())
    for e, r in zip(plan["ele_tab"][valid], plan["rowidx_tab"][valid]):
        seen[e] += (r != 0xFF)
    assert np.array_equal(seen, (conn < n_row).astype(np.int32))
    # interior units first: none of their elements has a ghost node
    for u in range(plan["n_units"]):
        P = int(plan["unit_planes"][u])
        eo = plan["unit_ele_off"][u]
        el = plan["ele_tab"][eo:eo + (P + 1) * 48]
        ghost = bool(np.any(conn[el[el >= 0]] >= n_row))
        assert ghost == (u >= plan["n_interior_units"])
    rng = np.random.default_rng(5)
    Ke = rng.standard_normal((len(conn), 8, 3, 8, 3))
    K_ref = direct_scatter(conn, n_row, nodeptr, nodeadj, Ke)
    K_plan = plan_scatter(plan, conn, n_row, nodeptr, nodeadj, Ke)
    assert np.allclose(K_plan, K_ref, rtol=0, atol=1e-12)


def test_non_lattice_mesh_has_no_plan(fc):
    g = fc.Grid(fc.FC_HEX8, (3, 3, 3))
    X = g.X.copy()
    X[21] += [0.05, 0.02, 0.0]  # one node off the rectilinear grid
    assert fc.capi.brick_plan(X, g.conn, g.n_row_nodes) is None
    conn = g.conn.copy()
    conn[3, 5] = conn[3, 4]     # collapsed edge: the element does not span a lattice cell
    assert fc.capi.brick_plan(g.X, conn, g.n_row_nodes) is None
    assert fc.capi.brick_plan(g.X, g.conn, g.n_row_nodes, max_row_nodes=40) is None  # rows longer than the 27-block window
    # a rectilinear grid with non-uniform spacing is a lattice
    Y = g.X.copy()
    Y[:, 1] = Y[:, 1] ** 2
    assert fc.capi.brick_plan(Y, g.conn, g.n_row_nodes) is not None


def test_structured_units(fc):
    """C2-like proportions: 7 x 5 node lines per unit, 48 element columns, element instances / elements = the redundancy
    DESIGN.md quotes"""
    g = fc.Grid(fc.FC_HEX8, (28, 21, 15))
    plan = fc.capi.brick_plan(g.X, g.conn, g.n_row_nodes, chunk=29)
    assert plan["n_units"] == 4 * 4 and plan["chunk"] == 29  # 22 x 16 node lines -> 4 x 4 units, one chunk of 29 planes
    assert plan["n_element_instances"] / g.n_col_eles < 1.45


def test_l2_sweep_order(fc):
    """large meshes: (y block, z, y, x) sweep; a permutation, blocks of W element rows, identity when a z layer fits"""
    g = fc.Grid(fc.FC_HEX8, (10, 12, 6))
    changed, perm = fc.capi.l2_sweep_order(g.X, g.conn, 40)  # W = max(4, 40 // 10) = 4 element rows per y block
    assert changed and sorted(perm.tolist()) == list(range(720))
    c = g.X[g.conn[perm]].mean(axis=1)
    iy = np.floor(c[:, 1] * 12).astype(int)
    iz = np.floor(c[:, 2] * 6).astype(int)
    ix = np.floor(c[:, 0] * 10).astype(int)
    key = ((iy // 4) * 6 + iz) * 12 * 10 + iy * 10 + ix
    assert np.all(np.diff(key) > 0)
    assert fc.capi.l2_sweep_order(g.X, g.conn, 120)[0] is False     # a layer (120 elements) fits
    assert fc.capi.l2_sweep_order(g.X, g.conn, 10 ** 6)[0] is False  # small mesh
