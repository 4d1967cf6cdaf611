"""Tile plan of the hex8 scatter (fc_tile_plan_*, host integer code): every element sits in exactly one tile slot, the
blocks of a tile are the distinct (owned row node, column node) pairs of its elements in CSR order, and scattering
element matrices through the plan (what k_hex8_nh_tile does in shared memory) reproduces the reference's per-element
assembly (SparseMatrix::assemble, 4C_linalg_sparsematrix.cpp:474-543) entry by entry."""
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


def plan_scatter(plan, nnz, Ke):
    """the tile kernel's data flow: per tile, element blocks summed into [3][3 nb] planes, each block added once"""
    K = np.zeros(nnz)
    for t in range(plan["n_tiles"]):
        r0, r1 = plan["tile_ptr"][t], plan["tile_ptr"][t + 1]
        nb = int(r1 - r0)
        planes = np.zeros((3, 3 * nb))
        for s in range(16):
            e = plan["tile_ele"][t, s]
            if e < 0:
                assert np.all(plan["tile_blk"][t, s] == 0xFFFFFFFF)
                continue
            for lane in range(32):
                a, q = lane >> 2, lane & 3
                bp = int(plan["tile_blk"][t, s, lane])
                if bp & 0xFFFF == 0xFFFF:
                    assert bp == 0xFFFFFFFF
                    continue
                for c, blk in enumerate((bp & 0xFFFF, bp >> 16)):
                    assert blk < nb
                    for i in range(3):
                        planes[i, 3 * blk:3 * blk + 3] += Ke[e, a, i, 2 * q + c, :]
        for blk in range(nb):
            rc = int(plan["rec"][r0 + blk])
            off, rl = rc & ((1 << 44) - 1), (rc >> 44) & 0xFFFF
            assert rc >> 60 == 0
            for i in range(3):
                K[off + i * rl:off + i * rl + 3] += planes[i, 3 * blk:3 * blk + 3]
    return K


@pytest.mark.parametrize("n,nranks,rank", [((8, 4, 4), 1, 0), ((5, 3, 2), 1, 0), ((6, 5, 4), 2, 1), ((7, 6, 5), 4, 2), ((1, 1, 1), 1, 0),
                                           ((4, 4, 4), 8, 7), ((9, 2, 3), 3, 0)])
def test_plan_reproduces_per_element_assembly(fc, n, nranks, rank):
    g = fc.Grid(fc.FC_HEX8, n, nranks=nranks, rank=rank)
    conn, n_row = g.conn, g.n_row_nodes
    nodeptr, nodeadj = node_graph(conn, n_row)
    plan = fc.capi.tile_plan(g.X, conn, n_row, nodeptr, nodeadj)
    assert plan is not None
    # every element exactly once
    eles = plan["tile_ele"].reshape(-1)
    assert sorted(eles[eles >= 0].tolist()) == list(range(len(conn)))
    # blocks of a tile: distinct (row, column) pairs, rows ascending, columns ascending inside a row (CSR order)
    for t in range(plan["n_tiles"]):
        r0, r1 = plan["tile_ptr"][t], plan["tile_ptr"][t + 1]
        te = [e for e in plan["tile_ele"][t] if e >= 0]
        pairs = sorted({(int(conn[e][a]), int(conn[e][b])) for e in te for a in range(8) for b in range(8) if conn[e][a] < n_row})
        assert r1 - r0 == len(pairs) <= plan["max_blocks"]
        assert plan["rownode"][r0:r1].tolist() == [p[0] for p in pairs]
        offs = [int(x) & ((1 << 44) - 1) for x in plan["rec"][r0:r1]]
        for (a, b), off in zip(pairs, offs):
            row = nodeadj[nodeptr[a]:nodeptr[a + 1]]
            assert off == 9 * nodeptr[a] + 3 * int(np.searchsorted(row, b))
    rng = np.random.default_rng(5)
    Ke = rng.standard_normal((len(conn), 8, 3, 8, 3))
    K_ref = direct_scatter(conn, n_row, nodeptr, nodeadj, Ke)
    K_plan = plan_scatter(plan, len(K_ref), Ke)
    assert np.allclose(K_plan, K_ref, rtol=0, atol=1e-12)


def test_structured_tiles_are_4x2x2(fc):
    """on a cube whose sides divide by 4, 2, 2 every tile is a full 4 x 2 x 2 brick: 45 nodes, 637 blocks instead of
    the 1024 element blocks (the reduction the kernel's roofline estimate in DESIGN.md uses)"""
    g = fc.Grid(fc.FC_HEX8, (12, 6, 4))
    nodeptr, nodeadj = node_graph(g.conn, g.n_row_nodes)
    plan = fc.capi.tile_plan(g.X, g.conn, g.n_row_nodes, nodeptr, nodeadj)
    assert plan["n_tiles"] == 12 * 6 * 4 // 16 and plan["max_blocks"] == 637
    assert np.all(np.diff(plan["tile_ptr"]) == 637)
    for t in range(plan["n_tiles"]):
        nodes = np.unique(g.conn[plan["tile_ele"][t]])
        assert len(nodes) == 45
        ext = g.X[nodes].max(0) - g.X[nodes].min(0)
        assert np.allclose(ext, [4 / 12, 2 / 6, 2 / 4])


def test_degenerate_element_has_no_plan(fc):
    g = fc.Grid(fc.FC_HEX8, (2, 2, 2))
    conn = g.conn.copy()
    conn[3, 5] = conn[3, 4]  # collapsed edge: two accumulator lanes of one element would hit the same block
    nodeptr, nodeadj = node_graph(conn, g.n_row_nodes)
    assert fc.capi.tile_plan(g.X, conn, g.n_row_nodes, nodeptr, nodeadj) is None


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
