"""Host-side logic of the product library that needs no GPU: exported C ABI, structured grid +
reference decomposition (closed form) against the oracle's replay of the reference procedure,
halo plan consistency, failure behaviour without a device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_every_declared_symbol_is_exported(fc):
    hdr = open(os.path.join(ROOT, "include", "fourc_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(fc_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 30
    lib = fc.capi.lib()
    missing = [n for n in sorted(declared) if not hasattr(lib, n)]
    assert not missing, missing
    assert declared == set(fc.capi.EXPORTS), declared ^ set(fc.capi.EXPORTS)


def test_no_cpu_fallback(fc):
    """without a CUDA device the compute entry point must fail loudly, not fall back"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(fc.FourCError) as ei:
        fc.Solid(0)
    assert ei.value.code == fc.FC_ERR_NO_DEVICE
    assert "no CPU fallback" in str(ei.value)


def test_product_never_touches_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "4c_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle_lib" not in src and "liboracle" not in src and "fc_oracle" not in src, f


@pytest.mark.parametrize("ct", [0, 1])
@pytest.mark.parametrize("n,nranks", [((3, 3, 3), 2), ((4, 2, 5), 3), ((5, 5, 5), 4), ((7, 3, 2), 6), ((6, 6, 6), 8),
                                      ((2, 2, 2), 12), ((1, 1, 1), 2), ((9, 4, 4), 5)])
def test_grid_and_partition_match_reference_procedure(fc, oracle, ct, n, nranks):
    job = oracle.Job.cuboid(ct, n, nranks=nranks)
    idx = {int(g): i for i, g in enumerate(job.node_gid)}
    total_row = 0
    for r in range(nranks):
        g = fc.Grid(ct, n, nranks=nranks, rank=r)
        o = job.rank(r)
        assert g.subdivisions == job.subdivisions()
        assert g.n_row_nodes == o["n_row_nodes"] and g.n_row_eles == o["n_row_eles"]
        assert np.array_equal(g.node_gid, o["col_node_gid"])       # owned ascending, ghosts by (owner, gid)
        assert np.array_equal(g.ele_gid, o["col_eles"])            # row elements first
        col = np.array([idx[int(x)] for x in g.node_gid], dtype=np.int64)
        assert np.array_equal(g.node_owner, job.node_owner[col]) if len(col) else True
        assert np.array_equal(g.X, job.X[col].reshape(-1, 3))
        assert np.array_equal(g.ele_owner, job.ele_owner[g.ele_gid])
        assert np.array_equal(g.node_gid[g.conn], job.conn[g.ele_gid]) if len(g.ele_gid) else True
        # DOF numbering 3 (gid - min gid) + j  (4C_fem_dofset.cpp:344-352)
        assert np.array_equal(g.dof_gid(g.node_gid), o["col_dof_gid"])
        total_row += g.n_row_nodes
    assert total_row == len(job.node_gid)


def test_halo_plan_is_consistent(fc):
    """what rank p sends to q is exactly what q expects from p, in q's ghost order"""
    ct, n, nranks = 0, (6, 5, 4), 6
    grids = [fc.Grid(ct, n, nranks=nranks, rank=r) for r in range(nranks)]
    for q, gq in enumerate(grids):
        ghosts = gq.node_gid[gq.n_row_nodes:]
        owners = gq.node_owner[gq.n_row_nodes:]
        off = 0
        for p in range(nranks):
            cnt = int(gq.recv_count[p])
            want = ghosts[off:off + cnt]
            assert np.all(owners[off:off + cnt] == p)
            off += cnt
            gp = grids[p]
            s0 = int(gp.send_count[:q].sum())
            sent = gp.node_gid[gp.send_nodes[s0:s0 + int(gp.send_count[q])]]
            assert np.array_equal(sent, want), (p, q)
        assert off == len(ghosts)


def test_c2_and_c5_sizes(fc):
    """BASELINE configs: C2 100^3 on 1 rank; C5 400^3 split 2x2x2 (sizes only, rank 0 of a 40^3 stand-in)"""
    g = fc.Grid(0, (100, 100, 100))
    assert (g.n_col_eles, g.n_col_nodes) == (1_000_000, 1_030_301)
    assert g.node_gid[-1] == 201 ** 3 - 1 and g.min_node_gid == 0
    g8 = fc.Grid(0, (40, 40, 40), nranks=8, rank=7)
    assert g8.subdivisions == (2, 2, 2)
    assert g8.n_row_eles + g8.n_col_eles > 0
    with pytest.raises(fc.FourCError):
        fc.Grid(2, (2, 2, 2))  # the reference generator has no tet cells
