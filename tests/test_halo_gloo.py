"""N>1 host path on CPU: world_size-2/-4 gloo processes run the forward halo import of
Discretization::set_state (src/core/fem/src/discretization/4C_fem_discretization.cpp:503-558) with the
product's halo plan (fc_grid_*): owned interface DOFs are gathered by the send list and land behind the
owned DOFs grouped by owner.  The pack kernel itself is GPU code (tests/test_gpu_multi.py); here the
gather is numpy so that plan + exchange are covered without a device."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def field(X):
    return np.stack([np.sin(3 * X[:, 0]) + X[:, 1], X[:, 2] ** 2 - X[:, 0], X[:, 0] * X[:, 1] * X[:, 2] + 1.0], axis=1)


def worker(rank, world, port, ct, n, out):
    sys.path.insert(0, ROOT)
    fc = importlib.import_module("4c_b200")
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    g = fc.Grid(ct, n, nranks=world, rank=rank)
    n_row = 3 * g.n_row_nodes
    u_col = np.full(3 * g.n_col_nodes, np.nan)
    u_col[:n_row] = field(g.X[:g.n_row_nodes]).reshape(-1)           # row-layout state
    send_lids = (3 * g.send_nodes[:, None] + np.arange(3)).reshape(-1)
    sendbuf = torch.from_numpy(u_col[send_lids].copy())               # == fc_solid_halo_pack
    recv = torch.empty(3 * int(g.recv_count.sum()), dtype=torch.float64)
    ops, so, ro = [], 0, 0
    for p in range(world):
        ns, nr = 3 * int(g.send_count[p]), 3 * int(g.recv_count[p])
        if ns:
            ops.append(dist.P2POp(dist.isend, sendbuf[so:so + ns], p))
        if nr:
            ops.append(dist.P2POp(dist.irecv, recv[ro:ro + nr], p))
        so, ro = so + ns, ro + nr
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    u_col[n_row:] = recv.numpy()                                      # no unpack: ghosts are grouped by owner
    ok = np.array_equal(u_col, field(g.X).reshape(-1))
    t = torch.tensor([1 if ok else 0])
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        out.put(int(t.item()))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,ct,n", [(2, 0, (6, 4, 4)), (4, 0, (6, 6, 3)), (2, 1, (3, 2, 2))])
def test_forward_halo_import_gloo(world, ct, n):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=worker, args=(r, world, port, ct, n, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert out.get(timeout=5) == 1
