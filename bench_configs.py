#!/usr/bin/env python
"""Secondary measurements: the other BASELINE.json configs (C1, C3, C4 and a C5-sized per-GPU box) on one
GPU, device-resident, CUDA events.  bench.py remains the contract benchmark (C2); this script fills the
per-config table of BASELINE.md / DESIGN.md.  One JSON line per config.

  python bench_configs.py [--only NAME] [--steps K]
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from bench import smooth_displacement  # noqa: E402


def tet10_cube(n):
    """n^3 cells x 6 tets (Kuhn split along the main diagonal) on the (2n+1)^3 lattice, vectorised.
    Node order of 4C_fem_general_utils_fem_shapefunctions.hpp:244-259."""
    N = 2 * n + 1
    i, j, k = np.meshgrid(np.arange(n), np.arange(n), np.arange(n), indexing="ij")
    o = np.stack([2 * i.ravel(), 2 * j.ravel(), 2 * k.ravel()], axis=1)  # cell origins on the lattice
    perms = [(0, 1, 2), (0, 2, 1), (1, 0, 2), (1, 2, 0), (2, 0, 1), (2, 1, 0)]
    conn = []
    for p in perms:
        c = [o.copy()]
        for ax in p:
            v = c[-1].copy()
            v[:, ax] += 2
            c.append(v)
        c = np.stack(c, axis=1).astype(np.int64)  # [cells, 4, 3]
        det = np.linalg.det((c[:, 1:] - c[:, :1]).astype(float))
        neg = det < 0
        c[neg] = c[neg][:, [0, 1, 3, 2]]
        mids = [(0, 1), (1, 2), (0, 2), (0, 3), (1, 3), (2, 3)]
        pts = [c[:, a] for a in range(4)] + [(c[:, a] + c[:, b]) // 2 for a, b in mids]
        ids = np.stack([(q[:, 2] * N + q[:, 1]) * N + q[:, 0] for q in pts], axis=1)
        conn.append(ids)
    conn = np.stack(conn, axis=1).reshape(-1, 10)  # the 6 tets of a cell are consecutive (cell-major element order)
    used, inv = np.unique(conn, return_inverse=True)
    X = np.stack([(used % N), (used // N) % N, used // (N * N)], axis=1) / (2.0 * n)
    return used.astype(np.int32), X, inv.reshape(conn.shape).astype(np.int32)


def nurbs27_patch(n):
    """one patch of n^3 nurbs27 elements on the unit cube: open uniform knot vectors of degree 2, (n + 2)^3 control points at the
    Greville abscissae (the geometry map is the identity), weights 1 +- 10 %; element (i, j, k) holds the control points
    [i, i + 3) x [j, j + 3) x [k, k + 3) with u fastest and the knots [i, i + 6) of each direction"""
    kv = np.concatenate([[0.0, 0.0], np.arange(n + 1) / n, [1.0, 1.0]])
    g = (kv[1:n + 3] + kv[2:n + 4]) / 2
    m = n + 2
    k, j, i = np.meshgrid(np.arange(m), np.arange(m), np.arange(m), indexing="ij")
    X = np.stack([g[i.ravel()], g[j.ravel()], g[k.ravel()]], axis=1)
    ek, ej, ei = [a.ravel() for a in np.meshgrid(np.arange(n), np.arange(n), np.arange(n), indexing="ij")]
    a = np.arange(27)
    conn = ((ek[:, None] + a // 9) * m + (ej[:, None] + (a // 3) % 3)) * m + (ei[:, None] + a % 3)
    knots = np.stack([kv[ei[:, None] + np.arange(6)], kv[ej[:, None] + np.arange(6)], kv[ek[:, None] + np.arange(6)]], axis=1)
    w = 1.0 + 0.1 * np.sin(7.0 * X[:, 0] + 3.0 * X[:, 1] + 5.0 * X[:, 2])
    return np.arange(len(X), dtype=np.int32), X, conn.astype(np.int32), knots, w


def run(fc, name, ct, form, mat, nue, n, steps, flags=0):
    import torch
    dev = torch.device("cuda", 0)
    t0 = time.time()
    if ct == fc.FC_TET10:
        gid, X, conn = tet10_cube(n[0])
        owner = np.zeros(len(gid), np.int32)
        egid = np.arange(len(conn), dtype=np.int32)
    elif ct == fc.FC_NURBS27:
        gid, X, conn, knots, weights = nurbs27_patch(n[0])
        owner = np.zeros(len(gid), np.int32)
        egid = np.arange(len(conn), dtype=np.int32)
    else:
        g = fc.Grid(ct, n)
        gid, X, conn, owner, egid = g.node_gid, g.X, g.conn, g.node_owner, g.ele_gid
    s = fc.Solid(0)
    s.set_stream(torch.cuda.current_stream().cuda_stream)
    s.set_mesh(ct, form, X, gid, owner, conn, egid)
    if ct == fc.FC_NURBS27:
        s.set_nurbs(knots, weights)
    s.set_material(mat, 10.0, nue)
    n_row, n_col, nnz = s.build_graph()
    setup_s = time.time() - t0
    u = torch.from_numpy(smooth_displacement(X, 1.0 / max(n)).reshape(-1)).to(dev)
    K = torch.empty(nnz, dtype=torch.float64, device=dev)
    f = torch.empty(n_row, dtype=torch.float64, device=dev)
    fl = flags | fc.FC_EVAL_ZERO_FIRST
    for _ in range(3):
        s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=fl, want_errors=False)
    torch.cuda.synchronize()
    assert s.error_state() == (0, -1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=fl, want_errors=False)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    zero_ms, ker_ms = s.last_kernel_ms()
    # residual-only action
    e0.record()
    for _ in range(steps):
        s.evaluate(fc.FC_ACTION_INTERNALFORCE, u, f, None, flags=fl, want_errors=False)
    e1.record()
    torch.cuda.synchronize()
    ms_f = e0.elapsed_time(e1) / steps
    nele, nen = len(egid), fc.NEN[ct]
    alg_bytes = 8 * nnz + 72 * len(gid) + 4 * nen * nele
    out = {"config": name, "elements": nele, "nodes": int(len(gid)), "nnz": int(nnz), "ms_per_eval": ms, "kernel_ms": ker_ms,
           "zero_fill_ms": zero_ms, "elements_per_s": nele / (ms * 1e-3), "assembled_nnz_per_s": nnz / (ms * 1e-3),
           "internalforce_ms": ms_f, "algorithmic_GBps_kernel": alg_bytes / (ker_ms * 1e-3) / 1e9, "setup_s": setup_s,
           "checksum_f": float(f.abs().sum())}
    print(json.dumps(out), flush=True)
    s.close()
    del K, f, u
    torch.cuda.empty_cache()


def run_newton(fc, edge, steps, method=None, mg_sweeps=2, mg_fp32=True, mg_damping=None):
    """SURVEY 8f F1 at C2 size: one load step of a clamped hex8 Neo-Hooke cube under a nodal load on the x+ face, full Newton with
    every stage on the device (element evaluation, Dirichlet application, CG with node-block Jacobi): times per stage, the
    y = K x kernel against the HBM roofline, CG iterations."""
    import torch
    dev = torch.device("cuda", 0)
    g = fc.Grid(fc.FC_HEX8, (edge, edge, edge))
    s = fc.Solid(0)
    s.set_stream(torch.cuda.current_stream().cuda_stream)
    s.set_mesh(fc.FC_HEX8, fc.FC_FORM_DISPLACEMENT, g.X, g.node_gid, g.node_owner, g.conn, g.ele_gid)
    s.set_material(fc.FC_MAT_ELASTHYPER_COUPNEOHOOKE, 10.0, 0.25)
    n_row, n_col, nnz = s.build_graph()
    fixed = np.flatnonzero(np.repeat(g.X[:, 0] < 1e-12, 3)).astype(np.int32)
    s.set_dirichlet_toggle(fixed)
    fext = np.zeros(n_row)
    loaded = np.flatnonzero(g.X[:, 0] > 1.0 - 1e-12)
    fext[3 * loaded + 2] = 0.02 / len(loaded)  # total transverse force 0.02 on the x+ face
    fext = torch.from_numpy(fext).to(dev)
    u = torch.zeros(n_col, dtype=torch.float64, device=dev)
    K = torch.empty(nnz, dtype=torch.float64, device=dev)
    f = torch.empty(n_row, dtype=torch.float64, device=dev)
    r = torch.empty(n_row, dtype=torch.float64, device=dev)
    x = torch.empty(n_row, dtype=torch.float64, device=dev)
    # y = K x alone
    s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=fc.FC_EVAL_ZERO_FIRST, want_errors=False)
    for _ in range(3):
        s.spmv(K, fext, x)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        s.spmv(K, fext, x)
    e1.record()
    torch.cuda.synchronize()
    spmv_ms = e0.elapsed_time(e1) / steps
    spmv_bytes = 8 * nnz + 4 * (nnz // 9) + 16 * n_row  # K once, one column-node id per 3x3 block, x and y once
    hist = []
    method = fc.FC_SOLVER_CG if method is None else method
    levels = s.mg_setup() if method == fc.FC_SOLVER_CG_MG else 1
    if method == fc.FC_SOLVER_CG_MG:
        s.mg_set_sweeps(mg_sweeps)
        s.mg_set_precision(mg_fp32)
        if mg_damping:
            s.mg_set_damping(mg_damping)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    for it in range(12):
        ev[0].record()
        s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=fc.FC_EVAL_ZERO_FIRST, want_errors=False)
        torch.sub(fext, f, out=r)
        s.apply_dirichlet(K)
        x.zero_()
        s.apply_dirichlet_to_system(x, r, None)
        if method == fc.FC_SOLVER_CG_MG:
            s.mg_update(u)  # coarse tangents at the injected state: part of the evaluation time
        ev[1].record()
        rn = float(r.norm())
        cg_it, rr = s.linear_solve(K, r, x, method=method, rel_tol=1e-8, max_iter=5000, check=False)
        u += x
        ev[2].record()
        torch.cuda.synchronize()
        hist.append({"residual": rn, "eval_ms": ev[0].elapsed_time(ev[1]), "solve_ms": ev[1].elapsed_time(ev[2]), "cg_iterations": cg_it, "cg_rel_residual": rr})
        if rn < 1e-9 * hist[0]["residual"]:
            break
    assert s.error_state() == (0, -1)
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    out = {"config": f"F1 device Newton step, hex8 NeoHooke {edge}^3, clamped x-, load on x+", "solver": ["CG + node-block Jacobi", "BiCGStab", f"CG + geometric multigrid V({mg_sweeps},{mg_sweeps}), {levels} levels, {'FP32' if mg_fp32 else 'FP64'} cycle matrices"][method], "elements": edge ** 3, "n_dof": n_row, "nnz": int(nnz),
           "spmv_ms": spmv_ms, "spmv_algorithmic_GBps": spmv_bytes / (spmv_ms * 1e-3) / 1e9, "measured_peaks": peak,
           "newton": hist, "max_displacement": float(u.abs().max())}
    if method == fc.FC_SOLVER_CG_MG:
        out["mg_levels"] = [dict(zip(("nodes", "jacobi_omega"), s.mg_level_info(l))) for l in range(levels)]
    print(json.dumps(out), flush=True)
    s.close()


def run_variant_b(fc, edge, steps):
    """SURVEY 8e variant B against variant A, measured under torchrun on N GPUs (weak scaling, edge^3 per GPU).
    A (built, bit-comparable with the reference's decomposition): forward halo import of u, every rank evaluates its owned
      AND ghost elements, assembles owned rows only - no reduction of f or K.
    B (north_star's wording): every element evaluated once by its owner; the matrix rows of ghost nodes that the rank's
      owned elements touch are exported to their owners and summed there.
    B is measured as a cost model on the real partition: the element kernel over the owned elements only, plus the export
    of EXACTLY the bytes B would send (9 doubles per (ghost row node, coupled column node) pair through owned elements,
    3 per ghost node for f) as one grouped exchange, plus the add of the received payload into K.  The index translation
    between the sender's and the owner's CSR positions (a gather list per peer) is not built: it only adds to B."""
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from bench import Problem, global_edge
    comm = fc.capi.nccl_comm_init(rank, world, local)
    P = Problem(fc, torch, edge, rank, world, local, comm)
    g, s = P.grid, P.s
    flags = fc.FC_EVAL_ZERO_FIRST
    # ---- variant A --------------------------------------------------------------------------------
    for _ in range(3):
        P.step(flags, False)
    ms_a, zero_a, ker_a, _ = P.timed(steps, flags, False)
    halo_bytes = 8 * (P.n_col_dof - P.n_row_dof)
    # ---- variant B payload from the partition ---------------------------------------------------------
    n_row_nodes, n_row_eles = g.n_row_nodes, g.n_row_eles
    own = g.conn[:n_row_eles]                       # row elements come first
    a = np.repeat(own, 8, axis=1).reshape(-1)       # row node of every (a, b) pair
    b = np.tile(own, (1, 8)).reshape(-1)
    m = a >= n_row_nodes                            # rows of ghost nodes
    pairs = np.unique(a[m].astype(np.int64) * g.n_col_nodes + b[m])
    ghost_rows = (pairs // g.n_col_nodes).astype(np.int64)
    owner = g.node_owner[ghost_rows]
    send_doubles = np.array([9 * int(np.sum(owner == p)) + 3 * len(np.unique(ghost_rows[owner == p])) for p in range(world)], np.int64)
    send_t = torch.tensor(send_doubles, device=dev)
    recv_t = torch.empty_like(send_t)
    dist.all_to_all_single(recv_t, send_t)
    recv_doubles = recv_t.cpu().numpy()
    sendbuf = torch.ones(int(send_doubles.sum()), dtype=torch.float64, device=dev)
    recvbuf = torch.empty(int(recv_doubles.sum()), dtype=torch.float64, device=dev)
    # the element kernel over the owned elements only: a second context on the row elements of this rank
    sB = fc.Solid(local, rank, world)
    sB.set_stream(torch.cuda.current_stream().cuda_stream)
    sB.set_mesh(fc.FC_HEX8, fc.FC_FORM_DISPLACEMENT, g.X, g.node_gid, g.node_owner, own, g.ele_gid[:n_row_eles])
    sB.set_material(fc.FC_MAT_ELASTHYPER_COUPNEOHOOKE, 10.0, 0.25)
    sB.build_graph()

    def step_b():
        P.halo_library()  # B needs the ghost displacements as well (its owned elements touch ghost nodes)
        sB.evaluate(fc.FC_ACTION_NLNSTIFF, P.u_col, P.f_row, P.K, flags=flags, want_errors=False)
        dist.all_to_all_single(recvbuf, sendbuf, [int(v) for v in recv_doubles], [int(v) for v in send_doubles])
        P.K[:recvbuf.numel()] += recvbuf  # add of the received rows (contiguous stand-in for the gather-add)

    for _ in range(3):
        step_b()
    P.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step_b()
    e1.record()
    P.barrier()
    t = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    z, k = sB.kernel_ms_history(min(steps, 64))
    tot = torch.tensor([float(send_doubles.sum()) * 8, float(halo_bytes)], dtype=torch.float64, device=dev)
    dist.all_reduce(tot)
    if rank == 0:
        print(json.dumps({"config": f"variant A vs variant B, hex8 NeoHooke {edge}^3 per GPU on {world} GPUs, partition " + "x".join(str(v) for v in g.subdivisions),
                          "variant_A": {"ms_per_step": ms_a, "kernel_ms": ker_a, "zero_fill_ms": zero_a, "exchange_bytes_all_ranks": float(tot[1].item()),
                                        "elements_evaluated_rank0": int(g.n_col_eles), "what": "halo import of u + owned and ghost elements, owner rows only"},
                          "variant_B": {"ms_per_step": float(t.item()), "kernel_ms": float(k.mean()), "zero_fill_ms": float(z.mean()),
                                        "exchange_bytes_all_ranks": float(tot[0].item()), "elements_evaluated_rank0": int(n_row_eles),
                                        "what": "halo import of u + owned elements only + export of the ghost-node rows (exact byte counts) + add; cost model, index translation not built"},
                          "bytes_ratio_B_over_A": float(tot[0].item() / tot[1].item())}), flush=True)
    sB.close()
    P.close()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default=None)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--mg-sweeps", type=int, default=2)
    ap.add_argument("--mg-damping", type=float, default=None, help="omega = this / lambda_max estimate (library default 4/3.3)")
    ap.add_argument("--mg-fp64", action="store_true", help="V-cycle on the FP64 tangents instead of their single-precision copies")
    args = ap.parse_args()
    fc = importlib.import_module("4c_b200")
    H8, H27, T10 = fc.FC_HEX8, fc.FC_HEX27, fc.FC_TET10
    D, FB = fc.FC_FORM_DISPLACEMENT, fc.FC_FORM_FBAR
    SVK, NH = fc.FC_MAT_STVENANTKIRCHHOFF, fc.FC_MAT_ELASTHYPER_COUPNEOHOOKE
    cfgs = [
        ("C1 hex8 SVK 10^3", H8, D, SVK, 0.3, (10, 10, 10), 0),
        ("C1b hex8 SVK 100^3", H8, D, SVK, 0.3, (100, 100, 100), 0),
        ("C2 hex8 NeoHooke 100^3 (tensor-core closed form)", H8, D, NH, 0.25, (100, 100, 100), 0),
        ("C2 hex8 NeoHooke 100^3 (owner-computed rows, tensor-core closed form)", H8, D, NH, 0.25, (100, 100, 100), fc.FC_EVAL_OWNER_COMPUTED),
        ("C2 hex8 NeoHooke 100^3 (FMA closed form)", H8, D, NH, 0.25, (100, 100, 100), fc.FC_EVAL_CLOSED_FORM_FMA),
        ("C2 hex8 NeoHooke 100^3 (generic B-operator)", H8, D, NH, 0.25, (100, 100, 100), fc.FC_EVAL_FORCE_GENERIC),
        ("C3a hex27 NeoHooke 48^3", H27, D, NH, 0.25, (48, 48, 48), 0),
        ("C3a-96 hex27 NeoHooke 96^3", H27, D, NH, 0.25, (96, 96, 96), 0),
        ("C3a-full hex27 NeoHooke 126^3 (K values 74.6 GB)", H27, D, NH, 0.25, (126, 126, 126), 0),
        ("C3b tet10 NeoHooke 70^3x6", T10, D, NH, 0.25, (70, 70, 70), 0),
        ("C3b tet10 NeoHooke 70^3x6 (tensor-core tiles)", T10, D, NH, 0.25, (70, 70, 70), fc.FC_EVAL_DMMA_TILES),
        ("C4s hex8 F-bar NeoHooke nu=0.499 100^3", H8, FB, NH, 0.499, (100, 100, 100), 0),
        ("C4 hex8 F-bar NeoHooke nu=0.499 200^3", H8, FB, NH, 0.499, (200, 200, 200), 0),
        ("C5/8 hex8 NeoHooke 200^3 (one GPU's share of 400^3)", H8, D, NH, 0.25, (200, 200, 200), 0),
        ("C5/8-oc hex8 NeoHooke 200^3, owner-computed rows", H8, D, NH, 0.25, (200, 200, 200), fc.FC_EVAL_OWNER_COMPUTED),
        # SURVEY 8f F4 formulations (on request): element histories in HBM (EAS 325 / 1009 doubles per element, MULF 144)
        ("F4 hex8 eas_mild NeoHooke 100^3", H8, fc.FC_FORM_EAS_MILD, NH, 0.25, (100, 100, 100), 0),
        ("F4 hex8 eas_full NeoHooke 100^3", H8, fc.FC_FORM_EAS_FULL, NH, 0.25, (100, 100, 100), 0),
        ("F4 hex8 mulf NeoHooke 100^3", H8, fc.FC_FORM_MULF, NH, 0.25, (100, 100, 100), 0),
        ("F4 hex8 linear SVK 100^3", H8, fc.FC_FORM_LINEAR, SVK, 0.3, (100, 100, 100), 0),
        # nurbs27 on the generic kernel with per-element shape tables (41.5 KB per element), next to hex27 on the same kernel
        ("F4 nurbs27 NeoHooke 46^3 patch (48^3 control points)", fc.FC_NURBS27, D, NH, 0.25, (46, 46, 46), 0),
        ("F4 hex27 NeoHooke 48^3 (generic B-operator)", H27, D, NH, 0.25, (48, 48, 48), fc.FC_EVAL_FORCE_GENERIC),
    ]
    if args.only and args.only.startswith("variantB"):
        run_variant_b(fc, int(args.only[8:] or 100), args.steps)
        return
    if args.only and args.only.startswith("newton"):
        mg = args.only.startswith("newtonmg")
        run_newton(fc, int(args.only[8 if mg else 6:] or 100), args.steps, fc.FC_SOLVER_CG_MG if mg else None, args.mg_sweeps, not args.mg_fp64, args.mg_damping)
        return
    names = args.only.split("|") if args.only else []
    exact = any(c[0] in names for c in cfgs)  # full config names (separated by |) select exactly those configs
    for c in cfgs:
        if args.only and (c[0] not in names if exact else args.only not in c[0]):
            continue
        if not args.only and ("C3a-96" in c[0] or "C3a-full" in c[0] or c[0].startswith("F4 ")):
            continue  # large hex27 cases and the F4 formulations only on request
        run(fc, c[0], c[1], c[2], c[3], c[4], c[5], args.steps, c[6])


if __name__ == "__main__":
    main()
