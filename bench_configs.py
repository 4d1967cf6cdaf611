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


def run(fc, name, ct, form, mat, nue, n, steps, flags=0):
    import torch
    dev = torch.device("cuda", 0)
    t0 = time.time()
    if ct == fc.FC_TET10:
        gid, X, conn = tet10_cube(n[0])
        owner = np.zeros(len(gid), np.int32)
        egid = np.arange(len(conn), dtype=np.int32)
    else:
        g = fc.Grid(ct, n)
        gid, X, conn, owner, egid = g.node_gid, g.X, g.conn, g.node_owner, g.ele_gid
    s = fc.Solid(0)
    s.set_stream(torch.cuda.current_stream().cuda_stream)
    s.set_mesh(ct, form, X, gid, owner, conn, egid)
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


def run_newton(fc, edge, steps):
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
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    for it in range(12):
        ev[0].record()
        s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=fc.FC_EVAL_ZERO_FIRST, want_errors=False)
        torch.sub(fext, f, out=r)
        s.apply_dirichlet(K)
        x.zero_()
        s.apply_dirichlet_to_system(x, r, None)
        ev[1].record()
        rn = float(r.norm())
        cg_it, rr = s.linear_solve(K, r, x, method=fc.FC_SOLVER_CG, rel_tol=1e-8, max_iter=5000, check=False)
        u += x
        ev[2].record()
        torch.cuda.synchronize()
        hist.append({"residual": rn, "eval_ms": ev[0].elapsed_time(ev[1]), "solve_ms": ev[1].elapsed_time(ev[2]), "cg_iterations": cg_it, "cg_rel_residual": rr})
        if rn < 1e-9 * hist[0]["residual"]:
            break
    assert s.error_state() == (0, -1)
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    out = {"config": f"F1 device Newton step, hex8 NeoHooke {edge}^3, clamped x-, load on x+", "elements": edge ** 3, "n_dof": n_row, "nnz": int(nnz),
           "spmv_ms": spmv_ms, "spmv_algorithmic_GBps": spmv_bytes / (spmv_ms * 1e-3) / 1e9, "measured_peaks": peak,
           "newton": hist, "max_displacement": float(u.abs().max())}
    print(json.dumps(out), flush=True)
    s.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default=None)
    ap.add_argument("--steps", type=int, default=5)
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
        ("C4s hex8 F-bar NeoHooke nu=0.499 100^3", H8, FB, NH, 0.499, (100, 100, 100), 0),
        ("C4 hex8 F-bar NeoHooke nu=0.499 200^3", H8, FB, NH, 0.499, (200, 200, 200), 0),
        ("C5/8 hex8 NeoHooke 200^3 (one GPU's share of 400^3)", H8, D, NH, 0.25, (200, 200, 200), 0),
        ("C5/8-oc hex8 NeoHooke 200^3, owner-computed rows", H8, D, NH, 0.25, (200, 200, 200), fc.FC_EVAL_OWNER_COMPUTED),
    ]
    if args.only and args.only.startswith("newton"):
        run_newton(fc, int(args.only[6:] or 100), args.steps)
        return
    for c in cfgs:
        if args.only and args.only not in c[0]:
            continue
        if not args.only and ("C3a-96" in c[0] or "C3a-full" in c[0]):
            continue  # large hex27 cases only on request
        run(fc, c[0], c[1], c[2], c[3], c[4], c[5], args.steps, c[6])


if __name__ == "__main__":
    main()
