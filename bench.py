#!/usr/bin/env python
"""bench.py — hex8 Neo-Hooke force+stiffness evaluation and assembly (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--edge E] [--impl ours|reference]

A "step" is one struct_calc_nlnstiff evaluation of the whole mesh: zero K and f, (N>1: halo import
of u through the library's own NCCL exchange), element evaluation, scatter into the CSR values.  Workload at N=1:
BASELINE config C2 (100^3 hex8 cube, MAT_ElastHyper + ELAST_CoupNeoHooke YOUNG 10 NUE 0.25, 1 evaluation).  N>1: weak
scaling, every rank keeps a 100^3 box of the reference's structured partition; the N=8 line also carries a "c5" record
(BASELINE C5: 400^3 = 64 M elements, 200^3 per GPU).
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
METRIC = "hex8_neohooke_force_stiffness_assembly_elements_per_s"
UNIT = "elements/s"
YOUNG, NUE = 10.0, 0.25


def smooth_displacement(X, h, seed=12345, amp=0.02, noise=1e-3, gid=None):
    """SURVEY 8d displacement state: affine + sine + uniform noise U(-noise, noise) h.  One rank: noise from PCG64 seed
    12345 in node order.  Several ranks (gid given): the noise is a function of the node's global id, so that every rank
    computes the same value for a node it shares - the ghost values the halo import delivers can then be checked
    bit for bit against the field."""
    A = np.array([[0.03, 0.01, -0.02], [0.00, -0.02, 0.015], [0.01, 0.02, 0.025]])
    u = X @ A.T + amp * np.sin(2 * np.pi * X[:, [1, 2, 0]])
    if gid is None:
        rng = np.random.Generator(np.random.PCG64(seed))
        u += rng.uniform(-noise, noise, size=X.shape) * h
    else:
        g = np.asarray(gid, np.uint64)[:, None] * np.uint64(3) + np.arange(3, dtype=np.uint64)[None, :] + np.uint64(seed)
        g = (g ^ (g >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)   # splitmix64 finaliser
        g = (g ^ (g >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        g = g ^ (g >> np.uint64(31))
        r = (g >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))   # uniform [0, 1)
        u += (2.0 * r - 1.0) * noise * h
    return u


def global_edge(edge, nranks):
    """global intervals so that the structured split gives every rank an edge^3 box"""
    n, sub, f = nranks, [1, 1, 1], []
    p = 2
    while n > 1:
        if n % p == 0:
            f.append(p)
            n //= p
        else:
            p += 1
    for fac in reversed(f):
        sub[int(np.argmin(sub))] *= fac
    return tuple(edge * s for s in sub)


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        super().__init__(daemon=True)
        self.device, self.rows, self.proc = device, [], None
        self.window = [None, None]  # wall-clock stamps of the timed region

    def run(self):
        # one long-lived nvidia-smi sampling every 20 ms (a fresh process per sample takes ~100 ms and would
        # miss a timed region of a few dozen ms).  stdout is a pty so that it is line-buffered; every line carries
        # nvidia-smi's own wall-clock timestamp, which places it relative to the timed region.
        import datetime
        import pty
        try:
            master, slave = pty.openpty()
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu=timestamp,{self.Q}", "--format=csv,noheader,nounits", "-i",
                                          str(self.device), "-lms", "20"], stdout=slave, stderr=subprocess.DEVNULL, close_fds=True)
            os.close(slave)
            with os.fdopen(master, "r", errors="ignore") as f:
                for line in f:
                    c = [x.strip() for x in line.strip().split(",")]
                    try:
                        ts = datetime.datetime.strptime(c[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                    except Exception:
                        continue
                    self.rows.append((ts, c[1:]))
        except Exception:
            pass

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        self.join(timeout=5)

    def summary(self):
        t0, t1 = self.window
        rows = [r for t, r in self.rows if len(r) > 8 and t0 is not None and t0 <= t <= t1]
        where = "timed region"
        if len(rows) < 3:  # region shorter than a few sampling periods: add the identical steps run right after it
            rows = [r for t, r in self.rows if len(r) > 8 and t0 is not None and t >= t0]
            where = "timed region + the identical steps after it"
        sm = [float(r[1]) for r in rows if r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in rows if r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "sampled": where}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def workload_string(edge, n):
    """the same text in both arms (the driver compares them)"""
    tag = "C2 " if (edge == 100 and n == (100, 100, 100)) else ""
    return (f"{tag}hex8 compressible Neo-Hooke cube (MAT_ElastHyper + ELAST_CoupNeoHooke YOUNG 10 NUE 0.25), {edge}^3 elements per GPU, "
            f"global {n[0]}x{n[1]}x{n[2]}, struct_calc_nlnstiff: zero K,f + halo import + element evaluation + CSR scatter")


def time_oracle(O, n, cores, warmup, steps):
    """seconds per struct_calc_nlnstiff evaluation of the n[0] x n[1] x n[2] cube on `cores` threads = structured subdomains"""
    job = O.Job.cuboid(O.HEX8, n, nranks=cores)
    job.set_problem(O.FORM_DISP, O.MAT_NEOHOOKE, YOUNG, NUE)
    job.set_u_nodal(smooth_displacement(job.X, 1.0 / max(n)))
    for _ in range(warmup):
        job.evaluate(O.ACTION_NLNSTIFF, max_threads=cores)
    times = []
    for _ in range(steps):
        bad, _ = job.evaluate(O.ACTION_NLNSTIFF, max_threads=cores)
        assert bad == 0
        times.append(job.seconds)
    return float(np.mean(times))


def reference_sample(edge, world, ref_edge):
    """mesh the CPU arm evaluates: the GPU arm's global mesh while it stays within 4 M elements (about 4 s per evaluation),
    a slab of it beyond that"""
    n = global_edge(edge, world)
    if ref_edge and ref_edge != edge:
        return (ref_edge,) * 3, f"a {ref_edge}^3 sample"
    if n[0] * n[1] * n[2] <= 4_000_000:
        return n, "the full workload"
    m = list(n)
    while m[0] * m[1] * m[2] > 4_000_000:
        m[int(np.argmax(m))] //= 2
    return tuple(m), f"a {m[0]}x{m[1]}x{m[2]} slab of it"


def run_reference(args):
    """The reference's CPU implementation of the path.  4C itself cannot be built on this box (no MPI/Trilinos,
    DESIGN.md), so this arm times the CPU restatement in oracle/ ("port"): its timing build (-O3 -march=native, FMA
    contraction allowed), one thread per reference-style structured subdomain (mirrors mpirun -np <cores>), on the GPU
    arm's mesh or a bounded sample of it."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    cflags = O.use_fast_build()
    cores = os.cpu_count() or 1
    n_gpu = global_edge(args.edge, args.gpus)
    n, what = reference_sample(args.edge, args.gpus, args.ref_edge)
    t = time_oracle(O, n, cores, min(args.warmup, 2), min(args.steps, 10))
    val = n[0] * n[1] * n[2] / t
    sample = (f"{what}: {n[0]}x{n[1]}x{n[2]} = {n[0]*n[1]*n[2]} elements per step, {cores} threads = {cores} structured subdomains, "
              f"evaluate + owner-row CSR assembly, oracle timing build (-O3 {[f for f in cflags.split() if f.startswith('-march')][0]}, FMA contraction allowed)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": min(args.steps, 10),
        "warmup": min(args.warmup, 2), "ms_per_step": 1e3 * t, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_string(args.edge, n_gpu)},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# dominant kernel per --kernel-flags, its source files and the committed ncu capture (profiles/extract_ncu.py) that holds
# dram__bytes_read.sum + dram__bytes_write.sum of one launch at C2.  The capture is keyed by the hash of the kernel's
# source: when the source has changed since, the capture describes another kernel and `traffic` is null.
KERNELS = {
    0: ("k_hex8_neohooke_mma<true,false,false>", "profiles/r02_hex8nh_mma_c2.json"),
    8: ("k_hex8_nh_brick<true,false>", "profiles/r02_hex8nh_brick_c2.json"),
    4: ("k_hex8_neohooke<true>", None),
    2: ("k_generic<HEX8,DISP,true>", None),
}


def committed_traffic(kernel_flags):
    """(dram bytes per launch, note) from the committed capture of the kernel, None when absent or stale"""
    import hashlib
    name, path = KERNELS.get(kernel_flags, (None, None))
    if not path or not os.path.exists(os.path.join(ROOT, path)):
        return None, "no committed ncu capture of this kernel"
    with open(os.path.join(ROOT, path)) as f:
        cap = json.load(f)
    h = hashlib.sha256()
    for src in cap.get("sources", []):
        with open(os.path.join(ROOT, src), "rb") as fh:
            h.update(fh.read())
    if not cap.get("sources") or h.hexdigest()[:16] != cap.get("sources_sha16"):
        return None, f"{path} is a capture of an older version of the kernel source"
    return cap["dram_bytes"], f"{path} ({cap['report']}, sources {cap['sources_sha16']})"


class Problem:
    """One rank's share of the edge^3-per-GPU cube, resident on the device: context, state and output buffers, halo plan."""

    def __init__(self, fc, torch, edge, rank, world, local, comm):
        dev = torch.device("cuda", local)
        self.fc, self.torch, self.world, self.rank, self.dev, self.comm = fc, torch, world, rank, dev, comm
        self.n = global_edge(edge, world)
        grid = fc.Grid(fc.FC_HEX8, self.n, nranks=world, rank=rank)
        self.grid = grid
        s = fc.Solid(local, rank, world)
        s.set_stream(torch.cuda.current_stream().cuda_stream)
        s.set_mesh(fc.FC_HEX8, fc.FC_FORM_DISPLACEMENT, grid.X, grid.node_gid, grid.node_owner, grid.conn, grid.ele_gid)
        s.set_material(fc.FC_MAT_ELASTHYPER_COUPNEOHOOKE, YOUNG, NUE)
        self.n_row_dof, self.n_col_dof, self.nnz = s.build_graph()
        self.s = s
        h = 1.0 / max(self.n)
        # several ranks: the field is a function of the node's global id, identical on every rank that holds the node
        self.u_host = smooth_displacement(grid.X, h, gid=grid.node_gid if world > 1 else None).reshape(-1)
        self.u_col = torch.from_numpy(self.u_host).to(dev)
        self.f_row = torch.empty(self.n_row_dof, dtype=torch.float64, device=dev)
        self.K = torch.empty(self.nnz, dtype=torch.float64, device=dev)
        self.main_stream = torch.cuda.current_stream()
        self.side_stream = torch.cuda.Stream() if world > 1 else None
        if world > 1:
            send_lids = (3 * grid.send_nodes[:, None] + np.arange(3)[None, :]).reshape(-1)
            s.halo_set_send(send_lids)
            self.send_split = [3 * int(c) for c in grid.send_count]
            self.recv_split = [3 * int(c) for c in grid.recv_count]
            s.halo_set_counts(self.send_split, self.recv_split)
            s.set_comm(comm)
            self.sendbuf = torch.empty(len(send_lids), dtype=torch.float64, device=dev)
        self.alg_bytes = 8 * self.nnz + 72 * grid.n_col_nodes + 32 * grid.n_col_eles  # K once, X/u/f once, conn once

    # forward import of u (Discretization::set_state): inside the library over its NCCL binding, or through torch
    def halo_library(self):
        if self.world > 1:
            self.s.halo_exchange(self.comm, self.u_col)

    def halo_torch(self):
        if self.world > 1:
            import torch.distributed as dist
            self.s.halo_pack(self.u_col, self.sendbuf)
            dist.all_to_all_single(self.u_col[self.n_row_dof:], self.sendbuf, self.recv_split, self.send_split)

    def halo_check(self):
        """ghosts poisoned with NaN, imported, compared bit for bit with the field every rank can evaluate itself"""
        torch = self.torch
        if self.world == 1:
            return None
        import torch.distributed as dist
        want = torch.from_numpy(self.u_host[self.n_row_dof:]).to(self.dev)
        res = {}
        for name, fn in (("library", self.halo_library), ("torch", self.halo_torch)):
            self.u_col[self.n_row_dof:] = float("nan")
            torch.cuda.synchronize()
            fn()
            torch.cuda.synchronize()
            ok = torch.tensor([1 if torch.equal(self.u_col[self.n_row_dof:], want) else 0], device=self.dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            res[name] = "bit-exact" if int(ok.item()) == 1 else "MISMATCH"
        nd = torch.tensor([self.n_col_dof - self.n_row_dof], device=self.dev)
        dist.all_reduce(nd)
        res["ghost_dofs"] = int(nd.item())
        res["timed_path"] = "library (fc_solid_halo_exchange: pack kernel + grouped ncclSend/ncclRecv)"
        return res

    def step(self, flags, overlap):
        fc, s = self.fc, self.s
        if self.world == 1 or not overlap:
            self.halo_library()
            s.evaluate(fc.FC_ACTION_NLNSTIFF, self.u_col, self.f_row, self.K, flags=flags, want_errors=False)
            return 1
        # halo import on a second stream while the elements without ghost nodes are evaluated, then the boundary elements
        self.side_stream.wait_stream(self.main_stream)
        s.set_stream(self.side_stream.cuda_stream)
        self.halo_library()
        s.set_stream(self.main_stream.cuda_stream)
        s.evaluate(fc.FC_ACTION_NLNSTIFF, self.u_col, self.f_row, self.K, flags=flags | fc.FC_EVAL_INTERIOR, want_errors=False)
        self.main_stream.wait_stream(self.side_stream)
        s.evaluate(fc.FC_ACTION_NLNSTIFF, self.u_col, self.f_row, self.K, flags=(flags & ~fc.FC_EVAL_ZERO_FIRST) | fc.FC_EVAL_BOUNDARY, want_errors=False)
        return 2

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        self.torch.cuda.synchronize()

    def timed(self, steps, flags, overlap):
        """(ms per step: max over ranks, zero-fill ms per step, element-kernel ms per step) of `steps` steps, CUDA events"""
        torch = self.torch
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        t_wall0 = time.time()
        ev0.record()
        calls = 0
        for _ in range(steps):
            calls += self.step(flags, overlap)
        ev1.record()
        self.barrier()
        t_wall1 = time.time()
        t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            import torch.distributed as dist
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        # device time of the zero-fill and of the element kernel inside the timed region (events of the library's ring)
        per = calls // steps
        nh = min(calls, 128) // per * per
        z, e = self.s.kernel_ms_history(nh)
        return float(t.item()) / steps, float(z.sum()) / (nh // per), float(e.sum()) / (nh // per), (t_wall0, t_wall1)

    def close(self):
        self.s.close()
        del self.K, self.f_row, self.u_col
        self.torch.cuda.empty_cache()


def run_e2e(args, fc, torch, P, flags):
    """The same step through the public call with HOST buffers: every step copies u from pinned host memory to the device
    and brings f and the K values back (the reference's caller holds them in host memory).  One GPU: the library's own
    host-buffer path (fc_solid_evaluate stages the copies).  Several GPUs: owned u H2D, halo import on the device, evaluate,
    f and K D2H, all asynchronous on one stream.  PCIe-bound by the K download."""
    import torch.distributed as dist
    world, s = P.world, P.s
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    u_pin = torch.from_numpy(P.u_host[:P.n_row_dof].copy()).pin_memory() if world > 1 else torch.from_numpy(P.u_host).pin_memory()
    f_pin = torch.empty(P.n_row_dof, dtype=torch.float64).pin_memory()
    K_pin = torch.empty(P.nnz, dtype=torch.float64).pin_memory()

    def step():
        if world == 1:
            s.evaluate(fc.FC_ACTION_NLNSTIFF, u_pin, f_pin, K_pin, flags=flags, want_errors=True)
            return
        P.u_col[:P.n_row_dof].copy_(u_pin, non_blocking=True)
        P.halo_library()
        s.evaluate(fc.FC_ACTION_NLNSTIFF, P.u_col, P.f_row, P.K, flags=flags, want_errors=False)
        f_pin.copy_(P.f_row, non_blocking=True)
        K_pin.copy_(P.K, non_blocking=True)
        torch.cuda.synchronize()

    step()
    P.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step()
    P.barrier()
    te = torch.tensor([(time.perf_counter() - t0) / e2e_steps], dtype=torch.float64, device=P.dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    n = P.n
    res = {"value": n[0] * n[1] * n[2] / float(te.item()), "unit": UNIT, "h2d_bytes_per_step": int(8 * (P.n_col_dof if world == 1 else P.n_row_dof)),
           "d2h_bytes_per_step": int(8 * (P.nnz + P.n_row_dof)), "steps": e2e_steps, "ms_per_step": 1e3 * float(te.item()),
           "checksum_f": float(f_pin.abs().sum())}
    # the wall this number sits on: a bare device-to-host copy of the K values, all ranks at once
    P.barrier()
    t0 = time.perf_counter()
    K_pin.copy_(P.K, non_blocking=True)
    torch.cuda.synchronize()
    tw = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=P.dev)
    if world > 1:
        dist.all_reduce(tw, op=dist.ReduceOp.MAX)
    res["d2h_probe"] = {"what": "cudaMemcpyAsync of the K values to pinned host memory, all ranks at once, nothing else running",
                        "ms": 1e3 * float(tw.item()), "aggregate_GBps": world * 8 * P.nnz / float(tw.item()) / 1e9}
    del u_pin, f_pin, K_pin
    return res


def run_c5(args, fc, torch, rank, world, local, comm, hbm_peak):
    """BASELINE C5: the 64 M-element cube (400^3) on 8 GPUs, 200^3 per GPU in the reference's 2x2x2 structured split;
    on fewer GPUs the same 200^3 share per GPU.  Device-resident, halo import overlapped with the interior elements."""
    P = Problem(fc, torch, 200, rank, world, local, comm)
    flags = fc.FC_EVAL_ZERO_FIRST | args.kernel_flags
    overlap = world > 1 and not args.no_overlap
    chk = P.halo_check()
    for _ in range(3):
        P.step(flags, overlap)
    P.barrier()
    assert P.s.error_state() == (0, -1)
    steps = 10
    ms, zero_ms, ele_ms, _ = P.timed(steps, flags, overlap)
    n = P.n
    ne = n[0] * n[1] * n[2]
    nnz_global = 9 * (3 * n[0] + 1) * (3 * n[1] + 1) * (3 * n[2] + 1)
    res = {"workload": f"C5 hex8 Neo-Hooke cube {n[0]}x{n[1]}x{n[2]} = {ne} elements on {world} GPU(s), 200^3 per GPU, partition "
                       + "x".join(str(v) for v in P.grid.subdivisions),
           "steps": steps, "warmup": 3, "ms_per_step": ms, "elements_per_s": ne / (ms * 1e-3), "assembled_nnz_per_s": nnz_global / (ms * 1e-3),
           "nnz": nnz_global, "kernel_ms": ele_ms, "zero_fill_ms": zero_ms, "halo_overlap": bool(overlap),
           "hbm_frac_step_per_gpu": P.alg_bytes / (ms * 1e-3) / 1e9 / hbm_peak, "hbm_frac_kernel_per_gpu": P.alg_bytes / (ele_ms * 1e-3) / 1e9 / hbm_peak,
           "algorithmic_bytes_per_gpu": P.alg_bytes, "north_star_target_frac": 0.60}
    if chk:
        res["halo_check"] = chk
    P.close()
    return res


def run_ours(args):
    import torch
    import torch.distributed as dist
    fc = importlib.import_module("4c_b200")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    comm = None
    if world > 1:
        import datetime
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))  # a mismatch fails in minutes, not in ten
        comm = fc.capi.nccl_comm_init(rank, world, local)  # the library's own communicator (ncclCommInitRank through ctypes)

    # ---- problem setup (not timed) -------------------------------------------------------------------
    P = Problem(fc, torch, args.edge, rank, world, local, comm)
    n, grid, s = P.n, P.grid, P.s
    flags = fc.FC_EVAL_ZERO_FIRST | args.kernel_flags
    overlap = world > 1 and not args.no_overlap and not (args.kernel_flags & (2 | 4))
    halo_check = P.halo_check()
    if halo_check and (halo_check["library"] != "bit-exact" or halo_check["torch"] != "bit-exact"):
        raise SystemExit(f"bench.py: halo import delivered wrong ghost values: {halo_check}")
    for _ in range(max(args.warmup, 3)):
        P.step(flags, overlap)
    P.barrier()
    assert s.error_state() == (0, -1), "element errors in the benchmark state"
    if world > 1:
        assert s.error_allreduce() == 0  # eval_error_check across the ranks (one 4-byte ncclAllReduce)

    # ---- timed region: K steps, CUDA events on the launching stream, max over ranks ----------------
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.1)  # let the sampler come up
    l0 = s.launch_count()
    ms_per_step, zero_ms, ele_ms, win = P.timed(args.steps, flags, overlap)
    sampler.window = list(win)
    launches = s.launch_count() - l0
    if args.steps * ms_per_step < 300.0:
        # a timed region shorter than a few sampling periods: keep the GPU on the identical steps for 0.3 s so that the sampler
        # has rows (the summary says which rows it used).  The count comes from the all-reduced step time: every rank runs the
        # same number of steps, hence the same number of halo exchanges.
        for _ in range(int(300.0 / ms_per_step) + 1):
            P.step(flags, overlap)
        P.barrier()
    sampler.stop()
    n_global_ele = n[0] * n[1] * n[2]
    value = n_global_ele / (ms_per_step * 1e-3)
    nnz_global = 9 * (3 * n[0] + 1) * (3 * n[1] + 1) * (3 * n[2] + 1)

    # ---- roofline of the dominant kernel (per launch = this rank's elements) -----------------------
    hbm_peak, peak_src = load_peaks()
    achieved = P.alg_bytes / (ele_ms * 1e-3) / 1e9
    traffic, traffic_src = (args.traffic, "--traffic") if args.traffic is not None else (None, "the committed capture is of C2 on one GPU; this is another workload")
    if args.traffic is None and args.edge == 100 and world == 1:
        traffic, traffic_src = committed_traffic(args.kernel_flags)
    roofline = {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "kernel": KERNELS.get(args.kernel_flags, (f"flags={args.kernel_flags}", None))[0],
                "kernel_ms": ele_ms, "zero_fill_ms": zero_ms, "timed": "CUDA events of the library inside the timed region, mean per step",
                "algorithmic_bytes_per_launch": P.alg_bytes,
                "step_frac": P.alg_bytes / (ms_per_step * 1e-3) / 1e9 / hbm_peak}
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
           "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
           "data": "synthetic",
           "config": {"workload": workload_string(args.edge, n),
                      "elements": n_global_ele, "nnz": nnz_global, "assembled_nnz_per_s": nnz_global / (ms_per_step * 1e-3),
                      "partition": "x".join(str(v) for v in grid.subdivisions), "halo_overlap": bool(overlap),
                      "l2": "inputs+outputs exceed L2 (K values alone are %.2f GB per GPU)" % (8 * P.nnz / 1e9)},
           "gpu_launches": int(launches), "roofline": roofline}
    if halo_check:
        out["halo_check"] = halo_check

    # FP64 rooflines (the binding ones): probes + algorithmic FLOP of the reference formulation
    if rank == 0:
        try:
            fma_tf, mma_tf = fc.probe_fp64(local)
            out["fp64"] = {"fma_peak_tflops": fma_tf, "tensor_peak_tflops": mma_tf,
                           "reference_flop_per_element": 94000, "closed_form_flop_per_element": 17000,
                           "achieved_tflops_reference_count": 94000 * grid.n_col_eles / (ele_ms * 1e-3) / 1e12,
                           "achieved_tflops_closed_form_count": 17000 * grid.n_col_eles / (ele_ms * 1e-3) / 1e12}
        except Exception as ex:  # measurement helper only
            out["fp64"] = {"error": str(ex)}
    out["clocks"] = sampler.summary()
    if not args.no_e2e:
        out["e2e"] = run_e2e(args, fc, torch, P, flags)
    else:
        out["e2e"] = None
    P.close()

    # ---- C5 record: 400^3 = 64 M elements on 8 GPUs (200^3 per GPU), after the contract measurement ------------
    if (world == 8 and not args.no_c5) or args.c5:
        out["c5"] = run_c5(args, fc, torch, rank, world, local, comm, hbm_peak)

    # ---- cpu_baseline: the oracle port on the host cores, bounded sample, rank 0 at N=1 only -------------
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as O
        cflags = O.use_fast_build()
        cores = os.cpu_count() or 1
        e = args.ref_edge or args.edge
        t_all = time_oracle(O, (e, e, e), cores, 1, 2)
        e1 = min(e, 50)  # one core: about 3 s per evaluation
        t_one = time_oracle(O, (e1, e1, e1), 1, 1, 2)
        march = [f for f in cflags.split() if f.startswith("-march")][0]
        out["cpu_baseline"] = {"value": e ** 3 / t_all, "unit": UNIT, "cores": cores, "kind": "port",
                               "sample": f"{e}^3 hex8 Neo-Hooke cube ({e**3} elements), {cores} threads = {cores} structured subdomains, "
                                         + ("the full workload" if e == args.edge else "bounded sample")
                                         + f", mean of 2 evaluations after 1 warm-up, oracle timing build (-O3 {march}, FMA contraction allowed)",
                               "one_core": {"value": e1 ** 3 / t_one, "unit": UNIT, "cores": 1,
                                            "sample": f"{e1}^3 cube ({e1**3} elements), faithful structure (per-element gather, full 24x24 product, "
                                                      "lower_bound CSR scatter), mean of 2 evaluations after 1 warm-up"}}
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--edge", type=int, default=100, help="hex8 elements per cube edge per GPU (C2: 100)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--ref-edge", type=int, default=None, help="edge of the cube the CPU arm evaluates (default: the GPU arm's mesh, capped at 4 M elements)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--kernel-flags", type=int, default=0, help="FC_EVAL_* kernel selection (0 default = per-element reductions, 8 owner-computed store-only kernel, 4 closed-form FMA, 2 generic)")
    ap.add_argument("--traffic", type=float, default=None, help="dram bytes per launch from the committed ncu capture")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--c5", action="store_true", help="add the C5 record (200^3 per GPU) on any GPU count; default: at --gpus 8 only")
    ap.add_argument("--no-c5", action="store_true")
    ap.add_argument("--no-overlap", action="store_true", help="N > 1: halo import before the evaluation instead of overlapped with the interior elements")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (large --edge: K alone is 15.7 GB of pinned memory per rank at 200)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
