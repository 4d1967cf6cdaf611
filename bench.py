#!/usr/bin/env python
"""bench.py — hex8 Neo-Hooke force+stiffness evaluation and assembly (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--edge E] [--impl ours|reference]

A "step" is one struct_calc_nlnstiff evaluation of the whole mesh: zero K and f, (N>1: halo import
of u), element evaluation, scatter into the CSR values.  Workload at N=1: BASELINE config C2
(100^3 hex8 cube, MAT_ElastHyper + ELAST_CoupNeoHooke YOUNG 10 NUE 0.25, 1 evaluation).  N>1: weak
scaling, every rank keeps a 100^3 box of the reference's structured partition.
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


def smooth_displacement(X, h, seed=12345, amp=0.02, noise=1e-3):
    """SURVEY 8d displacement state: affine + sine + uniform noise U(-noise, noise) h, seed 12345"""
    rng = np.random.Generator(np.random.PCG64(seed))
    A = np.array([[0.03, 0.01, -0.02], [0.00, -0.02, 0.015], [0.01, 0.02, 0.025]])
    u = X @ A.T + amp * np.sin(2 * np.pi * X[:, [1, 2, 0]])
    u += rng.uniform(-noise, noise, size=X.shape) * h
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


def run_reference(args):
    """The reference's CPU implementation of the path.  4C itself cannot be built on this box (no
    MPI/Trilinos), so this arm times the CPU restatement in oracle/ ("port"), one thread per
    reference-style structured subdomain (mirrors mpirun -np <cores>), on a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    cores = os.cpu_count() or 1
    e = args.ref_edge
    job = O.Job.cuboid(O.HEX8, (e, e, e), nranks=cores)
    job.set_problem(O.FORM_DISP, O.MAT_NEOHOOKE, YOUNG, NUE)
    job.set_u_nodal(smooth_displacement(job.X, 1.0 / e))
    for _ in range(args.warmup):
        job.evaluate(O.ACTION_NLNSTIFF, max_threads=cores)
    times = []
    for _ in range(args.steps):
        bad, _ = job.evaluate(O.ACTION_NLNSTIFF, max_threads=cores)
        assert bad == 0
        times.append(job.seconds)
    t = float(np.mean(times))
    val = e ** 3 / t
    sample = f"{e}^3 hex8 Neo-Hooke cube ({e**3} elements) per step, {cores} threads = {cores} structured subdomains, evaluate+assemble only"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"C2 hex8 Neo-Hooke cube {args.edge}^3, struct_calc_nlnstiff; reference arm timed on "
                               + ("the full workload" if e == args.edge else f"a {e}^3 sample")},
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
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    # ---- problem setup (not timed) -------------------------------------------------------------------
    n = global_edge(args.edge, world)
    grid = fc.Grid(fc.FC_HEX8, n, nranks=world, rank=rank)
    s = fc.Solid(local, rank, world)
    s.set_stream(torch.cuda.current_stream().cuda_stream)
    s.set_mesh(fc.FC_HEX8, fc.FC_FORM_DISPLACEMENT, grid.X, grid.node_gid, grid.node_owner, grid.conn, grid.ele_gid)
    s.set_material(fc.FC_MAT_ELASTHYPER_COUPNEOHOOKE, YOUNG, NUE)
    n_row_dof, n_col_dof, nnz = s.build_graph()
    u_host = smooth_displacement(grid.X, 1.0 / max(n)).reshape(-1)
    u_col = torch.from_numpy(u_host).to(dev)
    f_row = torch.empty(n_row_dof, dtype=torch.float64, device=dev)
    K = torch.empty(nnz, dtype=torch.float64, device=dev)
    flags = fc.FC_EVAL_ZERO_FIRST | args.kernel_flags

    # halo plan (N>1): owned interface DOFs -> peers; ghosts arrive grouped by owner behind the owned DOFs
    if world > 1:
        send_lids = (3 * grid.send_nodes[:, None] + np.arange(3)[None, :]).reshape(-1)
        s.halo_set_send(send_lids)
        sendbuf = torch.empty(len(send_lids), dtype=torch.float64, device=dev)
        send_split = [3 * int(c) for c in grid.send_count]
        recv_split = [3 * int(c) for c in grid.recv_count]

    main_stream = torch.cuda.current_stream()
    side_stream = torch.cuda.Stream() if world > 1 else None
    overlap = world > 1 and not args.no_overlap and not (args.kernel_flags & (2 | 4))

    def halo():
        if world > 1:
            s.halo_pack(u_col, sendbuf)
            dist.all_to_all_single(u_col[n_row_dof:], sendbuf, recv_split, send_split)

    def step_plain():
        halo()
        s.evaluate(fc.FC_ACTION_NLNSTIFF, u_col, f_row, K, flags=flags, want_errors=False)

    def step():
        if not overlap:
            return step_plain()
        # halo import on a second stream while the elements without ghost nodes are evaluated, then the boundary elements
        side_stream.wait_stream(main_stream)
        with torch.cuda.stream(side_stream):
            s.set_stream(side_stream.cuda_stream)
            s.halo_pack(u_col, sendbuf)
            s.set_stream(main_stream.cuda_stream)
            dist.all_to_all_single(u_col[n_row_dof:], sendbuf, recv_split, send_split)
        s.evaluate(fc.FC_ACTION_NLNSTIFF, u_col, f_row, K, flags=flags | fc.FC_EVAL_INTERIOR, want_errors=False)
        main_stream.wait_stream(side_stream)
        s.evaluate(fc.FC_ACTION_NLNSTIFF, u_col, f_row, K, flags=(flags & ~fc.FC_EVAL_ZERO_FIRST) | fc.FC_EVAL_BOUNDARY, want_errors=False)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    assert s.error_state() == (0, -1), "element errors in the benchmark state"

    # ---- timed region: K steps, CUDA events on the launching stream, max over ranks ----------------
    sampler = ClockSampler(local)
    sampler.start()
    l0 = s.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kernel_ms = []
    time.sleep(0.1)  # let the sampler come up
    barrier()
    sampler.window[0] = time.time()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    sampler.window[1] = time.time()
    total_ms = ev0.elapsed_time(ev1)
    launches = s.launch_count() - l0 + (args.steps if world > 1 else 0)  # + NCCL all-to-all per step
    # dominant-kernel duration: events bracketing the element kernel inside evaluate (ctx stream)
    t_more = time.monotonic()
    while len(kernel_ms) < min(args.steps, 10) or time.monotonic() - t_more < 0.25:
        step_plain()  # one launch over all elements: the dominant kernel's duration for the roofline
        kernel_ms.append(s.last_kernel_ms())
    sampler.stop()
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = float(t.item()) / args.steps
    n_global_ele = n[0] * n[1] * n[2]
    value = n_global_ele / (ms_per_step * 1e-3)
    nnz_global = 9 * (3 * n[0] + 1) * (3 * n[1] + 1) * (3 * n[2] + 1)

    # ---- roofline of the dominant kernel (per launch = this rank's elements) -----------------------
    hbm_peak, peak_src = load_peaks()
    zero_ms = float(np.mean([k[0] for k in kernel_ms]))
    ele_ms = float(np.mean([k[1] for k in kernel_ms]))
    alg_bytes = 8 * nnz + 72 * grid.n_col_nodes + 32 * grid.n_col_eles  # K once, X/u/f once, conn once
    achieved = alg_bytes / (ele_ms * 1e-3) / 1e9
    traffic, traffic_src = (args.traffic, "--traffic") if args.traffic is not None else (None, "capture is of C2 on one GPU; this is another workload")
    if args.traffic is None and args.edge == 100 and world == 1:
        traffic, traffic_src = committed_traffic(args.kernel_flags)
    roofline = {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "kernel": KERNELS.get(args.kernel_flags, (f"flags={args.kernel_flags}", None))[0],
                "kernel_ms": ele_ms, "zero_fill_ms": zero_ms, "algorithmic_bytes_per_launch": alg_bytes}
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
           "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
           "data": "synthetic",
           "config": {"workload": f"C2 hex8 compressible Neo-Hooke cube, {args.edge}^3 elements per GPU (global {n[0]}x{n[1]}x{n[2]}), "
                                  "struct_calc_nlnstiff: zero K,f + halo import + element evaluation + CSR scatter",
                      "elements": n_global_ele, "nnz": nnz_global, "assembled_nnz_per_s": nnz_global / (ms_per_step * 1e-3),
                      "partition": "x".join(str(v) for v in grid.subdivisions), "halo_overlap": bool(overlap), "l2": "inputs+outputs exceed L2 (K values alone are %.2f GB per GPU)" % (8 * nnz / 1e9)},
           "gpu_launches": int(launches), "roofline": roofline}

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

    # ---- e2e: same call with HOST buffers (pinned), H2D of u and D2H of f, K inside the timed region --
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    if args.no_e2e:
        out["e2e"] = None
        out["clocks"] = sampler.summary()
        if rank == 0:
            print(json.dumps(out))
        s.close()
        if world > 1:
            dist.destroy_process_group()
        return
    u_pin = torch.from_numpy(u_host).pin_memory()
    f_pin = torch.empty(n_row_dof, dtype=torch.float64).pin_memory()
    K_pin = torch.empty(nnz, dtype=torch.float64).pin_memory()

    def e2e_step():
        # host path: the library stages u (H2D), evaluates, and copies f and K back (D2H).  N>1: the
        # halo import runs on the device copy first, then the column vector goes through the host call.
        if world > 1:
            halo()
            u_pin.copy_(u_col, non_blocking=False)
        s.evaluate(fc.FC_ACTION_NLNSTIFF, u_pin, f_pin, K_pin, flags=flags, want_errors=True)

    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    te = torch.tensor([(time.perf_counter() - t0) / e2e_steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    out["e2e"] = {"value": n_global_ele / float(te.item()), "unit": UNIT, "h2d_bytes_per_step": int(8 * n_col_dof),
                  "d2h_bytes_per_step": int(8 * (nnz + n_row_dof)), "steps": e2e_steps, "ms_per_step": 1e3 * float(te.item()),
                  "checksum_f": float(f_pin.abs().sum())}
    out["clocks"] = sampler.summary()

    # ---- cpu_baseline: the oracle port on the host cores, bounded sample, rank 0 at N=1 only -------------
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as O
        cores = os.cpu_count() or 1
        e = args.ref_edge
        job = O.Job.cuboid(O.HEX8, (e, e, e), nranks=cores)
        job.set_problem(O.FORM_DISP, O.MAT_NEOHOOKE, YOUNG, NUE)
        job.set_u_nodal(smooth_displacement(job.X, 1.0 / e))
        job.evaluate(O.ACTION_NLNSTIFF, max_threads=cores)
        ts = []
        for _ in range(2):
            job.evaluate(O.ACTION_NLNSTIFF, max_threads=cores)
            ts.append(job.seconds)
        out["cpu_baseline"] = {"value": e ** 3 / float(np.mean(ts)), "unit": UNIT, "cores": cores, "kind": "port",
                               "sample": f"{e}^3 hex8 Neo-Hooke cube ({e**3} elements), {cores} threads = {cores} structured subdomains, "
                                         + ("the full workload" if e == args.edge else "bounded sample") + ", mean of 2 evaluations after 1 warm-up"}
    if rank == 0:
        print(json.dumps(out))
    s.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--edge", type=int, default=100, help="hex8 elements per cube edge per GPU (C2: 100)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--ref-edge", type=int, default=100, help="edge of the cube the CPU arm evaluates (100 = the full C2 workload)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--kernel-flags", type=int, default=0, help="FC_EVAL_* kernel selection (0 default = per-element reductions, 8 owner-computed store-only kernel, 4 closed-form FMA, 2 generic)")
    ap.add_argument("--traffic", type=float, default=None, help="dram bytes per launch from the committed ncu capture")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-overlap", action="store_true", help="N > 1: halo import before the evaluation instead of overlapped with the interior elements")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (large --edge: K alone is 15.7 GB of pinned memory per rank at 200)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
