#!/usr/bin/env python
"""Summarise `ncu --set full` captures (gpurun_out/*.ncu-rep) into the tracked profiles/ directory.

  python profiles/extract_ncu.py gpurun_out/prof_X.ncu-rep profiles/rNN_X [--units N --unit-name elements]

Writes <out>.json: the metrics the roofline uses (dram bytes, duration, pipe utilisation, instruction counts), the
kernel name and the sha256 of the kernel's source files at the time of extraction.  bench.py reads `traffic` from these
files and reports null when the hash of the source no longer matches (the capture is then of an older kernel)."""
import argparse
import csv
import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__grid_size", "launch__block_size",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
    "lts__t_sectors_srcunit_tex_op_red.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed_op_global_red.sum",
    "smsp__warps_eligible.avg.per_cycle_active", "smsp__warps_active.avg.per_cycle_active",
    "smsp__inst_executed_pipe_fp64.sum", "smsp__inst_executed_pipe_lsu.sum", "smsp__inst_executed_pipe_tensor_op_dmma.sum",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
]
UNIT_SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0, "msecond": 1e-3, "usecond": 1e-6, "nsecond": 1e-9, "second": 1.0}


def sha_of(files):
    h = hashlib.sha256()
    for f in files:
        with open(os.path.join(ROOT, f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("out")
    ap.add_argument("--units", type=float, default=None, help="work units of the captured launch (elements)")
    ap.add_argument("--algorithmic-bytes", type=float, default=None)
    ap.add_argument("--sources", nargs="*", default=[], help="kernel source files (repo-relative) whose hash keys the capture")
    ap.add_argument("--command", default="")
    ap.add_argument("--launch", type=int, default=0, help="which captured launch (row) to summarise")
    args = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", args.rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, row = rows[0], rows[1], rows[2 + args.launch]
    col = {h: i for i, h in enumerate(hdr)}
    out = {"kernel": row[col["Kernel Name"]], "report": os.path.basename(args.rep), "command": args.command, "metrics": {}}
    for m in METRICS:
        if m in col and row[col[m]] != "":
            try:
                out["metrics"][m] = {"value": float(row[col[m]].replace(",", "")), "unit": units[col[m]]}
            except ValueError:
                pass
    g = lambda m: out["metrics"][m]["value"] * UNIT_SCALE.get(out["metrics"][m]["unit"], 1.0)
    out["duration_s"] = g("gpu__time_duration.sum")
    out["dram_bytes"] = g("dram__bytes_read.sum") + g("dram__bytes_write.sum")
    if args.units:
        out["units"] = args.units
        out["warp_instructions_per_unit"] = out["metrics"]["smsp__inst_executed.sum"]["value"] / args.units
        out["dram_bytes_per_unit"] = out["dram_bytes"] / args.units
    if args.algorithmic_bytes:
        out["algorithmic_bytes"] = args.algorithmic_bytes
        out["dram_over_algorithmic"] = out["dram_bytes"] / args.algorithmic_bytes
    if args.sources:
        out["sources"] = args.sources
        out["sources_sha16"] = sha_of(args.sources)
    with open(args.out + ".json", "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps({k: v for k, v in out.items() if k != "metrics"}))


if __name__ == "__main__":
    main()
