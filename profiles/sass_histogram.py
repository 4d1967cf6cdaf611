#!/usr/bin/env python
"""Opcode histogram of the element kernels in libfourc_b200.so (cuobjdump -sass), written as a markdown table.

  python profiles/sass_histogram.py > profiles/rNN_sass_histogram.md

Counts are STATIC instructions of the kernel's SASS (loops count once).  The columns are the opcodes the roofline discussion
of DESIGN.md refers to: DMMA (FP64 tensor), DFMA/DMUL/DADD (FP64 pipe), RED (FP64 reductions into K and f), LDS/STS (shared
memory), LDG/STG, SHFL, BAR, BSSY/BSYNC (divergence bookkeeping)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "4c_b200", "lib", "libfourc_b200.so")
KERNELS = ["k_hex8_neohooke_mma", "k_hex8_nh_brick", "k_hex8_neohooke<", "k_hex8_mma_ext", "k_nh_mma_gen", "k_tet10_nh", "k_generic", "k_spmv", "k_mg_smooth"]
COLS = ["DMMA", "DFMA", "DMUL", "DADD", "RED", "LDS", "STS", "LDG", "STG", "SHFL", "BAR", "BSSY", "BSYNC", "BRA"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    demangle = lambda n: subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
    hist, name = {}, None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = demangle(m.group(1))
            hist[name] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and name:
            op = m.group(1)
            hist[name][op] += 1
            hist[name]["_total"] += 1
    print("| kernel | total | " + " | ".join(COLS) + " |")
    print("|---|---|" + "---|" * len(COLS))
    for k in sorted(hist):
        short = re.sub(r"^void fc::", "", k)
        short = re.sub(r"\((int|bool)\)", "", short)          # k_x<(int)0, (bool)1>(args) -> k_x<0, 1>
        short = re.sub(r"\(fc::.*$|\(const.*$|\(int.*$|\(double.*$|\(long.*$", "", short)
        if not any(short.startswith(p) or p in short for p in KERNELS):
            continue
        h = hist[k]
        red = h["REDG"] + h["RED"] + h["ATOMG"]
        vals = {c: h[c] for c in COLS}
        vals["RED"] = red
        vals["BAR"] = h["BAR"] + h["WARPSYNC"]
        print(f"| `{short}` | {h['_total']} | " + " | ".join(str(vals[c]) for c in COLS) + " |")


if __name__ == "__main__":
    sys.exit(main())
