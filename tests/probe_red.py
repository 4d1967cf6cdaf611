"""GPU microbenchmark driver (not a test): FP64 reduction / store throughput vs address pattern."""
import ctypes as C
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
fc = importlib.import_module("4c_b200")
L = fc.capi.lib()
print("mode span region_MB  Gops/s  GB/s(payload)")
for region in (64 << 20, 2048 << 20):
    for mode, name in ((0, "red"), (1, "st"), (2, "rmw")):
        for span in (1, 3, 4, 6, 8, 16, 32):
            g = C.c_double()
            iters = 2000 if region < (1 << 30) else 400
            rc = L.fc_probe_red(0, mode, span, C.c_longlong(region), iters, C.byref(g))
            print(f"{name:4s} {span:3d} {region >> 20:6d} {g.value:9.1f} {8 * g.value:9.1f}", flush=True)
