"""Driver (not a test): step time of C2 vs the slab count of the zero-fill pipeline."""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import smooth_displacement
fc = importlib.import_module("4c_b200")
n = (100, 100, 100)
g = fc.Grid(fc.FC_HEX8, n)
s = fc.Solid(0)
s.set_stream(torch.cuda.current_stream().cuda_stream)
s.set_mesh(fc.FC_HEX8, fc.FC_FORM_DISPLACEMENT, g.X, g.node_gid, g.node_owner, g.conn, g.ele_gid)
s.set_material(fc.FC_MAT_ELASTHYPER_COUPNEOHOOKE, 10.0, 0.25)
n_row, n_col, nnz = s.build_graph()
dev = torch.device("cuda", 0)
u = torch.from_numpy(smooth_displacement(g.X, 0.01).reshape(-1)).to(dev)
K = torch.empty(nnz, dtype=torch.float64, device=dev)
f = torch.empty(n_row, dtype=torch.float64, device=dev)
for ns, fl in [(1, 0), (2, 0), (3, 0), (4, 0), (8, 0), (16, 0), (8, fc.FC_EVAL_NO_SLAB_PIPELINE)]:
    s.set_slab_count(ns)
    flags = fc.FC_EVAL_ZERO_FIRST | fl
    for _ in range(5):
        s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=flags, want_errors=False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(30):
        s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=flags, want_errors=False)
    e1.record(); torch.cuda.synchronize()
    print(ns, fl, "ms/step", e0.elapsed_time(e1) / 30, s.last_kernel_ms(), float(K.abs().sum()))
