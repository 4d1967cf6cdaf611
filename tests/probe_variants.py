"""Driver (not a test): C2 element-kernel time under the FC_DEBUG / FC_DEBUG_EXTRA_SMEM experiment knobs."""
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
cases = [x.split(":") for x in (sys.argv[1:] or ["0:0"])]
for dbg, extra in cases:
    os.environ["FC_DEBUG"] = dbg
    if int(extra):
        os.environ["FC_DEBUG_EXTRA_SMEM"] = extra
    else:
        os.environ.pop("FC_DEBUG_EXTRA_SMEM", None)
    flags = fc.FC_EVAL_ZERO_FIRST
    for _ in range(5):
        s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=flags, want_errors=False)
    torch.cuda.synchronize()
    ks = []
    for _ in range(20):
        s.evaluate(fc.FC_ACTION_NLNSTIFF, u, f, K, flags=flags, want_errors=False)
        ks.append(s.last_kernel_ms()[1])
    print(f"FC_DEBUG={dbg} extra_smem={extra}: element kernel {np.median(ks):.3f} ms (min {min(ks):.3f})  sum|K|={float(K.abs().sum()):.6e}", flush=True)
