"""y = K x at C2 size for the kernel variants (FC_SPMV_VARIANT / FC_SPMV_BS): ms and algorithmic GB/s"""
import importlib, json, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
fc = importlib.import_module("4c_b200")
e = int(sys.argv[1]) if len(sys.argv) > 1 else 100
g = fc.Grid(fc.FC_HEX8, (e, e, e))
s = fc.Solid(0)
s.set_mesh(fc.FC_HEX8, 0, g.X, g.node_gid, g.node_owner, g.conn, g.ele_gid)
s.set_material(1, 10.0, 0.25)
n_row, n_col, nnz = s.build_graph()
dev = torch.device("cuda", 0)
K = torch.rand(nnz, dtype=torch.float64, device=dev)
x = torch.rand(n_col, dtype=torch.float64, device=dev)
y = torch.empty(n_row, dtype=torch.float64, device=dev)
s.set_stream(torch.cuda.current_stream().cuda_stream)
for _ in range(5):
    s.spmv(K, x, y)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50):
    s.spmv(K, x, y)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 50
b = 8 * nnz + 4 * (nnz // 9) + 16 * n_row
print(json.dumps({"variant": os.environ.get("FC_SPMV_VARIANT", "0"), "bs": os.environ.get("FC_SPMV_BS", "256"), "ms": ms, "GBps": b / ms / 1e6, "checksum": float(y.sum())}))
