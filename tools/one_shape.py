"""One dense renorm launch of a given shape (for ncu): python tools/one_shape.py rows D k f32|f64 [reps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from llm_mcp_b200 import embedgpu as eg
rows, D, k = (int(v) for v in sys.argv[1:4])
f64 = sys.argv[4] == "f64"
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
tdt, code = (torch.float64, eg.F64) if f64 else (torch.float32, eg.F32)
torch.cuda.set_device(0)
L = eg.lib(); st = torch.cuda.current_stream().cuda_stream
x = torch.empty((rows, D), dtype=tdt, device="cuda")
eg.check(L.etn_fill_synthetic(3, 0, rows, D, D, D, code, x.data_ptr(), st), "fill")
y = torch.empty((rows, min(k, D)), dtype=tdt, device="cuda")
for _ in range(reps):
    eg.check(L.embed_truncate_norm(rows, D, k, code, x.data_ptr(), y.data_ptr(), st), "renorm")
torch.cuda.synchronize()
print("ok", rows, D, k, sys.argv[4])
