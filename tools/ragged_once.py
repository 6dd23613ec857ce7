"""Runs the C5 ragged float64 renorm a few times (for ncu).  `python tools/ragged_once.py c1x64`
runs the gather launch over 64 merged C1 requests (65,536 rows of 8 KiB) instead."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from llm_mcp_b200 import embedgpu as eg
torch.cuda.set_device(0)
rng = np.random.default_rng(5)
merged = len(sys.argv) > 1 and sys.argv[1] == "c1x64"
rows, D = (65536, 4096) if merged else (1048576, 4096)
parts, n = [], 0
while n < rows:
    m = min(int(rng.integers(1, 2049)), rows - n)
    parts.append(np.full(m, 1024 if merged else rng.choice([64, 128, 256, 512, 1024]), dtype=np.int32)); n += m
kpr = np.concatenate(parts)
off = np.zeros(rows + 1, dtype=np.int64); np.cumsum(kpr, out=off[1:])
L = eg.lib(); st = torch.cuda.current_stream().cuda_stream
x = torch.empty((rows, D), dtype=torch.float64, device="cuda")
eg.check(L.etn_fill_synthetic(5, 0, rows, D, D, D, eg.F64, x.data_ptr(), st), "fill")
dk, do = torch.from_numpy(kpr).cuda(), torch.from_numpy(off[:-1].copy()).cuda()
out = torch.empty(int(off[-1]), dtype=torch.float64, device="cuda")
di = torch.arange(rows, dtype=torch.int64, device="cuda") * D
for _ in range(4):
    if merged:
        eg.check(L.embed_truncate_norm_gather(rows, di.data_ptr(), dk.data_ptr(), do.data_ptr(), eg.F64, 1, x.data_ptr(), out.data_ptr(), st), "gather")
    else:
        eg.check(L.embed_truncate_norm_ragged(rows, D, dk.data_ptr(), do.data_ptr(), eg.F64, 1, x.data_ptr(), out.data_ptr(), st), "ragged")
torch.cuda.synchronize()
print("ok", float(out[:10].sum()), "algorithmic bytes per launch", 16 * int(off[-1]))
