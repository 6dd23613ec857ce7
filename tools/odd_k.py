"""Times dense float64/float32 renorm for shapes off the 16-byte grid: odd k (prefix not 16-byte
tileable) and odd D (rows that start off a 16-byte boundary).  Both go through the page ring,
staged as the aligned span around the prefix; ETN_PDL etc. as usual."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from llm_mcp_b200 import embedgpu as eg
torch.cuda.set_device(0)
L = eg.lib(); st = torch.cuda.current_stream().cuda_stream
for rows, D, k, tdt, code, es in ((1048576, 4096, 257, torch.float64, eg.F64, 8), (524288, 4096, 1001, torch.float64, eg.F64, 8),
                                  (1048576, 4096, 1000, torch.float64, eg.F64, 8), (1048576, 770, 257, torch.float32, eg.F32, 4),
                                  (1048576, 3072, 1001, torch.float32, eg.F32, 4), (1048576, 3072, 1002, torch.float32, eg.F32, 4),
                                  # rows off the 16-byte grid: odd D
                                  (1000000, 771, 257, torch.float32, eg.F32, 4), (1000000, 771, 257, torch.float64, eg.F64, 8),
                                  (1000000, 771, 256, torch.float32, eg.F32, 4), (500000, 4097, 1024, torch.float64, eg.F64, 8)):
    x = torch.empty((rows, D), dtype=tdt, device="cuda")
    eg.check(L.etn_fill_synthetic(3, 0, rows, D, D, D, code, x.data_ptr(), st), "fill")
    y = torch.empty((rows, k), dtype=tdt, device="cuda")
    for _ in range(3):
        eg.check(L.embed_truncate_norm(rows, D, k, code, x.data_ptr(), y.data_ptr(), st), "renorm")
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        eg.check(L.embed_truncate_norm(rows, D, k, code, x.data_ptr(), y.data_ptr(), st), "renorm")
    e1.record(); torch.cuda.synchronize()
    t = e0.elapsed_time(e1) * 1e-3 / 20
    print("%d x %d -> %d %s: %.3f ms  %.0f GB/s" % (rows, D, k, str(tdt)[6:], t * 1e3, rows * 2 * k * es / t / 1e9), flush=True)
    del x, y
