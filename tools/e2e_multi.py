"""End-to-end host path with 1..N devices inside ONE process (embed_truncate_norm_host's own
row sharding: one host thread and four streams per device), C3-shaped rows."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from llm_mcp_b200 import embedgpu as eg
nd_all = eg.init()
rows_per, D, k = 524288, 4096, 1024
for nd in [n for n in (1, 2, 4, 8) if n <= nd_all]:
    rows = rows_per * nd
    pin_in, pin_out = eg.PinnedBuffer(rows * D * 4), eg.PinnedBuffer(rows * k * 4)
    x, y = pin_in.array((rows, D), np.float32), pin_out.array((rows, k), np.float32)
    x[:, :k] = 0.01
    for _ in range(2):
        eg.truncate_norm_host(x, k, out=y, n_devices=nd)
    t0 = time.perf_counter()
    for _ in range(4):
        eg.truncate_norm_host(x, k, out=y, n_devices=nd)
    t = (time.perf_counter() - t0) / 4
    print("in-process n_devices=%d: %d rows in %.1f ms -> %.2f M vectors/s, %.1f GB/s each way in total" % (nd, rows, t * 1e3, rows / t / 1e6, rows * k * 4 / t / 1e9), flush=True)
    del x, y
    pin_in.free(); pin_out.free()
eg.shutdown()
