"""Single-caller latency of the host path under the runtime's knobs (EMBED_GPU_TIMING,
EMBED_GPU_PIECES are read by etn_init): C1 request and a 64-row request, p50 of 300 calls."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from llm_mcp_b200 import embedgpu as eg
D, k = 4096, 1024
pi, po = eg.PinnedBuffer(1024 * D * 8), eg.PinnedBuffer(1024 * k * 8)
x, y = pi.array((1024, D), np.float64), po.array((1024, k), np.float64)
x[:, :k] = 0.01
L = eg.lib()
for timing, pieces in ((16, 16), (1, 16), (0, 16), (0, 4), (0, 1)):
    os.environ["EMBED_GPU_TIMING"], os.environ["EMBED_GPU_PIECES"] = str(timing), str(pieces)
    eg.init([0])
    out = []
    for rows in (1024, 64):
        lat = []
        for i in range(320):
            t0 = time.perf_counter()
            L.embed_truncate_norm_host(rows, D, k, eg.F64, 1, x.ctypes.data, D * 8, y.ctypes.data, 1)
            lat.append(time.perf_counter() - t0)
        a = np.sort(np.array(lat[20:])) * 1e6
        out.append("rows=%4d p50 %6.1f us p99 %6.1f us" % (rows, a[len(a) // 2], a[int(len(a) * 0.99)]))
    print(f"timing={timing} pieces={pieces:2d}: " + "   ".join(out), flush=True)
    eg.shutdown()
