"""Times the request-sized shapes (C1, C2) and C3-f64 through the C-ABI; run with ETN_PDL=0/1."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from llm_mcp_b200 import embedgpu as eg
torch.cuda.set_device(0)
peak, _ = bench.measured_peak()
for name in sys.argv[1:] or ["c1", "c2", "c4", "c3f64"]:
    r = bench.quick_measure(torch, eg, bench.WORKLOADS[name], peak, steps=200, warmup=20)
    print(os.environ.get("ETN_PDL", "default"), name, "%.2f us  %.0f GB/s  %.3g vec/s" % (r["us_per_launch"], r["achieved_GBps"], r["vectors_per_s"]), flush=True)
    torch.cuda.empty_cache()

# C5: ragged mixed-request batch, float64, D = 4096, per-request k in {64,...,1024}
import numpy as np
rng = np.random.default_rng(5)
rows, D = 1048576, 4096
parts, n = [], 0
while n < rows:
    m = min(int(rng.integers(1, 2049)), rows - n)
    parts.append(np.full(m, rng.choice([64, 128, 256, 512, 1024]), dtype=np.int32)); n += m
kpr = np.concatenate(parts)
off = np.zeros(rows + 1, dtype=np.int64); np.cumsum(kpr, out=off[1:])
L = eg.lib(); st = torch.cuda.current_stream().cuda_stream
x = torch.empty((rows, D), dtype=torch.float64, device="cuda")
eg.check(L.etn_fill_synthetic(5, 0, rows, D, D, D, eg.F64, x.data_ptr(), st), "fill")
dk, do = torch.from_numpy(kpr).cuda(), torch.from_numpy(off[:-1].copy()).cuda()
out = torch.empty(int(off[-1]), dtype=torch.float64, device="cuda")
for mode in (1, 0):
    for _ in range(5):
        eg.check(L.embed_truncate_norm_ragged(rows, D, dk.data_ptr(), do.data_ptr(), eg.F64, mode, x.data_ptr(), out.data_ptr(), st), "ragged")
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        eg.check(L.embed_truncate_norm_ragged(rows, D, dk.data_ptr(), do.data_ptr(), eg.F64, mode, x.data_ptr(), out.data_ptr(), st), "ragged")
    e1.record(); torch.cuda.synchronize()
    t = e0.elapsed_time(e1) * 1e-3 / 50
    print("c5 ragged f64 mode=%d mean k=%.1f: %.3f ms  %.0f GB/s  %.3g vec/s" % (mode, kpr.mean(), t * 1e3, 2 * 8 * int(off[-1]) / t / 1e9, rows / t), flush=True)

# request-sized end-to-end latency through the host entry points (pinned buffers)
import time
eg.init([0])
for rows_, D_, k_, dt_ in ((1024, 4096, 1024, np.float64), (64, 4096, 1024, np.float64), (2048, 3072, 256, np.float32)):
    pin_in, pin_out = eg.PinnedBuffer(rows_ * D_ * np.dtype(dt_).itemsize), eg.PinnedBuffer(rows_ * k_ * np.dtype(dt_).itemsize)
    hx, hy = pin_in.array((rows_, D_), dt_), pin_out.array((rows_, k_), dt_)
    hx[:] = 0.01
    for fn, label in ((lambda: eg.truncate_norm_host(hx, k_, 1, out=hy, n_devices=1), "host"),
                      (lambda: L.embed_truncate_norm_host_static(rows_, D_, k_, eg.dtype_code(dt_), 1, hx.ctypes.data, 0, hy.ctypes.data, None), "host_static")):
        for _ in range(5):
            fn()
        t0 = time.perf_counter()
        for _ in range(50):
            fn()
        t = (time.perf_counter() - t0) / 50
        print("e2e %-11s %5d x %d -> %d %s: %.1f us per call (%.1f GB/s each way)" % (label, rows_, D_, k_, np.dtype(dt_).name, t * 1e6, rows_ * k_ * np.dtype(dt_).itemsize / t / 1e9), flush=True)
    del hx, hy
    pin_in.free(); pin_out.free()
eg.shutdown()

# C2 and C1 replayed from a CUDA graph of 24 launches over rotating buffers
for name in ("c2", "c1"):
    wl = bench.WORKLOADS[name]
    dt = bench.NPDT[wl["dtype"]]; tdt = torch.float32 if wl["dtype"] == "f32" else torch.float64
    es = np.dtype(dt).itemsize; rows_, D_, k_, code = wl["rows"], wl["D"], wl["k"], eg.dtype_code(dt)
    nset = max(6, -(-4 * bench.L2_BYTES // (rows_ * 2 * k_ * es)))
    xs = [torch.empty((rows_, D_), dtype=tdt, device="cuda") for _ in range(nset)]
    ys = [torch.empty((rows_, k_), dtype=tdt, device="cuda") for _ in range(nset)]
    for i, x in enumerate(xs):
        eg.check(L.etn_fill_synthetic(3 + i, 0, rows_, D_, D_, D_, code, x.data_ptr(), torch.cuda.current_stream().cuda_stream), "fill")
    eg.check(L.embed_truncate_norm(rows_, D_, k_, code, xs[0].data_ptr(), ys[0].data_ptr(), torch.cuda.current_stream().cuda_stream), "warm")
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        s_ = torch.cuda.current_stream().cuda_stream
        for i in range(24):
            eg.check(L.embed_truncate_norm(rows_, D_, k_, code, xs[i % nset].data_ptr(), ys[i % nset].data_ptr(), s_), "cap")
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    t = e0.elapsed_time(e1) * 1e-3 / (20 * 24)
    print("graph replay %s: %.2f us per launch  %.0f GB/s" % (name, t * 1e6, rows_ * 2 * k_ * es / t / 1e9), flush=True)
    del xs, ys
