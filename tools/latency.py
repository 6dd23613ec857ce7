"""Per-call latency of the host path for C1-sized requests (1,024 x 4096 float64 -> 1024, 8 MiB
each way) with 1 / 8 / 64 concurrent callers, on 1 device and on every device of the box.

    gpurun [--gpus 8] -- python tools/latency.py > gpurun_out/latency.txt

Every caller is a thread in a blocking embed_truncate_norm_host call with its own pinned
buffers — the reference's calling pattern, one goroutine per connection
(core/internal/api/server.go:54).  ctypes drops the GIL for the length of a call, so the
callers really are concurrent.  Also a small request (64 x 4096 -> 1024) for the fixed costs."""
import os
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from llm_mcp_b200 import embedgpu as eg  # noqa: E402


def run(n_callers, calls_each, rows, D, k, bufs):
    lat = [[] for _ in range(n_callers)]
    start = threading.Barrier(n_callers + 1)

    def caller(i):
        x, y = bufs[i]
        eg.truncate_norm_host(x[:rows], k, out=y[:rows])
        start.wait()
        for _ in range(calls_each):
            t0 = time.perf_counter()
            eg.truncate_norm_host(x[:rows], k, out=y[:rows])
            lat[i].append(time.perf_counter() - t0)

    th = [threading.Thread(target=caller, args=(i,)) for i in range(n_callers)]
    [t.start() for t in th]
    start.wait()
    t0 = time.perf_counter()
    [t.join() for t in th]
    wall = time.perf_counter() - t0
    a = np.sort(np.concatenate([np.array(v) for v in lat])) * 1e6
    return {"p50_us": float(a[len(a) // 2]), "p99_us": float(a[min(len(a) - 1, int(len(a) * 0.99))]), "max_us": float(a[-1]),
            "requests_per_s": n_callers * calls_each / wall, "GBps_each_way": n_callers * calls_each * rows * k * 8 / wall / 1e9}


def main():
    import torch
    ndev_all = torch.cuda.device_count()
    D, k, R = 4096, 1024, 1024
    pins, bufs = [], []
    for _ in range(64):
        pi, po = eg.PinnedBuffer(R * D * 8), eg.PinnedBuffer(R * k * 8)
        x, y = pi.array((R, D), np.float64), po.array((R, k), np.float64)
        x[:, :k] = 0.01
        pins += [pi, po]
        bufs.append((x, y))
    print(f"# host-path latency, {ndev_all} GPU(s) visible; C1 request = {R} x {D} float64 -> {k} (8 MiB each way); small = 64 rows")
    for ndev in sorted({1, ndev_all}):
        eg.init(list(range(ndev)))
        for rows, label in ((R, "C1"), (64, "small")):
            for callers in (1, 8, 64):
                r = run(callers, max(4, 256 // callers) if rows == R else max(20, 2000 // callers), rows, D, k, bufs)
                print(f"devices={ndev} request={label:5s} callers={callers:2d}  p50 {r['p50_us']:9.1f} us  p99 {r['p99_us']:9.1f} us  max {r['max_us']:9.1f} us  "
                      f"{r['requests_per_s']:9.0f} req/s  {r['GBps_each_way']:6.1f} GB/s each way", flush=True)
        c = eg.counters()
        h = eg.histogram(eg.HIST_REQS_PER_LAUNCH)
        print(f"# devices={ndev}: coalesced requests so far {c['llmcore_embed_gpu_coalesced_requests_total']}, chunks by requests-per-launch "
              f"{dict(zip(['<=1', '<=2', '<=4', '<=8', '<=16', '<=32', '<=64', '<=128', 'more'], h['counts']))}", flush=True)
        eg.shutdown()


if __name__ == "__main__":
    main()
