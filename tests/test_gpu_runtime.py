"""-m gpu tests of the host runtime added in round 2 (llm-mcp_b200/csrc/etn_host.cu): the
ragged host entry point, the coalescing of concurrent small requests into one launch, the
ticket API, the gather launch on device pointers, and the histograms.

Reference anchors: dims is resolved per request (core/internal/api/handlers.go:1874-1877) and
every vector is cut on its own (handlers.go:2068-2076), so a batch that merges several
responses is ragged by construction; handlers run one goroutine per connection
(server.go:54), so several of them arrive at the library at once.
The checker is oracle/ (binary128 renormalisation, strided copy for truncation)."""
import ctypes
import os
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

from llm_mcp_b200 import embedgpu as eg  # noqa: E402
from tests.gpu_util import to_dev  # noqa: E402


@pytest.fixture(scope="module", autouse=True)
def _init():
    # a worker that holds one small request waits up to 3 ms for company: makes the merging
    # of concurrent requests deterministic enough to assert on
    os.environ["EMBED_GPU_MAX_WAIT_US"] = "3000"
    os.environ["EMBED_GPU_TIMING"] = "1"             # time every chunk: the histogram test counts them
    eg.shutdown()
    assert eg.init() >= 1
    yield
    eg.shutdown()
    os.environ.pop("EMBED_GPU_MAX_WAIT_US", None)
    os.environ.pop("EMBED_GPU_TIMING", None)


@pytest.fixture
def one_device():
    """Requests only queue up behind each other — and get merged — when they meet on one device;
    with several devices bound the library spreads them instead.  The merging tests bind one."""
    eg.shutdown()
    assert eg.init([0]) == 1
    yield
    eg.shutdown()
    assert eg.init() >= 1


def _ulp(oracle, got, want, bar=2):
    d, i = oracle.max_ulp(got, want)
    assert 0 <= d <= bar, (d, i)


def _check_ragged(oracle, x, kpr, n_devices=0):
    xc = np.ascontiguousarray(x)
    for mode in (eg.TRUNCATE_RENORM, eg.TRUNCATE):
        want, off = oracle.ragged(xc, kpr, mode)
        got = eg.truncate_norm_ragged_host(x, kpr, mode, n_devices=n_devices)
        assert got.shape == want.shape == (int(off[-1]),)
        if mode == eg.TRUNCATE:
            assert np.array_equal(got, want)
        else:
            _ulp(oracle, got, want)


# ------------------------------------------------------------------ ragged host ----
@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_ragged_host_against_oracle(oracle, dt):
    rng = np.random.default_rng(3)
    D = 768
    x = oracle.fill(31, 4000, D, dt)
    # requests of 1..300 vectors, one k each; odd k, k == D, k > D, k == 1
    parts, n = [], 0
    while n < 4000:
        m = min(int(rng.integers(1, 301)), 4000 - n)
        parts.append(np.full(m, rng.choice([1, 64, 100, 256, 257, 512, 767, 768, 900, 3])))
        n += m
    kpr = np.concatenate(parts).astype(np.int32)
    x[17] = 0                                        # all-zero vectors are copied, never NaN
    x[18, :3] = [0.0, -0.0, 0.0]
    _check_ragged(oracle, x, kpr)
    # k that changes on every row (blocks of mixed width)
    kpr2 = rng.integers(1, 400, size=4000).astype(np.int32)
    _check_ragged(oracle, x, kpr2)
    # pitched input: D columns of a wider matrix
    wide = oracle.fill(32, 500, 1000, dt)
    _check_ragged(oracle, wide[:, :D], kpr[:500])
    # degenerate batches
    _check_ragged(oracle, x[:1], np.array([300], dtype=np.int32))
    assert eg.truncate_norm_ragged_host(x[:0], np.zeros(0, dtype=np.int32)).size == 0
    # k <= 0 is refused like in the dense calls (the handler only calls with dims > 0, handlers.go:1891)
    with pytest.raises(eg.EtnError) as ei:
        eg.truncate_norm_ragged_host(x[:5], np.array([4, 4, 0, 4, 4], dtype=np.int32))
    assert ei.value.code == eg.ERR_SHAPE


def test_ragged_host_large_multi_chunk_sharded_by_kept_bytes(oracle):
    """C5 in small: float64, D = 4096, per-request k in {64,...,1024}; 160 MiB of kept prefixes,
    so several chunks per device, cut over min(2, devices) devices by kept bytes."""
    rng = np.random.default_rng(5)
    rows, D = 48000, 4096
    parts, n = [], 0
    while n < rows:
        m = min(int(rng.integers(1, 2049)), rows - n)
        parts.append(np.full(m, rng.choice([64, 128, 256, 512, 1024]), dtype=np.int32))
        n += m
    kpr = np.concatenate(parts)
    pin = eg.PinnedBuffer(rows * D * 8)
    x = pin.array((rows, D), np.float64)
    x[:] = oracle.fill(5, rows, D, np.float64)
    nd = min(2, eg.lib().etn_device_count())
    want, off = oracle.ragged(x, kpr, eg.TRUNCATE_RENORM)
    for n_devices in sorted({1, nd}):
        got = eg.truncate_norm_ragged_host(x, kpr, eg.TRUNCATE_RENORM, n_devices=n_devices)
        _ulp(oracle, got, want)
    # what each shard moves is balanced to within one request's rows
    sh = eg.shard_ragged(kpr, D, max(nd, 2))
    loads = [int(off[b] - off[a]) for a, b in sh]
    assert max(loads) - min(loads) <= 1024
    del x
    pin.free()


# ------------------------------------------------------------------- coalescing ----
def test_concurrent_small_requests_share_launches(oracle, one_device):
    """Sixteen C1-like requests submitted together: the worker merges the queued ones into
    shared chunks (one H2D set, ONE kernel, one D2H set) — same k for all, so the dense
    kernel runs on the merged matrix."""
    xs = [oracle.fill(60 + i, 256, 2048, np.float64) for i in range(16)]
    wants = [oracle.renorm_exact(x, 512) for x in xs]
    before, h0 = eg.counters(), eg.histogram(eg.HIST_REQS_PER_LAUNCH)
    tickets = [eg.Ticket(x, 512) for x in xs]
    outs = [t.wait() for t in tickets]
    after, h1 = eg.counters(), eg.histogram(eg.HIST_REQS_PER_LAUNCH)
    for got, want in zip(outs, wants):
        _ulp(oracle, got, want)
    launches = after["llmcore_embed_gpu_kernel_launches_total"] - before["llmcore_embed_gpu_kernel_launches_total"]
    merged = after["llmcore_embed_gpu_coalesced_requests_total"] - before["llmcore_embed_gpu_coalesced_requests_total"]
    assert merged >= 8 and launches < 16, (merged, launches)
    chunks = h1["count"] - h0["count"]
    assert chunks == launches                        # one kernel per staged chunk
    # the histogram saw at least one chunk with more than one request in it
    assert sum(h1["counts"][1:]) - sum(h0["counts"][1:]) >= 1


def test_concurrent_requests_with_different_k_and_D_share_a_launch(oracle, one_device):
    """Requests of different models and different `dimensions` (handlers.go:1874-1877) in one
    launch: per-row metadata and the gathering page-ring kernel.  Odd k and pitched rows too."""
    shapes = [(300, 768, 256), (64, 4096, 1024), (500, 1024, 100), (33, 3072, 257), (200, 768, 768), (90, 1536, 5000),
              (128, 4096, 64), (77, 1000, 333)]
    for dt in (np.float32, np.float64):
        xs = [oracle.fill(80 + i, r, D + 8, dt)[:, :D] for i, (r, D, k) in enumerate(shapes)]
        before = eg.counters()
        tickets = [eg.Ticket(x, k) for x, (_, _, k) in zip(xs, shapes)]
        outs = [t.wait() for t in tickets]
        after = eg.counters()
        for got, x, (_, D, k) in zip(outs, xs, shapes):
            assert got.shape == (x.shape[0], min(k, D))
            _ulp(oracle, got, oracle.renorm_exact(np.ascontiguousarray(x), k))
        assert after["llmcore_embed_gpu_coalesced_requests_total"] > before["llmcore_embed_gpu_coalesced_requests_total"]
        # truncation of a merged batch: bit-exact
        tickets = [eg.Ticket(x, k, eg.TRUNCATE) for x, (_, _, k) in zip(xs, shapes)]
        for t, x, (_, D, k) in zip(tickets, xs, shapes):
            assert np.array_equal(t.wait(), x[:, :min(k, D)])


def test_many_handler_threads_blocking_calls(oracle):
    """The reference's calling pattern: many goroutines, each in a blocking call."""
    inputs = [oracle.fill(140 + t, 200 + 31 * t, 1024, np.float32) for t in range(12)]
    wants = [oracle.renorm_exact(x, 256) for x in inputs]
    errs = []

    def handler(t):
        try:
            for _ in range(6):
                _ulp(oracle, eg.truncate_norm_host(inputs[t], 256), wants[t])
        except Exception as e:  # noqa: BLE001
            errs.append((t, repr(e)))

    th = [threading.Thread(target=handler, args=(t,)) for t in range(12)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs


def test_tickets_survive_shutdown_and_unknown_tickets_are_refused(oracle):
    x = oracle.fill(7, 3000, 1024, np.float32)
    want = oracle.renorm_exact(x, 256)
    tickets = [eg.Ticket(x, 256) for _ in range(4)]
    eg.shutdown()                                    # calls in flight keep their workers alive
    assert eg.lib().etn_device_count() == 0
    for t in tickets:
        _ulp(oracle, t.wait(), want)
    assert eg.lib().etn_wait(tickets[0].ticket) == eg.ERR_TICKET
    assert eg.init() >= 1                            # and the library can be brought up again
    _ulp(oracle, eg.truncate_norm_host(x, 256), want)


# ------------------------------------------------------ gather on device pointers ----
@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_gather_launch_on_device_pointers(oracle, dt):
    """embed_truncate_norm_gather: three requests that live in one device allocation at
    16-byte aligned element offsets, cut to their own k, in one launch."""
    es = np.dtype(dt).itemsize
    reqs = [(oracle.fill(90, 100, 768, dt), 256), (oracle.fill(91, 40, 4096, dt), 1024), (oracle.fill(92, 77, 1000, dt), 333)]
    flat, in_off, kpr, out_off, pos, opos = [], [], [], [], 0, 0
    for x, k in reqs:
        rows, D = x.shape
        pitch = -(-D * es // 16) * 16 // es          # row starts on 16-byte boundaries
        buf = np.zeros((rows, pitch), dtype=dt)
        buf[:, :D] = x
        flat.append(buf.ravel())
        for r in range(rows):
            in_off.append(pos + r * pitch)
            kpr.append(min(k, D))
            out_off.append(opos)
            opos += min(k, D)
        pos += rows * pitch
    d_in = to_dev(np.concatenate(flat))
    d_out = torch.zeros(opos, dtype=d_in.dtype, device="cuda")
    d_io, d_k, d_oo = to_dev(np.array(in_off, dtype=np.int64)), to_dev(np.array(kpr, dtype=np.int32)), to_dev(np.array(out_off, dtype=np.int64))
    st = torch.cuda.current_stream().cuda_stream
    for mode in (eg.TRUNCATE_RENORM, eg.TRUNCATE):
        eg.check(eg.lib().embed_truncate_norm_gather(len(kpr), d_io.data_ptr(), d_k.data_ptr(), d_oo.data_ptr(), eg.dtype_code(dt),
                                                     mode, d_in.data_ptr(), d_out.data_ptr(), st), "gather")
        torch.cuda.synchronize()
        got, o = d_out.cpu().numpy(), 0
        for x, k in reqs:
            kk = min(k, x.shape[1])
            part = got[o:o + x.shape[0] * kk].reshape(x.shape[0], kk)
            if mode == eg.TRUNCATE:
                assert np.array_equal(part, x[:, :kk])
            else:
                _ulp(oracle, part, oracle.renorm_exact(x, k))
            o += x.shape[0] * kk


# ------------------------------------------------------------------ histograms ----
def test_histogram_counts_add_up_to_the_counters(oracle):
    """N3: every host-path call lands in the batch-rows and call-latency histograms, every
    staged chunk in the copy histograms, every kernel launch in the kernel-time histogram."""
    H = (eg.HIST_BATCH_ROWS, eg.HIST_KERNEL_SECONDS, eg.HIST_H2D_SECONDS, eg.HIST_D2H_SECONDS, eg.HIST_CALL_SECONDS,
         eg.HIST_REQS_PER_LAUNCH)
    h0, c0 = [eg.histogram(w) for w in H], eg.counters()
    x = oracle.fill(1, 20000, 1024, np.float32)
    for rows in (1, 300, 20000):
        eg.truncate_norm_host(x[:rows], 256)
    eg.truncate_norm_host(x[:500], 256, eg.TRUNCATE)             # copies only: no kernel sample
    h1, c1 = [eg.histogram(w) for w in H], eg.counters()
    d = [b["count"] - a["count"] for a, b in zip(h0, h1)]
    launches = c1["llmcore_embed_gpu_kernel_launches_total"] - c0["llmcore_embed_gpu_kernel_launches_total"]
    calls = c1["llmcore_embed_gpu_host_calls_total"] - c0["llmcore_embed_gpu_host_calls_total"]
    assert calls == 4 and d[0] == 4 and d[4] == 4
    assert d[1] == launches >= 3                                 # kernel samples == kernels launched
    assert d[2] == d[3] == d[5] == launches + 1                  # chunks: one per launch + the copy-only one
    rows_hist = h1[0]
    assert rows_hist["sum"] - h0[0]["sum"] == 1 + 300 + 20000 + 500
    ksum = h1[1]["sum"] - h0[1]["sum"]
    assert 0 < ksum < 0.05                                       # seconds of kernel time, from CUDA events
    assert "llmcore_embed_gpu_kernel_seconds_bucket" in eg.metrics_text()


# ---------------------------------------------------------------- stress, lifetime ----
def test_mixed_traffic_from_many_threads(oracle):
    """Sixteen handler threads for a few seconds, each issuing its own mix of dense calls, ragged
    calls, tickets and plain truncation with its own shapes, while the workers merge, split and
    pipeline whatever is queued.  Every caller must get exactly its own rows back."""
    import time
    rng0 = np.random.default_rng(2026)
    shapes = [(int(rng0.integers(1, 3000)), int(rng0.choice([768, 1024, 1536, 4096])), int(rng0.choice([64, 100, 256, 257, 1024, 5000])),
               rng0.choice([np.float32, np.float64])) for _ in range(16)]
    inputs = [oracle.fill(300 + t, r, D, dt) for t, (r, D, k, dt) in enumerate(shapes)]
    wants = [oracle.renorm_exact(x, k) for x, (_, _, k, _) in zip(inputs, shapes)]
    kprs = [np.random.default_rng(t).integers(1, D + 50, size=r).astype(np.int32) for t, (r, D, _, _) in enumerate(shapes)]
    rag_wants = [oracle.ragged(x, kp, eg.TRUNCATE_RENORM)[0] for x, kp in zip(inputs, kprs)]
    errs, calls = [], [0] * 16
    deadline = time.time() + 3.0

    def handler(t):
        x, (_, D, k, _) = inputs[t], shapes[t]
        try:
            i = 0
            while time.time() < deadline:
                kind = (i + t) % 4
                if kind == 0:
                    _ulp(oracle, eg.truncate_norm_host(x, k), wants[t])
                elif kind == 1:
                    _ulp(oracle, eg.truncate_norm_ragged_host(x, kprs[t]), rag_wants[t])
                elif kind == 2:
                    tk = [eg.Ticket(x, k) for _ in range(3)]
                    for q in tk:
                        _ulp(oracle, q.wait(), wants[t])
                else:
                    assert np.array_equal(eg.truncate_norm_host(x, k, eg.TRUNCATE), x[:, :min(k, D)])
                i += 1
            calls[t] = i
        except Exception as e:  # noqa: BLE001
            errs.append((t, shapes[t], repr(e)[:300]))

    th = [threading.Thread(target=handler, args=(t,)) for t in range(16)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    assert min(calls) >= 2, calls


def test_init_shutdown_cycles_give_the_memory_back(oracle):
    """etn_init takes 4 x (64 + 64) MiB per device; etn_shutdown has to return all of it, every
    time, whether or not calls were made in between."""
    x = oracle.fill(4, 2000, 1024, np.float32)
    want = oracle.renorm_exact(x, 256)
    eg.shutdown()
    torch.cuda.set_device(0)
    torch.cuda.synchronize()
    _ulp(oracle, (eg.init([0]), eg.truncate_norm_host(x, 256))[1], want)       # first cycle loads the kernels
    eg.shutdown()
    free0, _ = torch.cuda.mem_get_info()
    for cycle in range(4):
        assert eg.init([0]) == 1
        during, _ = torch.cuda.mem_get_info()
        assert free0 - during >= 4 * 128 * (1 << 20)
        if cycle % 2 == 0:
            _ulp(oracle, eg.truncate_norm_host(x, 256), want)
        eg.shutdown()
        after, _ = torch.cuda.mem_get_info()
        assert after >= free0 - (8 << 20), (cycle, free0, after)
    assert eg.init() >= 1
