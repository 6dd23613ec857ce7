"""-m gpu test of embed_truncate_norm_host_static, the host entry point for callers that
cannot allocate device memory: a llmcore that shares its GPU with a model server which has
already reserved the device's memory, or a short-lived probe that never runs etn_init.  It
stages through an array compiled into the library instead of cudaMalloc'ed buffers.

Collected last on purpose (pytest orders files by name): a `-x` run has then been through
every other test before this one can stop it.  It must also pass alone in a fresh process,
so it sets up everything it needs itself and calls the new entry point before anything
else has touched the library's staging.

Check: the output of the new entry point equals that of the existing one,
embed_truncate_norm_host, on two calls in a row (the second call finds the first call's
leftovers in the shared array), within the 2 ulp the other GPU tests allow; truncation must
be bit-identical."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

from llm_mcp_b200 import embedgpu as eg  # noqa: E402


def _static(x, k, mode, stream=None):
    batch, D = x.shape
    out = np.empty((batch, min(k, D)), dtype=x.dtype)
    rc = eg.lib().embed_truncate_norm_host_static(batch, D, k, eg.dtype_code(x.dtype), mode, x.ctypes.data,
                                                  x.strides[0], out.ctypes.data, stream)
    return rc, out


def test_static_entry_point_matches_allocating_one(oracle):
    torch.cuda.set_device(0)
    # two different calls in a row through the built-in array, before etn_init is ever needed
    x1 = oracle.fill(21, 1024, 4096, np.float64)               # C1: 8 MiB of prefixes
    x2 = oracle.fill(22, 3000, 772, np.float32)[:, :768]        # pitched rows, smaller footprint
    rc1, a1 = _static(x1, 1024, eg.TRUNCATE_RENORM)
    rc2, a2 = _static(x2, 256, eg.TRUNCATE_RENORM)
    rc3, t2 = _static(x2, 256, eg.TRUNCATE)
    assert (rc1, rc2, rc3) == (0, 0, 0), [eg.strerror(r) for r in (rc1, rc2, rc3)]

    was_init = eg.lib().etn_device_count() > 0
    eg.init([0])
    try:
        b1 = eg.truncate_norm_host(x1, 1024, eg.TRUNCATE_RENORM)
        b2 = eg.truncate_norm_host(x2, 256, eg.TRUNCATE_RENORM)
    finally:
        if not was_init:
            eg.shutdown()
    for got, want in ((a1, b1), (a2, b2)):
        d, i = oracle.max_ulp(got, want)
        assert 0 <= d <= 2, (d, i)
    assert np.array_equal(t2, x2[:, :256])
    # and against the oracle itself
    d, _ = oracle.max_ulp(a1, oracle.renorm_exact(x1, 1024))
    assert 0 <= d <= 2


def test_static_entry_point_limits_and_errors():
    torch.cuda.set_device(0)
    L = eg.lib()
    x = np.zeros((4, 8), dtype=np.float32)
    o = np.ones((4, 8), dtype=np.float32)
    big_rows = eg.STATIC_STAGE_BYTES // (1024 * 4) + 1          # one row too many
    xb = np.zeros((big_rows, 1024), dtype=np.float32)
    rc, _ = _static(xb, 1024, eg.TRUNCATE_RENORM)
    assert rc == eg.ERR_TOO_LARGE and "staging" in eg.strerror(rc)
    rc, out = _static(xb[:-1], 1024, eg.TRUNCATE_RENORM)        # exactly the limit fits
    assert rc == 0 and not out.any()
    p, q = x.ctypes.data, o.ctypes.data
    assert L.embed_truncate_norm_host_static(0, 8, 4, eg.F32, 1, None, 0, None, None) == 0
    assert L.embed_truncate_norm_host_static(4, 8, 4, eg.F32, 1, None, 0, q, None) == eg.ERR_NULL
    assert L.embed_truncate_norm_host_static(4, 8, 4, 9, 1, p, 0, q, None) == eg.ERR_ENUM
    assert L.embed_truncate_norm_host_static(4, 8, 0, eg.F32, 1, p, 0, q, None) == eg.ERR_SHAPE
    assert L.embed_truncate_norm_host_static(4, 8, 4, eg.F32, 1, p, 16, q, None) == eg.ERR_SHAPE
    assert (o == 1).all()
    # an explicit stream works too
    s = torch.cuda.Stream()
    xs = np.arange(32, dtype=np.float64).reshape(4, 8) + 1
    rc, out = _static(xs, 4, eg.TRUNCATE_RENORM, ctypes.c_void_p(s.cuda_stream))
    assert rc == 0
    assert np.allclose(out, xs[:, :4] / np.linalg.norm(xs[:, :4], axis=1, keepdims=True), rtol=1e-15)


def test_static_entry_point_concurrent_callers_and_no_allocation(oracle):
    """Callers on different threads and streams share the one built-in array: the library
    queues them behind a per-device lock, so each gets its own rows back.  And the point of
    the entry point — free device memory is the same before and after a call (the first call
    on the device, which loads the kernels, is made beforehand)."""
    import threading
    torch.cuda.set_device(0)
    xs = [oracle.fill(70 + t, 1500 + 100 * t, 1024, np.float32) for t in range(4)]
    wants = [oracle.renorm_exact(x, 256) for x in xs]
    rc, _ = _static(xs[0], 256, eg.TRUNCATE_RENORM)
    assert rc == 0
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    errs = []

    def caller(t):
        try:
            torch.cuda.set_device(0)
            s = torch.cuda.Stream()
            for _ in range(8):
                rc, got = _static(xs[t], 256, eg.TRUNCATE_RENORM, ctypes.c_void_p(s.cuda_stream))
                assert rc == 0, eg.strerror(rc)
                d, _ = oracle.max_ulp(got, wants[t])
                assert 0 <= d <= 2, d
        except Exception as e:  # noqa: BLE001
            errs.append((t, repr(e)))

    before = eg.counters()
    th = [threading.Thread(target=caller, args=(t,)) for t in range(4)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    after = eg.counters()
    assert after["llmcore_embed_gpu_host_calls_total"] - before["llmcore_embed_gpu_host_calls_total"] == 32
    assert after["llmcore_embed_gpu_host_rows_total"] - before["llmcore_embed_gpu_host_rows_total"] == 8 * sum(len(x) for x in xs)
    # one more plain call with nothing else running: free memory unchanged
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    rc, _ = _static(xs[1], 256, eg.TRUNCATE_RENORM)
    free2, _ = torch.cuda.mem_get_info()
    assert rc == 0 and free2 == free1, (free0, free1, free2)
    # a refused size is counted as "take the other path", not as an error
    f0, e0 = eg.counters()["llmcore_embed_gpu_fallback_total"], eg.counters()["llmcore_embed_gpu_errors_total"]
    big = np.zeros((eg.STATIC_STAGE_BYTES // 4096 + 1, 1024), dtype=np.float32)
    assert _static(big, 1024, eg.TRUNCATE_RENORM)[0] == eg.ERR_TOO_LARGE
    assert eg.counters()["llmcore_embed_gpu_fallback_total"] == f0 + 1 and eg.counters()["llmcore_embed_gpu_errors_total"] == e0
