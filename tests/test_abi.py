"""CPU tests of the drop-in boundary: the shared library loads, exports every symbol that
include/embed_truncate_norm.h declares, and rejects bad arguments before touching CUDA.
No compute call is made here (there is no GPU in the CPU container)."""
import ctypes
import os
import re

import pytest

from llm_mcp_b200 import embedgpu as eg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "embed_truncate_norm.h")


def declared_symbols():
    src = open(HEADER).read()
    return re.findall(r"ETN_API\s+[\w\s\*]+?\b(\w+)\s*\(", src)


def test_library_exports_every_declared_symbol():
    names = declared_symbols()
    assert len(names) >= 12 and "embed_truncate_norm" in names
    L = eg.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in the header but not exported"
    assert sorted(names) == sorted(eg.EXPORTS)


def test_header_cites_the_reference_per_entry_point():
    src = open(HEADER).read()
    assert "handlers.go:2062-2078" in src and "handlers.go:1891-1893" in src
    for block in re.split(r"\n(?=/\*)", src):
        if "ETN_API" in block and "etn_free_pinned" not in block.split("ETN_API")[1][:40]:
            assert re.search(r"(handlers|main|helpers|metrics|router|server)\.go:\d+|[Nn]o reference[\s\*]+counterpart|no[\s\*]+counterpart in the reference", block), block[:200]


def test_version_and_strerror():
    L = eg.lib()
    assert L.etn_version() == 200
    assert eg.strerror(0) == "ok"
    for code in (eg.ERR_NULL, eg.ERR_SHAPE, eg.ERR_ENUM, eg.ERR_ALIGN, eg.ERR_NOT_INIT, eg.ERR_DEVICE):
        assert eg.strerror(code) not in ("ok", "unknown error")
    assert eg.strerror(-99) == "unknown error"
    assert "memory" in eg.strerror(2).lower()      # cudaErrorMemoryAllocation


def test_argument_errors_need_no_device():
    L = eg.lib()
    buf = (ctypes.c_double * 16)()
    p = ctypes.addressof(buf)
    assert L.embed_truncate_norm(0, 8, 4, eg.F64, None, None, None) == 0
    assert L.embed_truncate_norm(2, 8, 4, eg.F64, None, p, None) == eg.ERR_NULL
    assert L.embed_truncate_norm(-2, 8, 4, eg.F64, p, p, None) == eg.ERR_SHAPE
    assert L.embed_truncate_norm(2, -8, 4, eg.F64, p, p, None) == eg.ERR_SHAPE
    assert L.embed_truncate(2, 8, 0, eg.F64, p, p, None) == eg.ERR_SHAPE
    assert L.embed_truncate(2, 8, 4, 2, p, p, None) == eg.ERR_ENUM
    assert L.embed_truncate_norm(2, 8, 4, eg.F64, p + 4, p, None) == eg.ERR_ALIGN
    assert L.embed_truncate_norm_ragged(2, 8, None, None, eg.F64, 1, p, p, None) == eg.ERR_NULL
    assert L.embed_truncate_norm_ragged(2, 8, p, p, eg.F64, 5, p, p, None) == eg.ERR_ENUM
    assert L.embed_truncate_norm_host(2, 8, 4, eg.F64, 3, p, 0, p, 0) == eg.ERR_ENUM
    assert L.embed_truncate_norm_host(2, 8, 4, eg.F64, 1, p, 8, p, 0) == eg.ERR_SHAPE   # pitch < D*8
    assert L.embed_truncate_norm_host(2, 8, 4, eg.F64, 1, p, 0, p, -1) == eg.ERR_DEVICE
    assert L.embed_truncate_norm_host(0, 8, 4, eg.F64, 1, None, 0, None, 0) == 0
    assert L.embed_truncate_norm_host(2, 8, 4, eg.F64, 1, p, 0, p, 0) == eg.ERR_NOT_INIT
    assert L.etn_fill_synthetic(1, 0, 4, 8, 9, 9, eg.F32, p, None) == eg.ERR_SHAPE
    assert L.etn_device_count() == 0
    L.etn_shutdown()                                 # harmless before init


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(eg, "_lib", None)
    monkeypatch.setattr(eg, "LIB_PATH", "/nonexistent/libembed_truncate_norm.so")
    with pytest.raises(FileNotFoundError, match="no CPU fallback"):
        eg.lib()


def test_host_mirror_config_surface():
    assert eg.resolve_dims(256, {"CLOUD_EMBED_DIMENSIONS": "512"}) == 256
    assert eg.resolve_dims(0, {"CLOUD_EMBED_DIMENSIONS": "512"}) == 512
    assert eg.resolve_dims(0, {"CLOUD_EMBED_DIMENSIONS": "5x"}) == 0 and eg.resolve_dims(0, {}) == 0
    assert eg.resolve_dims(-1, {"CLOUD_EMBED_DIMENSIONS": "512"}) == -1
    assert eg.renorm_enabled({}) is False and eg.renorm_enabled({"EMBED_RENORMALIZE": "1"}) is True
    # no device needed when nothing has to be cut: same no-ops as handlers.go:2064-2074
    r = {"data": [{"embedding": [1.0, 2.0]}, "x", {"embedding": "b64"}]}
    assert eg.truncate_embeddings(r, 2)["gpu_vectors"] == 0 and r["data"][0]["embedding"] == [1.0, 2.0]
    assert eg.truncate_embeddings({"data": 3}, 2)["gpu_vectors"] == 0
    assert eg.truncate_embeddings(r, 0)["gpu_vectors"] == 0 and eg.truncate_embeddings(r, -5)["gpu_vectors"] == 0


def test_counters_and_typed_mirror_need_no_device():
    c0 = eg.counters()
    assert list(c0) == [eg.lib().etn_counter_name(i).decode() for i in range(10)] and all(k.startswith("llmcore_embed_gpu_") for k in c0)
    assert "llmcore_embed_gpu_fallback_total" in c0 and "llmcore_embed_gpu_errors_total" in c0
    eg.lib().embed_truncate_norm(2, 8, 0, eg.F64, None, None, None)          # shape error: counted nowhere but...
    p = (ctypes.c_double * 4)()
    assert eg.lib().embed_truncate_norm_host(2, 8, 4, eg.F64, 1, ctypes.addressof(p), 0, ctypes.addressof(p), 0) == eg.ERR_NOT_INIT
    assert "# TYPE llmcore_embed_gpu_errors_total counter" in eg.metrics_text()
    # N2 mirror: nothing to cut -> no device needed, input returned as is
    assert eg.truncate_typed([[1.0, 2.0], [3.0]], 2) == [[1.0, 2.0], [3.0]]
    assert eg.truncate_typed([[1.0, 2.0, 3.0]], 0) == [[1.0, 2.0, 3.0]]


def test_new_host_entry_points_reject_bad_arguments_without_a_device():
    L = eg.lib()
    buf = (ctypes.c_double * 16)()
    kpr = (ctypes.c_int32 * 2)(4, 4)
    p, pk = ctypes.addressof(buf), ctypes.addressof(kpr)
    assert L.embed_truncate_norm_ragged_host(0, 8, None, eg.F64, 1, None, 0, None, 0) == 0
    assert L.embed_truncate_norm_ragged_host(2, 8, None, eg.F64, 1, p, 0, p, 0) == eg.ERR_NULL
    assert L.embed_truncate_norm_ragged_host(2, 8, pk, 7, 1, p, 0, p, 0) == eg.ERR_ENUM
    assert L.embed_truncate_norm_ragged_host(2, 0, pk, eg.F64, 1, p, 0, p, 0) == eg.ERR_SHAPE
    assert L.embed_truncate_norm_ragged_host(2, 8, pk, eg.F64, 1, p, 4, p, 0) == eg.ERR_SHAPE
    assert L.embed_truncate_norm_ragged_host(2, 8, pk, eg.F64, 1, p, 0, p, 0) == eg.ERR_NOT_INIT
    t = ctypes.c_int64(0)
    assert L.etn_submit_host(2, 8, 4, None, eg.F64, 1, p, 0, p, 0, None) == eg.ERR_NULL
    assert L.etn_submit_host(2, 8, 4, None, eg.F64, 1, p, 0, p, 0, ctypes.byref(t)) == eg.ERR_NOT_INIT and t.value == 0
    assert L.etn_submit_host(0, 8, 4, None, eg.F64, 1, None, 0, None, 0, ctypes.byref(t)) == 0 and t.value > 0
    assert L.etn_wait(t.value) == 0                      # empty batch: nothing to wait for
    assert L.etn_wait(t.value) == eg.ERR_TICKET and L.etn_wait(123456789) == eg.ERR_TICKET
    assert "ticket" in eg.strerror(eg.ERR_TICKET)
    assert L.embed_truncate_norm_gather(2, None, pk, p, eg.F64, 1, p, p, None) == eg.ERR_NULL
    assert L.embed_truncate_norm_gather(0, None, None, None, eg.F64, 1, None, None, None) == 0
    assert L.embed_truncate_norm(2, 8, 4, eg.F64, p, p, None) == eg.ERR_SHAPE      # in place needs k >= D
    assert L.etn_shard_ragged(2, 8, pk, 0, p) == eg.ERR_SHAPE and L.etn_shard_ragged(2, 8, None, 2, p) == eg.ERR_NULL


def test_histograms_and_prometheus_exposition():
    """N3: fixed-bucket histograms beside metrics.go:18-26, exported as cumulative `le` series."""
    import math
    names = []
    for w in range(eg.HIST_COUNT):
        h = eg.histogram(w)
        names.append(h["name"])
        assert h["name"].startswith("llmcore_embed_gpu_") and h["help"]
        assert h["le"][-1] == math.inf and h["le"] == sorted(h["le"]) and len(h["le"]) == len(h["counts"])
    assert names == ["llmcore_embed_gpu_batch_rows", "llmcore_embed_gpu_kernel_seconds", "llmcore_embed_gpu_h2d_seconds",
                     "llmcore_embed_gpu_d2h_seconds", "llmcore_embed_gpu_call_seconds", "llmcore_embed_gpu_requests_per_launch"]
    assert eg.lib().etn_get_histogram(99, None, None, 0, None) == eg.ERR_ENUM
    le = (ctypes.c_double * 2)()
    cnt = (ctypes.c_int64 * 2)()
    assert eg.lib().etn_get_histogram(0, le, cnt, 2, None) == eg.ERR_SHAPE
    text = eg.metrics_text()
    for n in names:
        assert f"# TYPE {n} histogram" in text and f'{n}_bucket{{le="+Inf"}}' in text
        assert f"{n}_sum " in text and f"{n}_count " in text
    # cumulative buckets never decrease and the +Inf bucket equals _count
    import re
    for n in names:
        vals = [int(v) for v in re.findall(rf"^{n}_bucket{{[^}}]*}} (\d+)$", text, re.M)]
        assert vals == sorted(vals) and vals[-1] == int(re.search(rf"^{n}_count (\d+)$", text, re.M).group(1))
    before = eg.counters()["llmcore_embed_gpu_fallback_total"]
    eg.lib().etn_note_fallback(3)
    assert eg.counters()["llmcore_embed_gpu_fallback_total"] == before + 3
