"""-m gpu tests of the host-buffer path (embed_truncate_norm_host) and of the host-side
mirror of the reference call site, truncate_embeddings(result, dims)
(reference: core/internal/api/handlers.go:1891-1893, 2062-2078)."""
import copy
import json

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

from llm_mcp_b200 import embedgpu as eg  # noqa: E402


@pytest.fixture(scope="module", autouse=True)
def _init():
    assert eg.init() >= 1
    yield
    eg.shutdown()


def _ulp(oracle, got, want, bar=2):
    d, i = oracle.max_ulp(got, want)
    assert 0 <= d <= bar, (d, i)


@pytest.mark.parametrize("rows,D,k,dt", [
    (1024, 4096, 1024, np.float64),        # C1 through the handler-facing call
    (70000, 768, 256, np.float32),         # several chunks, ragged last chunk
    (5000, 770, 257, np.float32), (3, 40, 39, np.float64), (17, 48, 100, np.float32),
])
def test_host_path_against_oracle(oracle, rows, D, k, dt):
    xh = oracle.fill(3, rows, D, dt)
    for pinned in (False, True):
        if pinned:
            pin = eg.PinnedBuffer(xh.nbytes)
            x = pin.array(xh.shape, dt)
            x[:] = xh
        else:
            x = xh
        got = eg.truncate_norm_host(x, k, eg.TRUNCATE_RENORM)
        _ulp(oracle, got, oracle.renorm_exact(xh, k))
        got = eg.truncate_norm_host(x, k, eg.TRUNCATE)
        assert np.array_equal(got, oracle.truncate(xh, k))
        if pinned:
            del x
            pin.free()


def test_host_path_row_pitch(oracle):
    wide = oracle.fill(4, 4000, 1100, np.float32)
    x = wide[:, :1024]                        # pitch 1100*4 > D*4
    got = eg.truncate_norm_host(x, 256)
    _ulp(oracle, got, oracle.renorm_exact(np.ascontiguousarray(x), 256))


def test_host_path_large_multi_chunk(oracle):
    rows, D, k = 262144, 1024, 256            # 256 MiB prefix -> 8 staging chunks
    pin_in, pin_out = eg.PinnedBuffer(rows * D * 4), eg.PinnedBuffer(rows * k * 4)
    x = pin_in.array((rows, D), np.float32)
    x[:] = oracle.fill(6, rows, D, np.float32)
    y = eg.truncate_norm_host(x, k, out=pin_out.array((rows, k), np.float32))
    idx = np.r_[0:64, rows // 2:rows // 2 + 64, rows - 64:rows]
    _ulp(oracle, y[idx], oracle.renorm_exact(x[idx], k))
    n = np.sqrt((y.astype(np.float64) ** 2).sum(1))
    assert np.abs(n - 1).max() < 2e-7
    del x, y
    pin_in.free(); pin_out.free()


# ------------------------------------------------ reference call-site mirror ----
def _response(oracle, n=37, D=1024, seed=8):
    v = oracle.fill(seed, n, D, np.float64)
    return {"object": "list", "model": "qwen/qwen3-embedding-8b",
            "data": [{"object": "embedding", "index": i, "embedding": v[i].tolist()} for i in range(n)],
            "usage": {"prompt_tokens": 11, "total_tokens": 11}}


def test_mirror_truncate_equals_reference(oracle):
    """Default mode: byte-for-byte what truncateEmbeddings produces, incl. the JSON text."""
    r = _response(oracle)
    r["data"][3]["embedding"] = r["data"][3]["embedding"][:100]       # shorter than dims: left whole
    r["data"][5]["embedding"] = r["data"][5]["embedding"][:300]       # ragged lengths in one response
    r["data"][7]["embedding"] = "AAAAAA=="                            # base64: untouched
    r["data"][9]["embedding"] = None
    r["data"].append("not-a-map")
    r["data"].append({"index": 99})
    r["data"][11]["embedding"] = r["data"][11]["embedding"][:299] + ["x", 1.0, 2.0]   # unpackable: resliced
    want = copy.deepcopy(r)
    oracle.truncate_embeddings_ref(want, 256)
    stats = eg.truncate_embeddings(r, 256, renorm=False)
    assert r == want
    assert json.dumps(r, sort_keys=True) == json.dumps(want, sort_keys=True)
    # truncation alone is the reference's O(1) reslice, as in the Go binding: no GPU round trip
    assert stats["gpu_vectors"] == 0 and stats["resliced_vectors"] == 37 - 3 and stats["launch_groups"] == 0
    # ... and the GPU's own TRUNCATE mode gives the same bits when asked for directly
    x = oracle.fill(8, 37, 1024, np.float64)
    assert np.array_equal(eg.truncate_norm_host(x, 256, eg.TRUNCATE), x[:, :256])


def test_mirror_renorm(oracle):
    r = _response(oracle, n=20, D=768)
    ref = copy.deepcopy(r)
    oracle.truncate_embeddings_ref(ref, 256)
    eg.truncate_embeddings(r, 256, renorm=True)
    for a, b in zip(r["data"], ref["data"]):
        assert a["index"] == b["index"] and len(a["embedding"]) == 256
        want = oracle.renorm_exact(np.array([b["embedding"]]), 256)[0]
        _ulp(oracle, np.array(a["embedding"]), want)
    assert r["usage"] == ref["usage"] and r["model"] == ref["model"]


def test_mirror_dims_not_positive_is_noop(oracle):
    r = _response(oracle, n=3, D=64)
    keep = copy.deepcopy(r)
    for dims in (0, -3):                      # handlers.go:1891: the op runs only when dims > 0
        eg.truncate_embeddings(r, dims)
        assert r == keep


def test_host_path_row_sharded_over_devices(oracle):
    """n_devices > 1: contiguous row blocks, one host thread and its own streams per device,
    no exchange between devices."""
    n = eg.lib().etn_device_count()
    rows, D, k = 100003, 1024, 256
    xh = oracle.fill(9, rows, D, np.float32)
    want = oracle.renorm_exact(xh, k)
    for nd in sorted({1, min(2, n), n}):              # on a one-GPU box: the one device, all four slots
        got = eg.truncate_norm_host(xh, k, eg.TRUNCATE_RENORM, n_devices=nd)
        _ulp(oracle, got, want)
        assert np.array_equal(eg.truncate_norm_host(xh, k, eg.TRUNCATE, n_devices=nd), xh[:, :k])
    assert eg.lib().embed_truncate_norm_host(4, 8, 4, eg.F32, 1, xh.ctypes.data, 0, xh.ctypes.data, n + 1) == eg.ERR_DEVICE


# ------------------------------------------------------------- JSON packer (N1) ----
def test_json_packer_renorm_against_reference(oracle):
    """Response text in, response text out: cut vectors are parsed prefix-only into pinned
    memory, renormalised on the GPU and formatted the way encoding/json writes float64."""
    import re
    x = oracle.fill(1, 48, 4096, np.float64)
    data = [{"object": "embedding", "index": i, "embedding": r.tolist()} for i, r in enumerate(x)]
    data[5]["embedding"] = data[5]["embedding"][:700]            # shorter than dims: left whole
    data[6]["embedding"] = "base64=="                             # not an array
    data[7]["embedding"] = data[7]["embedding"][:1023] + ["x"] + data[7]["embedding"][1024:]   # unpackable: cut only
    data[8]["embedding"] = [0.0] * 2000                           # zero prefix: copied, no NaN
    body = json.dumps({"object": "list", "data": data, "usage": {"prompt_tokens": 3}}).encode()
    out, stats = eg.truncate_response_json(body, 1024, renorm=True)
    got, want = json.loads(out), oracle.response_ref(body, 1024, renorm=True)
    assert stats == {"gpu_vectors": 45, "resliced_vectors": 1, "arrays": 47}
    assert got["usage"] == want["usage"] and len(got["data"]) == len(want["data"])
    for a, b in zip(got["data"], want["data"]):
        assert a.keys() == b.keys() and a["index"] == b["index"]
        if isinstance(b["embedding"], list) and all(type(v) is float for v in b["embedding"]):
            _ulp(oracle, np.array(a["embedding"], dtype=np.float64), np.array(b["embedding"], dtype=np.float64))
        else:
            assert a["embedding"] == b["embedding"]
    # every number the packer wrote is encoding/json's spelling of the value it parses back to
    m = re.search(rb'"index": 0, "embedding": \[([^\]]*)\]|"embedding": \[([^\]]*)\], "index": 0', out)
    toks = (m.group(1) or m.group(2)).decode().split(",")
    assert len(toks) == 1024 and all(t == oracle.go_format_float(float(t)) for t in toks)


# ------------------------------------------------------ typed vectors (N2), counters (N3) ----
def test_typed_vectors_ollama_branch(oracle):
    """handlers.go:1976-1984: Embeddings [][]float64; same cutting rules as truncateEmbeddings."""
    x = oracle.fill(12, 9, 768, np.float64)
    embs = [r.tolist() for r in x]
    embs[2] = embs[2][:100]
    embs[4] = embs[4][:300]
    before = eg.counters()
    got = eg.truncate_typed(embs, 256, renorm=False)
    assert [len(v) for v in got] == [256, 256, 100, 256, 256, 256, 256, 256, 256]
    assert all(g == e[:256] for g, e in zip(got, embs))
    assert eg.counters() == before                    # a slice needs no GPU
    got = eg.truncate_typed(embs, 256, renorm=True)
    for g, e in zip(got, embs):
        if len(e) > 256:
            _ulp(oracle, np.array(g), oracle.renorm_exact(np.array([e]), 256)[0])
        else:
            assert g == e
    after = eg.counters()
    d = {k: after[k] - before[k] for k in after}
    assert d["llmcore_embed_gpu_host_calls_total"] == 2 and d["llmcore_embed_gpu_host_rows_total"] == 8
    assert d["llmcore_embed_gpu_h2d_bytes_total"] == 8 * 256 * 8 == d["llmcore_embed_gpu_d2h_bytes_total"]
    # request-sized groups stay on one device: one launch each
    assert d["llmcore_embed_gpu_kernel_launches_total"] == 2 and d["llmcore_embed_gpu_errors_total"] == 0


def test_json_packer_base64_renorm(oracle):
    """N4: base64 float32 vectors reach the GPU without any number parsing."""
    import base64
    x = oracle.fill(3, 40, 4096, np.float32)
    items = [{"index": i, "embedding": base64.b64encode(r.astype("<f4").tobytes()).decode()} for i, r in enumerate(x)]
    items[3]["embedding"] = base64.b64encode(x[3, :100].tobytes()).decode()      # shorter than dims
    items.append({"index": 40, "embedding": oracle.fill(3, 1, 4096, np.float64)[0].tolist()})   # arrays still work
    body = json.dumps({"data": items}).encode()
    out, stats = eg.truncate_response_json(body, 1024, renorm=True, base64_vectors=True)
    got = json.loads(out)["data"]
    assert stats["gpu_vectors"] == 40 and stats["arrays"] == 41
    for g, it in zip(got[:40], items[:40]):
        want = oracle.b64_vector_ref(it["embedding"], 1024, True)
        a = np.frombuffer(base64.b64decode(g["embedding"]), dtype="<f4")
        b = np.frombuffer(base64.b64decode(want), dtype="<f4")
        _ulp(oracle, a.copy(), b.copy())
    _ulp(oracle, np.array(got[40]["embedding"]), oracle.renorm_exact(np.array([items[40]["embedding"]]), 1024)[0])


def test_host_path_concurrent_handlers(oracle):
    """Many handler threads at once (one goroutine per connection in the reference): calls on
    one device queue behind its lock and every caller gets its own rows back."""
    import threading
    inputs = [oracle.fill(40 + t, 3000 + 17 * t, 1024, np.float32) for t in range(6)]
    wants = [oracle.renorm_exact(x, 256) for x in inputs]
    errs = []

    def worker(t):
        try:
            for _ in range(5):
                got = eg.truncate_norm_host(inputs[t], 256)
                d, _ = oracle.max_ulp(got, wants[t])
                assert 0 <= d <= 2, d
        except Exception as e:  # noqa: BLE001
            errs.append((t, e))

    th = [threading.Thread(target=worker, args=(t,)) for t in range(6)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs


def test_job_results_async_path(oracle):
    """worker/llm_worker/main.py:246-261: one vector per finished ollama.embed job."""
    x = oracle.fill(14, 6, 768, np.float64)
    jobs = [{"status": 200, "ms": 12, "data": {"embedding": r.tolist()}} for r in x]
    jobs[1]["data"]["embedding"] = jobs[1]["data"]["embedding"][:64]
    jobs += [{"status": 500, "ms": 3, "data": {}}, {"status": 200, "ms": 1, "data": {"embedding": None}}, "junk"]
    keep = copy.deepcopy(jobs)
    st = eg.truncate_job_results(jobs, 256, renorm=True)
    assert st == {"gpu_vectors": 5, "resliced_vectors": 0, "launch_groups": 1}
    for j, k0 in zip(jobs[:6], keep[:6]):
        e0 = k0["data"]["embedding"]
        if len(e0) > 256:
            _ulp(oracle, np.array(j["data"]["embedding"]), oracle.renorm_exact(np.array([e0]), 256)[0])
        else:
            assert j == k0
    assert jobs[6:] == keep[6:]
    assert eg.truncate_job_results(copy.deepcopy(keep), 0)["gpu_vectors"] == 0
