"""-m gpu tests of the JSON packer's device path (llm-mcp_b200/csrc/etn_json_gpu.cu,
etn_numfmt.cuh): the numbers of the kept prefixes are parsed, renormalised and formatted on the
GPU.  Bar: byte-identical with the host path (std::from_chars / std::to_chars, the guarantees
of strconv.ParseFloat / AppendFloat that encoding/json uses at handlers.go:2048 and
helpers.go:14) and with the oracle's Go spelling."""
import ctypes
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

from llm_mcp_b200 import embedgpu as eg  # noqa: E402
from tests.gpu_util import to_dev  # noqa: E402
from tests.test_json_packer import GO_KNOWN  # noqa: E402


@pytest.fixture(scope="module", autouse=True)
def _init():
    assert eg.init() >= 1
    yield
    eg.shutdown()


def _random_doubles(n, seed):
    rng = np.random.default_rng(seed)
    bits = rng.integers(0, 2 ** 64, size=n, dtype=np.uint64)
    v = bits.view(np.float64).copy()
    v = v[np.isfinite(v)]
    emb = rng.standard_normal(n) * 0.04
    ints = rng.integers(-10 ** 15, 10 ** 15, size=n // 4).astype(np.float64)
    edge = np.array([0.0, -0.0, 1e-6, 9.999999999999999e-7, 1e21, 999999999999999934463.0, 5e-324, 1.7976931348623157e308,
                     2.2250738585072014e-308, 0.1, 0.3, 1 / 3, 100.0, 1e22, 123456789.125] + [x for x, _ in GO_KNOWN])
    return np.concatenate([v, emb, ints, edge, -edge])


def _host_format(v):
    buf = ctypes.create_string_buffer(40)
    out = []
    for x in v:
        a = np.array([x])
        n = eg.lib().etn_json_format_array(a.ctypes.data, 1, buf, 40)
        assert n > 2
        out.append(buf.raw[1:n - 1])
    return out


def test_device_format_equals_host_format_and_go_spelling(oracle):
    v = _random_doubles(25000, 1)
    d_v = to_dev(v)
    d_out = torch.zeros(len(v) * 32, dtype=torch.uint8, device="cuda")
    d_len = torch.zeros(len(v), dtype=torch.int32, device="cuda")
    eg.check(eg.lib().etn_json_format_numbers_device(d_v.data_ptr(), len(v), d_out.data_ptr(), d_len.data_ptr(), None), "format")
    torch.cuda.synchronize()
    text, lens = d_out.cpu().numpy().reshape(-1, 32), d_len.cpu().numpy()
    got = [text[i, :lens[i]].tobytes() for i in range(len(v))]
    want = _host_format(v)
    bad = [(float(x), g, w) for x, g, w in zip(v, got, want) if g != w]
    assert not bad, bad[:5]
    for x, g in list(zip(v, got))[::97]:                       # and the oracle's restatement of encoding/json
        assert g.decode() == oracle.go_format_float(float(x)), x
    spelled = dict(zip((int(b) for b in v.view(np.uint64)), got))       # by bit pattern: -0.0 == 0.0 as floats
    for x, want_text in GO_KNOWN:                              # Go's documented spellings
        assert spelled[int(np.float64(x).view(np.uint64))].decode() == want_text, x
    assert lens.max() <= 25


def test_device_parse_equals_host_parse():
    v = _random_doubles(25000, 2)
    toks = _host_format(v)                                     # shortest spellings ...
    rng = np.random.default_rng(3)
    toks += [("%.17g" % x).encode() for x in v[::3]]           # ... 17 digits ...
    toks += [("%.*e" % (int(rng.integers(0, 19)), x)).encode() for x in v[::5]]
    toks += [b"1e999", b"-1e999", b"12345678901234567890", b"0.12345678901234567891", b"4.9e-324", b"1e-999", b"1E5", b"0e0", b"-0.0"]
    toks = [t for t in toks if b"n" not in t and b"i" not in t]
    blob = b",".join(toks)
    beg = np.cumsum([0] + [len(t) + 1 for t in toks[:-1]]).astype(np.int64)
    ln = np.array([len(t) for t in toks], dtype=np.int32)
    d_text = to_dev(np.frombuffer(blob, dtype=np.uint8).copy())
    d_beg, d_len = to_dev(beg), to_dev(ln)
    d_out = torch.zeros(len(toks), dtype=torch.float64, device="cuda")
    d_ok = torch.zeros(len(toks), dtype=torch.uint8, device="cuda")
    eg.check(eg.lib().etn_json_parse_tokens_device(d_text.data_ptr(), d_beg.data_ptr(), d_len.data_ptr(), len(toks), d_out.data_ptr(),
                                                   d_ok.data_ptr(), None), "parse")
    torch.cuda.synchronize()
    got, ok = d_out.cpu().numpy(), d_ok.cpu().numpy().astype(bool)
    want = np.zeros(len(toks))
    host_ok = np.zeros(len(toks), dtype=bool)
    one = np.zeros(1)
    for i, t in enumerate(toks):
        arr = b"[" + t + b"]"
        host_ok[i] = eg.lib().etn_json_parse_array(arr, len(arr), one.ctypes.data, 1) == 1
        want[i] = one[0]
    assert not (ok & ~host_ok).any()                           # never answers what the host refuses (overflow)
    both = ok & host_ok
    assert np.array_equal(got[both].view(np.uint64), want[both].view(np.uint64))
    assert both.mean() > 0.97                                  # the fall-back is the rare case


def _body(x, sep=(", ", ": "), indent=None):
    return json.dumps({"object": "list", "data": [{"object": "embedding", "index": i, "embedding": r} for i, r in enumerate(x)],
                       "usage": {"prompt_tokens": 4}}, separators=sep, indent=indent).encode()


def test_response_rewrite_on_the_gpu_is_byte_identical_to_the_host_path(oracle):
    rng = np.random.default_rng(9)
    x = [list(map(float, r)) for r in oracle.fill(1, 48, 2048, np.float64)]
    x += [list(map(float, rng.standard_normal(2048) * 10.0 ** rng.integers(-8, 8))) for _ in range(8)]
    x[3][5] = 0.0; x[3][6] = -0.0; x[4] = [0.0] * 2048                        # zeros; an all-zero prefix is copied
    x[5][:4] = [1e-300, -2.5e300, 123456789012345680000.0, 5e-324]             # scaling path of the renorm kernel, subnormal in
    x[6] = x[6][:700]                                                          # shorter than dims: left whole
    for sep, indent in (((", ", ": "), None), ((",", ":"), None), ((" ,\n", " : "), 1)):
        body = _body(x, sep, indent)
        os.environ["ETN_JSON_GPU"] = "0"
        host, hs = eg.truncate_response_json(body, 1024, renorm=True)
        os.environ["ETN_JSON_GPU"] = "1"
        before = eg.counters()["llmcore_embed_gpu_kernel_launches_total"]
        dev, ds = eg.truncate_response_json(body, 1024, renorm=True)
        launches = eg.counters()["llmcore_embed_gpu_kernel_launches_total"] - before
        os.environ.pop("ETN_JSON_GPU")
        assert dev == host and ds == hs
        assert launches == 3                                   # parse, renormalise, format: one launch each
        want = oracle.response_ref(body, 1024, renorm=True)
        got = json.loads(dev)
        for a, b in zip(got["data"], want["data"]):
            d, i = oracle.max_ulp(np.array(a["embedding"], dtype=np.float64), np.array(b["embedding"], dtype=np.float64))
            assert 0 <= d <= 2, (d, i)


def test_numbers_the_fast_parser_declines_are_redone_by_the_host(oracle):
    x = [list(map(float, r)) for r in oracle.fill(2, 6, 300, np.float64)]
    body = _body(x, (",", ":"))
    # 25 significant digits in vector 1, a half-way case in vector 4: Eisel-Lemire says "ask the host"
    body = body.replace(b'"index":1,"embedding":[', b'"index":1,"embedding":[0.1234567890123456789012345,', 1)
    body = body.replace(b'"index":4,"embedding":[', b'"index":4,"embedding":[9007199254740993,', 1)
    os.environ["ETN_JSON_GPU"] = "0"
    host, _ = eg.truncate_response_json(body, 128, renorm=True)
    os.environ["ETN_JSON_GPU"] = "1"
    dev, _ = eg.truncate_response_json(body, 128, renorm=True)
    os.environ.pop("ETN_JSON_GPU")
    assert dev == host
    assert json.loads(dev) == json.loads(host)


@pytest.mark.parametrize("dims,n_src", [(1, 5), (3, 7), (255, 300), (4096, 4100), (7000, 7001)])
def test_device_path_at_the_ends_of_its_range(oracle, dims, n_src):
    """One kept number per vector, thousands of them, and the largest prefix whose text still
    fits a block's shared memory (7,000 numbers; beyond that the host path takes over): the
    answer is the host path's, byte for byte."""
    rng = np.random.default_rng(dims)
    x = [list(map(float, rng.standard_normal(n_src) * 0.05)) for _ in range(5)]
    x[2] = x[2][:dims]                                           # exactly dims long: left whole
    body = _body(x, ("\n , ", ": "))
    os.environ["ETN_JSON_GPU"] = "0"
    host, hs = eg.truncate_response_json(body, dims, renorm=True)
    os.environ["ETN_JSON_GPU"] = "1"
    l0 = eg.counters()["llmcore_embed_gpu_kernel_launches_total"]
    dev, ds = eg.truncate_response_json(body, dims, renorm=True)
    launches = eg.counters()["llmcore_embed_gpu_kernel_launches_total"] - l0
    os.environ.pop("ETN_JSON_GPU")
    assert dev == host and ds == hs == {"gpu_vectors": 4, "resliced_vectors": 0, "arrays": 5}
    assert launches == 3, launches                               # parse, renormalise, format: it did run on the device
    got = json.loads(dev)["data"]
    assert [len(d["embedding"]) for d in got] == [dims] * 5
    for d, v in zip(got, x):
        want = oracle.renorm_exact(np.array([v]), dims)[0] if len(v) > dims else np.array(v)
        dd, i = oracle.max_ulp(np.array(d["embedding"], dtype=np.float64), want)
        assert 0 <= dd <= 2, (dd, i)


def test_vectors_that_only_get_the_cut_are_counted_as_fallbacks(oracle):
    x = [list(map(float, r)) for r in oracle.fill(3, 4, 64, np.float64)]
    x[1][3] = "not a number"                                     # cannot be renormalised: the reference reslice, on the text
    body = _body(x)
    f0 = eg.counters()["llmcore_embed_gpu_fallback_total"]
    out, st = eg.truncate_response_json(body, 16, renorm=True)
    assert st == {"gpu_vectors": 3, "resliced_vectors": 1, "arrays": 4}
    assert json.loads(out)["data"][1]["embedding"] == x[1][:16]
    assert eg.counters()["llmcore_embed_gpu_fallback_total"] == f0 + 1


def test_handler_with_renormalisation_on(oracle):
    """POST /v1/embeddings, cloud branch, EMBED_RENORMALIZE=1: the provider ignores `dimensions`,
    the handler mirror cuts and renormalises through the packer's device path."""
    from llm_mcp_b200 import handler as h
    x = oracle.fill(6, 5, 512, np.float64)
    body = json.dumps({"object": "list", "data": [{"index": i, "embedding": r.tolist()} for i, r in enumerate(x)],
                       "usage": {"prompt_tokens": 12}}).encode()
    env = {"OPENROUTER_API_KEY": "k", "EMBED_RENORMALIZE": "1", "CLOUD_EMBED_DIMENSIONS": "128"}
    m = h.Metrics()
    status, _, out = h.handle_embeddings("POST", b'{"model":"qwen/qwen3-embedding-8b","input":["a","b","c","d","e"]}', env, lambda u, hd, b: (200, body), m)
    assert status == 200 and m.input_tokens == {("qwen/qwen3-embedding-8b", "openrouter"): 12.0}
    got = json.loads(out)
    assert [d["index"] for d in got["data"]] == list(range(5)) and got["usage"] == {"prompt_tokens": 12}
    want = oracle.renorm_exact(x, 128)
    for d, w in zip(got["data"], want):
        dd, i = oracle.max_ulp(np.array(d["embedding"], dtype=np.float64), w)
        assert 0 <= dd <= 2, (dd, i)


def test_fuzz_device_path_equals_host_path():
    """Random responses — vector counts and lengths, white space, number spellings (integers,
    exponents, 20-digit mantissas, negative zero, tiny and huge magnitudes), vectors that cannot
    be renormalised — through the device path and the host path: same bytes, same statistics."""
    rng = np.random.default_rng(424242)

    def spell(v):
        c = rng.integers(0, 8)
        if c == 0: return repr(float(v))
        if c == 1: return "%.17g" % v
        if c == 2: return "%.*e" % (int(rng.integers(0, 18)), v)
        if c == 3: return str(int(v * 1000))
        if c == 4: return "%.20f" % v
        if c == 5: return ("%.6E" % v).replace("E", "e" if rng.integers(0, 2) else "E")
        if c == 6: return "-0.0" if rng.integers(0, 2) else "0"
        return repr(float(v) * 10.0 ** int(rng.integers(-200, 200)))

    for trial in range(60):
        dims = int(rng.choice([1, 2, 7, 64, 300]))
        seps = [",", ", ", " ,", "\n,\t", " , "]
        items = []
        for i in range(int(rng.integers(1, 9))):
            n = int(rng.integers(max(1, dims - 2), dims + 40))
            toks = [spell(x) for x in rng.standard_normal(n) * 0.05]
            if rng.integers(0, 6) == 0 and n > 1:
                toks[int(rng.integers(0, min(n, dims)))] = '"x"'          # not renormalisable: cut as text
            sep = str(rng.choice(seps))
            items.append('{"index": %d, "embedding": [%s%s]}' % (i, " " * int(rng.integers(0, 3)), sep.join(toks)))
        body = ('{"object": "list", "data": [' + ", ".join(items) + '], "usage": {"prompt_tokens": %d}}' % trial).encode()
        os.environ["ETN_JSON_GPU"] = "0"
        try:
            host = eg.truncate_response_json(body, dims, renorm=True)
        except eg.EtnError as e:
            host = ("error", e.code)
        os.environ["ETN_JSON_GPU"] = "1"
        try:
            dev = eg.truncate_response_json(body, dims, renorm=True)
        except eg.EtnError as e:
            dev = ("error", e.code)
        os.environ.pop("ETN_JSON_GPU")
        assert dev == host, (trial, dims, body[:300])
