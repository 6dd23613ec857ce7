"""CPU tests of the JSON packer: number parsing/formatting against
encoding/json's rules, and the response rewrite in TRUNCATE mode (no GPU involved: the kept
tokens are spliced as text) against the map-level restatement of the reference
(handlers.go:2047-2050, 2062-2078; helpers.go:11-15)."""
import ctypes
import json
import math
import struct

import numpy as np
import pytest

from llm_mcp_b200 import embedgpu as eg

# encoding/json outputs that are part of Go's documented/observable behaviour
GO_KNOWN = [
    (0.1, "0.1"), (100.0, "100"), (1.5, "1.5"), (-0.0, "-0"), (0.0, "0"), (123456789.0, "123456789"),
    (1e-6, "0.000001"), (1e-7, "1e-7"), (1.234e-7, "1.234e-7"), (1e-10, "1e-10"), (0.000001234, "0.000001234"),
    (1e20, "100000000000000000000"), (1e21, "1e+21"), (1.5e300, "1.5e+300"), (5e-324, "5e-324"),
    (1.7976931348623157e308, "1.7976931348623157e+308"), (-2.5e-9, "-2.5e-9"), (0.30000000000000004, "0.30000000000000004"),
    (123456789012345680000.0, "123456789012345680000"), (3.0e-5, "0.00003"), (-1234.5678, "-1234.5678"),
]


def fmt(vals):
    v = np.ascontiguousarray(vals, dtype=np.float64)
    buf = ctypes.create_string_buffer(34 * len(v) + 2)
    n = eg.lib().etn_json_format_array(v.ctypes.data, len(v), buf, len(buf))
    assert n > 0, n
    return buf.raw[:n].decode()


def parse(text, max_n):
    dst = np.full(max_n, np.nan)
    n = eg.lib().etn_json_parse_array(text.encode(), len(text.encode()), dst.ctypes.data, max_n)
    return n, dst


def test_format_matches_go_known_answers(oracle):
    for x, want in GO_KNOWN:
        assert oracle.go_format_float(x) == want, (x, oracle.go_format_float(x))
        assert fmt([x]) == f"[{want}]", (x, fmt([x]))
    assert fmt([]) == "[]" and fmt([1.0, -2.0, 0.5]) == "[1,-2,0.5]"
    bad = np.array([1.0, np.inf])
    buf = ctypes.create_string_buffer(100)
    assert eg.lib().etn_json_format_array(bad.ctypes.data, 2, buf, 100) == eg.ERR_JSON
    assert eg.lib().etn_json_format_array(bad.ctypes.data, 1, buf, 3) == eg.ERR_SHAPE


def test_format_random_doubles_round_trip_and_match_restatement(oracle):
    rng = np.random.default_rng(3)
    bits = rng.integers(0, 2**63 - 1, size=20000, dtype=np.int64) | (rng.integers(0, 2, size=20000, dtype=np.int64) << 63)
    x = bits.view(np.float64)
    x = x[np.isfinite(x)]
    x = np.concatenate([x, rng.standard_normal(5000) * 0.03, np.float64(10.0) ** rng.integers(-30, 30, 500),
                        np.array([9.999999e-7, 1.0000001e-6, 9.99999999999999e20, 1.0000000000000001e21, 2.2250738585072014e-308])])
    toks = fmt(x)[1:-1].split(",")
    assert len(toks) == len(x)
    for v, t in zip(x[::7], toks[::7]):
        assert t == oracle.go_format_float(float(v)), (v, t)
    back = np.array([float(t) for t in toks])
    assert np.array_equal(back.view(np.uint64), x.view(np.uint64))          # shortest digits round-trip exactly
    n, got = parse("[" + ",".join(toks) + "]", len(x))
    assert n == len(x) and np.array_equal(got.view(np.uint64), x.view(np.uint64))


def test_parse_is_correctly_rounded_and_strict():
    texts = ["0.1", "1e-7", "-0", "123456789012345678901234567890", "2.2250738585072011e-308", "4.9e-324", "1E5", "0.30000000000000004",
             "179769313486231570000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000"]
    n, got = parse("[" + ", ".join(texts) + " ]", len(texts))
    assert n == len(texts)
    assert np.array_equal(got.view(np.uint64), np.array([float(t) for t in texts]).view(np.uint64))
    assert parse("[1,2,3,4]", 2)[0] == 4 and parse("[]", 4)[0] == 0
    for bad in ("[1,]", "[01]", "[1.]", "[.5]", "[1e]", "[+1]", "[NaN]", "[1 2]", "1,2", "[1e999]", '["x"]'):
        assert parse(bad, 4)[0] == eg.ERR_JSON, bad


def _body(vectors, extra_items=(), **top):
    data = [{"object": "embedding", "index": i, "embedding": v} for i, v in enumerate(vectors)] + list(extra_items)
    d = {"object": "list", "data": data, "model": "qwen/qwen3-embedding-8b", "usage": {"prompt_tokens": 5, "total_tokens": 5}}
    d.update(top)
    return json.dumps(d).encode()


def test_truncate_mode_equals_reference_value_and_keeps_tokens(oracle):
    rng = np.random.default_rng(8)
    vecs = [list(map(float, rng.standard_normal(n) * 0.05)) for n in (40, 40, 7, 16, 17)]
    body = _body(vecs, extra_items=["not-a-map", {"embedding": "AAAA"}, {"embedding": None}, {"index": 9},
                                    {"embedding": [1, "x", 3.5, None, {"a": [1, 2]}, 6]}], note="<&> é \\\" ]} [")
    out, stats = eg.truncate_response_json(body, 16, renorm=False)
    assert json.loads(out) == oracle.response_ref(body, 16, renorm=False)
    assert stats == {"gpu_vectors": 0, "resliced_vectors": 3, "arrays": 6}
    # the kept numbers are the provider's own tokens, byte for byte
    for v in vecs[:2]:
        kept = json.dumps(v[:16])[1:-1].encode()
        assert kept in out
    # short vectors and everything else untouched: nothing but the cut tails went missing
    assert out.count(b'"embedding"') == body.count(b'"embedding"')
    for dims in (0, -3):
        assert eg.truncate_response_json(body, dims, renorm=False)[0] == body
    assert eg.truncate_response_json(body, 1000, renorm=False)[0] == body


def test_rewrite_handles_whitespace_escapes_and_odd_shapes(oracle):
    body = b' {\n "data" : [ { "index":0 , "embedding" : [ 1.0 , 2.5e-3,\t-7 , 4 ] , "k\\"ey":"]}" } , [1,2,3], {"embedding":[ ]} ],\n "data2":{"embedding":[1,2,3]} }\n'
    out, _ = eg.truncate_response_json(body, 2, renorm=False)
    assert json.loads(out) == oracle.response_ref(body, 2, renorm=False)
    assert b"[ 1.0 , 2.5e-3]" in out                       # tokens and their spacing kept
    for b2 in (b'{"data": 3}', b'{"data": {"embedding": [1,2,3]}}', b'{}', b'[1,2,3]', b'{"x": {"data": [{"embedding": [1,2,3]}]}}'):
        assert eg.truncate_response_json(b2, 1, renorm=False)[0] == b2
    for bad, code in ((b'{"data": [', eg.ERR_JSON), (b'{"data": [{"embedding": [1,2}]}', eg.ERR_JSON), (b'{"a":1} x', eg.ERR_JSON),
                      (b'{"data": [], "data": []}', eg.ERR_JSON_DUP), (b'{"data": [{"embedding": [1], "embedding": [2]}]}', eg.ERR_JSON_DUP)):
        with pytest.raises(eg.EtnError) as ei:
            eg.truncate_response_json(bad, 1, renorm=False)
        assert ei.value.code == code, bad
    small = ctypes.create_string_buffer(4)
    assert eg.lib().etn_json_truncate_response(b'{"a": 1}', 8, 1, 0, small, 4, 0, None) == eg.ERR_SHAPE


def test_c1_sized_response_truncate(oracle):
    """The reference's CPU-runnable case: 1,024 vectors of 4096 numbers cut to 1024."""
    x = oracle.fill(1, 64, 4096, np.float64)               # 64 vectors keep the CPU suite quick
    body = _body([r.tolist() for r in x])
    out, stats = eg.truncate_response_json(body, 1024, renorm=False)
    got = json.loads(out)
    assert [len(d["embedding"]) for d in got["data"]] == [1024] * 64 and stats["resliced_vectors"] == 64
    assert np.array_equal(np.array([d["embedding"] for d in got["data"]]), x[:, :1024])


def _b64(v):
    import base64
    return base64.b64encode(np.asarray(v, dtype="<f4").tobytes()).decode()


def test_base64_vectors_truncate(oracle):
    """N4: with ETN_JSON_BASE64 a base64 float32 embedding is cut like an array; without the
    flag it is left alone, as the reference does (handlers.go:2073)."""
    rng = np.random.default_rng(4)
    vs = [rng.standard_normal(n).astype(np.float32) for n in (40, 16, 17, 3, 41)]
    items = [{"object": "embedding", "index": i, "embedding": _b64(v)} for i, v in enumerate(vs)]
    items += [{"embedding": "not base64!"}, {"embedding": "QUJD"}, {"embedding": ""}, {"embedding": [1.0, 2.0, 3.0]}]
    body = json.dumps({"data": items, "model": "m"}).encode()
    assert eg.truncate_response_json(body, 16, renorm=False)[0].count(b'"embedding":') == 9
    untouched, _ = eg.truncate_response_json(body, 16, renorm=False)
    assert json.loads(untouched)["data"][0]["embedding"] == items[0]["embedding"]           # flag off: reference behaviour
    for dims in (16, 1, 2, 5):
        out, stats = eg.truncate_response_json(body, dims, renorm=False, base64_vectors=True)
        got = json.loads(out)["data"]
        for g, it in zip(got, items):
            if isinstance(it["embedding"], str):
                assert g["embedding"] == oracle.b64_vector_ref(it["embedding"], dims, False), dims
        assert got[8]["embedding"] == [1.0, 2.0, 3.0][:dims]
        assert stats["gpu_vectors"] == 0


def test_fuzz_random_json_matches_reference(oracle):
    """Random JSON documents shaped like (and unlike) provider responses: the rewrite must
    equal decode -> truncateEmbeddings -> encode as a JSON value, for every dims."""
    rng = np.random.default_rng(77)

    def rnd_scalar():
        c = rng.integers(0, 7)
        if c == 0: return None
        if c == 1: return bool(rng.integers(0, 2))
        if c == 2: return int(rng.integers(-10**9, 10**9))
        if c == 3: return float(rng.standard_normal() * 10.0 ** rng.integers(-12, 12))
        if c == 4: return "".join(rng.choice(list('ab]["\\\\{},: é\n'), size=rng.integers(0, 8)))
        if c == 5: return float(rng.integers(-5, 5))
        return 1e-7 * float(rng.integers(1, 100))

    def rnd_value(depth):
        c = rng.integers(0, 10)
        if depth > 3 or c < 5: return rnd_scalar()
        if c < 8: return [rnd_value(depth + 1) for _ in range(rng.integers(0, 5))]
        return {rnd_key(): rnd_value(depth + 1) for _ in range(rng.integers(0, 4))}

    def rnd_key():
        return str(rng.choice(["data", "embedding", "index", "object", "x", "usage", "data ", "Embedding"]))

    def rnd_embedding():
        c = rng.integers(0, 8)
        if c == 0: return rnd_value(2)
        n = int(rng.integers(0, 12))
        v = [float(rng.standard_normal()) for _ in range(n)]
        if c == 1 and n: v[int(rng.integers(0, n))] = rnd_value(2)
        return v

    def rnd_item():
        c = rng.integers(0, 8)
        if c == 0: return rnd_value(1)
        it = {"index": int(rng.integers(0, 99))}
        if c != 1: it["embedding"] = rnd_embedding()
        if c == 2: it["extra"] = rnd_value(1)
        return it

    for trial in range(400):
        doc = {rnd_key() + str(i): rnd_value(1) for i in range(rng.integers(0, 3))}
        c = rng.integers(0, 10)
        if c < 8: doc["data"] = [rnd_item() for _ in range(rng.integers(0, 6))]
        elif c == 8: doc["data"] = rnd_value(1)
        body = json.dumps(doc, indent=int(rng.integers(0, 3)) or None, ensure_ascii=bool(rng.integers(0, 2))).encode()
        for dims in (1, 3, 8):
            out, _ = eg.truncate_response_json(body, dims, renorm=False)
            assert json.loads(out) == oracle.response_ref(body, dims, renorm=False), (trial, dims, body)


def test_prompt_tokens_like_the_handler():
    """handlers.go:2052-2057: usage.prompt_tokens is read as float64 and converted with int()."""
    f = eg.prompt_tokens
    assert f(b'{"data":[],"usage":{"prompt_tokens":17,"total_tokens":17}}') == 17
    assert f(b'{"usage":{"total_tokens":3, "prompt_tokens" : 12.9 }, "data":[{"embedding":[1,2]}]}') == 12
    assert f(b'{"usage":{"prompt_tokens":"17"}}') == 0 and f(b'{"usage":17}') == 0 and f(b'{"data":[]}') == 0 and f(b'[1]') == 0
    assert f(b'{"x":{"usage":{"prompt_tokens":5}}}') == 0                       # only the top-level usage counts
    assert f(b'{"usage":{"prompt_tokens":1},"usage":{"prompt_tokens":2}}') == 2   # last duplicate wins, as in Go
    assert f(b'{"usage":{"prompt_tokens":-3.5}}') == -3
    with pytest.raises(eg.EtnError):
        f(b'{"usage":{"prompt_tokens":1}')


def test_keys_spelled_with_escapes_are_the_keys_they_spell(oracle):
    """Go compares object keys after unescaping (handlers.go:2064, :2073 index the decoded map), so
    "d\\u0061ta" IS "data" to the reference — and to the packer: the vector is cut, a twin of
    "data" written that way is a duplicate, and an escape Go rejects fails the scan."""
    for body in (b'{"d\\u0061ta": [{"embedding": [1,2,3]}]}',
                 b'{"data": [{"emb\\u0065dd\\u0069ng": [1,2,3]}]}',
                 b'{"\\u0064\\u0061\\u0074\\u0061": [{"index": 0, "embedding": [1,2,3], "k\\"ey": 1, "\\u00e9": 2, "\\ud83d\\ude00": 3}]}'):
        out, stats = eg.truncate_response_json(body, 2, renorm=False)
        assert json.loads(out) == oracle.response_ref(body, 2, renorm=False) and stats["resliced_vectors"] == 1, body
    for body in (b'{"data": [], "d\\u0061ta": [{"embedding": [1,2,3]}]}',
                 b'{"data": [{"embedding": [1,2,3], "embeddin\\u0067": [4,5,6]}]}'):
        with pytest.raises(eg.EtnError) as ei:
            eg.truncate_response_json(body, 2, renorm=False)
        assert ei.value.code == eg.ERR_JSON_DUP, body
    for body in (b'{"d\\u00G1ta": 1}', b'{"d\\xta": 1}', b'{"data\\u00": 1}'):
        with pytest.raises(eg.EtnError) as ei:
            eg.truncate_response_json(body, 2, renorm=False)
        assert ei.value.code == eg.ERR_JSON, body
    assert eg.prompt_tokens(b'{"us\\u0061ge": {"prompt\\u005ftokens": 4}}') == 4
    # near misses stay misses
    body = b'{"dat\\u0061s": [{"embedding": [1,2,3]}], "Dat\\u0061": [{"embedding": [1,2,3]}]}'
    assert eg.truncate_response_json(body, 2, renorm=False)[0] == body


def test_numbers_outside_float64_fail_like_the_reference_decode():
    """json.Decode into map[string]any parses every number of the body with ParseFloat(s, 64);
    an overflow fails the request (handlers.go:2047-2050 -> 502).  TRUNCATE mode copies tokens
    without parsing them, so the scan itself has to notice — wherever the token stands."""
    for body in (b'{"data": [{"embedding": [1e999, 2, 3]}]}',
                 b'{"data": [{"embedding": [1, 2, 3, -1e400]}]}',               # in the tail that is cut away
                 b'{"data": [{"embedding": [1, 2, 3]}], "usage": {"x": 2e308}}',
                 b'{"data": [{"embedding": [1, 2, 3]}], "big": ' + b"9" * 400 + b'}',
                 b'{"data": [{"embedding": [0.0001e313, 2, 3]}]}'):
        with pytest.raises(eg.EtnError) as ei:
            eg.truncate_response_json(body, 2, renorm=False)
        assert ei.value.code == eg.ERR_JSON, body
    # the largest finite values, long digit strings that stay in range, and underflow (no error in Go) pass
    ok = b'{"data": [{"embedding": [1.7976931348623157e308, ' + b"1" * 300 + b', 1e-999, 0.' + b"0" * 400 + b'1, 4]}]}'
    out, _ = eg.truncate_response_json(ok, 4, renorm=False)
    assert json.loads(out)["data"][0]["embedding"][0] == 1.7976931348623157e308 and len(json.loads(out)["data"][0]["embedding"]) == 4
    buf = np.zeros(4)
    assert eg.lib().etn_json_parse_array(b"[1e-999, 4.9e-324, 2e-324, -1e-400]", 35, buf.ctypes.data, 4) == 4
    assert buf.tolist() == [0.0, 5e-324, 0.0, -0.0] and np.signbit(buf[3])
