"""HTTP-level known answers for the cloud branch of POST /v1/embeddings (SURVEY.md §8(c) KAT 11),
against llm-mcp_b200/handler.py.  Expected values are written from the reference source
(core/internal/api/handlers.go:1821-1896, 2018-2060; helpers.go:10-20) and, for the success
path, produced by the oracle's restatement (oracle.handle_embeddings_ref).  CPU only."""
import json

from llm_mcp_b200 import handler as h

ENV = {"OPENROUTER_API_KEY": "k", "OPENROUTER_BASE_URL": "http://mock/v1"}
VEC = [[0.1 * (i + 1) * (j + 1) for j in range(8)] for i in range(3)]


def provider(vectors=VEC, honour=False, status=200, raw=None):
    seen = []

    def post(url, headers, body):
        fwd = json.loads(body)
        seen.append((url, headers, body))
        if raw is not None:
            return status, raw
        vs = [v[:fwd["dimensions"]] for v in vectors] if honour and "dimensions" in fwd else vectors
        return status, json.dumps({"object": "list", "data": [{"object": "embedding", "index": i, "embedding": v} for i, v in enumerate(vs)],
                                   "model": fwd["model"], "usage": {"prompt_tokens": 9, "total_tokens": 9}}).encode()
    return post, seen


def test_error_contract_byte_for_byte():
    post, seen = provider()
    call = lambda m, b, env=ENV: h.handle_embeddings(m, b, env, post)
    assert call("GET", b"") == (405, {"Content-Type": "application/json"}, b'{"error":"method_not_allowed","message":"Only POST allowed"}\n')
    for bad in (b"{", b"[1]", b'{"model":5}', b'{"model":"a/b","dimensions":2.0}', b'{"model":"a/b","dimensions":"2"}'):
        assert call("POST", bad)[::2] == (400, b'{"error":"invalid_json","message":"Invalid JSON body"}\n'), bad
    assert call("POST", b'{"input":"x"}')[2] == b'{"error":"model_required","message":"Field \'model\' is required"}\n'
    assert call("POST", b'{"model":"a/b","input":5}')[2] == b'{"error":"invalid_input","message":"Field \'input\' must be a string or array of strings"}\n'
    assert call("POST", b'{"model":"a/b"}')[0] == 400
    assert call("POST", b'{"model":"a/b","input":[1,2]}')[2] == b'{"error":"empty_input","message":"Input is empty"}\n'
    for env in ({}, {"OPENROUTER_API_KEY": "not-used"}):
        assert call("POST", b'{"model":"a/b","input":"x"}', env)[::2] == (503, b'{"error":"no_api_key","message":"OPENROUTER_API_KEY not configured"}\n')
    assert not seen                                           # nothing above ever reached the provider
    assert h.handle_embeddings("POST", b'{"model":"nomic-embed-text","input":"x"}', ENV, post) is h.LOCAL_BRANCH


def test_forwarded_request_and_dims_resolution():
    for body, env, want in (
            (b'{"model":"q/e","input":["a",7,"b<&>"],"dimensions":4}', ENV, b'{"dimensions":4,"input":["a","b\\u003c\\u0026\\u003e"],"model":"q/e"}'),
            (b'{"MODEL":"q/e","Input":"solo"}', ENV, b'{"input":["solo"],"model":"q/e"}'),                     # keys match case-insensitively
            (b'{"model":"q/e","input":"x"}', dict(ENV, CLOUD_EMBED_DIMENSIONS="3"), b'{"dimensions":3,"input":["x"],"model":"q/e"}'),
            (b'{"model":"q/e","input":"x","dimensions":5}', dict(ENV, CLOUD_EMBED_DIMENSIONS="3"), b'{"dimensions":5,"input":["x"],"model":"q/e"}'),
            (b'{"model":"q/e","input":"x","dimensions":-2}', dict(ENV, CLOUD_EMBED_DIMENSIONS="3"), b'{"input":["x"],"model":"q/e"}'),   # negative: passes through, disables both
            (b'{"model":"q/e","input":"x"}', dict(ENV, CLOUD_EMBED_DIMENSIONS="junk"), b'{"input":["x"],"model":"q/e"}')):
        post, seen = provider()
        status, _, out = h.handle_embeddings("POST", body, env, post)
        assert status == 200 and out.endswith(b"\n")
        url, headers, fwd = seen[0]
        assert url == "http://mock/v1/embeddings" and headers == {"Content-Type": "application/json", "Authorization": "Bearer k"}
        assert fwd == want
    post, seen = provider()
    h.handle_embeddings("POST", b'{"model":"q/e","input":"x"}', {"OPENROUTER_API_KEY": "k"}, post)
    assert seen[0][0] == "https://openrouter.ai/api/v1/embeddings"


def _same_as_reference(oracle, req, env, **pkw):
    post, _ = provider(**pkw)
    m = h.Metrics()
    status, _, out = h.handle_embeddings("POST", json.dumps(req).encode(), env, post, m)
    want_status, want, _ = oracle.handle_embeddings_ref(req, env, lambda fwd: post("", {}, json.dumps(fwd).encode()))
    assert status == want_status and json.loads(out) == want
    return json.loads(out), m


def test_provider_ignores_dimensions_so_the_handler_cuts(oracle):
    got, m = _same_as_reference(oracle, {"model": "q/e", "input": ["a", "b", "c"], "dimensions": 3}, ENV)
    assert [d["embedding"] for d in got["data"]] == [v[:3] for v in VEC] and got["usage"]["prompt_tokens"] == 9
    assert m.requests == {("q/e", "openrouter", "ok"): 1} and m.input_tokens == {("q/e", "openrouter"): 9.0}
    assert len(m.duration[("q/e", "openrouter")]) == 1


def test_provider_honours_dimensions_no_op(oracle):
    got, _ = _same_as_reference(oracle, {"model": "q/e", "input": "a", "dimensions": 3}, ENV, honour=True)
    assert all(len(d["embedding"]) == 3 for d in got["data"])


def test_no_dimensions_nothing_is_cut(oracle):
    got, _ = _same_as_reference(oracle, {"model": "q/e", "input": "a"}, ENV)
    assert [d["embedding"] for d in got["data"]] == VEC


def test_base64_strings_ragged_lengths_and_odd_items(oracle):
    raw = json.dumps({"data": [{"index": 0, "embedding": "AAAAAAAAgD8="}, {"index": 1, "embedding": VEC[0]}, {"index": 2, "embedding": VEC[1][:2]},
                               "junk", {"index": 4, "embedding": None}, {"index": 5, "embedding": [1, "x", 2, 3, 4]}],
                      "usage": {"prompt_tokens": 2.9}}).encode()
    got, m = _same_as_reference(oracle, {"model": "q/e", "input": "a", "dimensions": 4}, ENV, raw=raw)
    assert got["data"][0]["embedding"] == "AAAAAAAAgD8=" and got["data"][1]["embedding"] == VEC[0][:4]
    assert got["data"][2]["embedding"] == VEC[1][:2] and got["data"][5]["embedding"] == [1, "x", 2, 3]
    assert m.input_tokens == {("q/e", "openrouter"): 2.0}     # int(pt): truncation


def test_bodies_only_the_reference_path_can_judge(oracle):
    dup = b'{"data":[{"embedding":[1,2,3]}],"data":[{"embedding":[4,5,6,7]}],"usage":{"prompt_tokens":1}}'
    got, _ = _same_as_reference(oracle, {"model": "q/e", "input": "a", "dimensions": 2}, ENV, raw=dup)
    assert got["data"] == [{"embedding": [4, 5]}]             # Go keeps the last duplicate
    # ... and that path re-encodes like json.Marshal: keys sorted, HTML-unsafe characters escaped, Go's float spelling
    odd = b'{"b":[1,2.5,-0.0,1e-7,3],"a":{"z":"<x>","y":null,"k":true},"data":[],"data":[{"embedding":[4,5,6,7]}]}'
    post, _ = provider(raw=odd)
    _, _, out = h.handle_embeddings("POST", b'{"model":"q/e","input":"a","dimensions":2}', ENV, post)
    assert out == b'{"a":{"k":true,"y":null,"z":"\\u003cx\\u003e"},"b":[1,2.5,-0,1e-7,3],"data":[{"embedding":[4,5]}]}\n'
    got, _ = _same_as_reference(oracle, {"model": "q/e", "input": "a", "dimensions": 2}, ENV, raw=b" null ")
    assert got is None


def test_provider_failures_become_502(oracle):
    m = h.Metrics()
    post, _ = provider(status=429, raw=b"slow down " * 200)
    status, _, out = h.handle_embeddings("POST", b'{"model":"q/e","input":"a"}', ENV, post, m)
    msg = json.loads(out)
    assert status == 502 and msg["error"] == "cloud_embed_failed"
    assert msg["message"] == "cloud embed returned 429: " + ("slow down " * 200)[:1024]
    for raw in (b"{not json", b"[1,2]", b'{"data":[{"embedding":[1e999]}]}'):
        post, _ = provider(raw=raw)
        status, _, out = h.handle_embeddings("POST", b'{"model":"q/e","input":"a","dimensions":1}', ENV, post, m)
        assert status == 502 and json.loads(out)["message"].startswith("decode cloud embed response"), raw

    def down(url, headers, body):
        raise ConnectionRefusedError("connection refused")
    status, _, out = h.handle_embeddings("POST", b'{"model":"q/e","input":"a"}', ENV, down, m)
    assert status == 502 and json.loads(out)["message"].startswith("cloud embed request failed: ")
    assert m.requests == {("q/e", "openrouter", "error"): 5} and not m.duration and not m.input_tokens


def test_over_real_http_with_a_mock_provider(oracle):
    """The default transport against a provider on 127.0.0.1 (OPENROUTER_BASE_URL is how the
    reference would be pointed at a mock, handlers.go:1869)."""
    import threading
    from http.server import BaseHTTPRequestHandler, HTTPServer

    class Mock(BaseHTTPRequestHandler):
        def do_POST(self):
            fwd = json.loads(self.rfile.read(int(self.headers["Content-Length"])))
            assert self.path == "/api/v1/embeddings" and self.headers["Authorization"] == "Bearer secret"
            body = json.dumps({"data": [{"index": 0, "embedding": VEC[2]}], "usage": {"prompt_tokens": len(fwd["input"])}}).encode()
            self.send_response(200 if fwd["model"] != "q/broken" else 500)
            self.send_header("Content-Type", "application/json")
            self.send_header("Content-Length", str(len(body)))
            self.end_headers()
            self.wfile.write(body)

        def log_message(self, *a):
            pass

    srv = HTTPServer(("127.0.0.1", 0), Mock)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    try:
        env = {"OPENROUTER_API_KEY": "secret", "OPENROUTER_BASE_URL": "http://127.0.0.1:%d/api/v1" % srv.server_port}
        status, _, out = h.handle_embeddings("POST", b'{"model":"q/e","input":["a","b"],"dimensions":5}', env)
        assert status == 200 and json.loads(out) == {"data": [{"index": 0, "embedding": VEC[2][:5]}], "usage": {"prompt_tokens": 2}}
        status, _, out = h.handle_embeddings("POST", b'{"model":"q/broken","input":"a"}', env)
        assert status == 502 and json.loads(out)["message"].startswith("cloud embed returned 500: ")
    finally:
        srv.shutdown()
