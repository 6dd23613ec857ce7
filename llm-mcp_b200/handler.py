"""handler — the cloud branch of ``Server.HandleEmbeddings`` around the GPU path.

A mirror of the reference handler, statement by statement
(``core/internal/api/handlers.go:1821-1896`` and ``proxyOpenAIEmbed``, ``:2018-2060``), with the
three lines that matter replaced: the provider's body is NOT decoded into ``map[string]any``
(``:2047-2050``), cut by ``truncateEmbeddings`` (``:1891-1893``) and re-encoded
(``helpers.go:11-15``); it goes through ``etn_json_truncate_response`` as text, and
``usage.prompt_tokens`` comes from ``etn_json_prompt_tokens``.  Everything else — request
validation, input normalisation, ``dims`` resolution, the forwarded body, the error contract,
the three embedding metrics — is the reference's behaviour restated, so that the HTTP-level
known-answer tests of SURVEY.md §8(c) item 11 can run against it (the Go handler itself cannot
be built here: no Go toolchain).  The local/Ollama branch is not mirrored here: it is returned
to the caller untouched (``LOCAL_BRANCH``).

When the packer says "take the reference path" (ETN_ERR_JSON / ETN_ERR_JSON_DUP: duplicate keys,
input only Go's decoder can judge) the response is produced the reference's way: decode,
reslice, re-encode.  That path does no arithmetic; with EMBED_RENORMALIZE=1 its vectors are
merely cut and the fallback counter says so.
"""
from __future__ import annotations

import json
import time
import urllib.error
import urllib.request
from typing import Callable

from . import embedgpu as eg

EMBED_TIMEOUT_S = 120.0                     # handlers.go:1816
LOCAL_BRANCH = object()                     # model without "/": the reference's Ollama branch, not this module's business

# post(url, headers, body) -> (status, body bytes); raises OSError on transport failure
Post = Callable[[str, dict, bytes], "tuple[int, bytes]"]


def _go_json_string(s: str) -> str:
    """encoding/json's string encoding: HTML-unsafe characters and U+2028/2029 escaped."""
    out = json.dumps(s, ensure_ascii=False)
    return (out.replace("<", "\\u003c").replace(">", "\\u003e").replace("&", "\\u0026")
               .replace("\u2028", "\\u2028").replace("\u2029", "\\u2029"))


def write_json(status: int, payload_text: str) -> tuple[int, dict, bytes]:
    """WriteJSON (helpers.go:10-15): Content-Type, status, the encoder's trailing newline."""
    return status, {"Content-Type": "application/json"}, payload_text.encode() + b"\n"


def write_error(status: int, code: str, message: str) -> tuple[int, dict, bytes]:
    """WriteError (helpers.go:18-20) with models.ErrorResp (types.go:14-18): empty message omitted."""
    body = '{"error":' + _go_json_string(code) + (',"message":' + _go_json_string(message) if message else "") + "}"
    return write_json(status, body)


class _Obj(list):
    """A JSON object as its (key, value) pairs, in document order, duplicates kept."""


class _Num(str):
    """A JSON number as its text; `is_int` tells an integer literal from one with a fraction or exponent."""
    is_int = False


def _num(text: str, is_int: bool) -> _Num:
    n = _Num(text)
    n.is_int = is_int
    return n


def _plain(v):
    """What Go decodes into an `any`: float64 numbers, map[string]any (the last duplicate wins), []any."""
    if isinstance(v, _Num):
        return float(v)
    if isinstance(v, _Obj):
        return {k: _plain(x) for k, x in v}
    if isinstance(v, list):
        return [_plain(x) for x in v]
    return v


def _decode_request(raw: bytes):
    """json.NewDecoder(r.Body).Decode(&req) into {Model string; Input any; EncodingFormat string;
    Dimensions int} (handlers.go:1828-1837).  Keys match case-insensitively (the last one wins),
    null leaves a field at its zero value, a value of the wrong JSON type is an error, and an
    int field only takes an integer literal (no fraction, no exponent).  None = decode error."""
    try:
        dec = json.JSONDecoder(object_pairs_hook=_Obj, parse_float=lambda t: _num(t, False), parse_int=lambda t: _num(t, True))
        value, _ = dec.raw_decode(raw.decode("utf-8", errors="replace").lstrip(" \t\r\n"))
    except (ValueError, RecursionError):
        return None
    req = {"model": "", "input": None, "encoding_format": "", "dimensions": 0}
    if value is None:
        return req
    if not isinstance(value, _Obj):
        return None
    for key, v in value:
        lk = key.lower()
        if lk in ("model", "encoding_format"):
            if v is None:
                continue
            if not isinstance(v, str) or isinstance(v, _Num):
                return None
            req[lk] = v
        elif lk == "dimensions":
            if v is None:
                continue
            if not (isinstance(v, _Num) and v.is_int):
                return None
            n = int(v)
            if not -(1 << 63) <= n < (1 << 63):
                return None
            req[lk] = n
        elif lk == "input":
            req[lk] = _plain(v)
    return req


def _default_post(url: str, headers: dict, body: bytes) -> tuple[int, bytes]:
    req = urllib.request.Request(url, data=body, headers=headers, method="POST")
    try:
        with urllib.request.urlopen(req, timeout=EMBED_TIMEOUT_S) as resp:
            return resp.status, resp.read()
    except urllib.error.HTTPError as e:              # a non-2xx answer is an answer, not a transport failure
        return e.code, e.read()


class Metrics:
    """The three collectors the branch touches (metrics.go:10-26, 64-70), as plain dicts."""

    def __init__(self):
        self.requests, self.duration, self.input_tokens = {}, {}, {}

    def inc(self, model, device, status):
        self.requests[(model, device, status)] = self.requests.get((model, device, status), 0) + 1

    def observe(self, model, device, seconds):
        self.duration.setdefault((model, device), []).append(seconds)

    def add_tokens(self, model, device, n):
        self.input_tokens[(model, device)] = self.input_tokens.get((model, device), 0.0) + float(n)


def _go_float(text: str) -> float:
    v = float(text)
    if v in (float("inf"), float("-inf")):             # strconv.ParseFloat: out of range -> Decode fails
        raise ValueError("number out of float64 range")
    return v


def _not_json(name):
    raise ValueError(name + " is not JSON")            # Python's decoder would take NaN / Infinity


def _loads(body: bytes):
    """json.Decode into map[string]any, as far as values go: numbers are float64 or an error."""
    return json.loads(body, parse_float=_go_float, parse_int=_go_float, parse_constant=_not_json)


def _go_encode(v) -> str:
    """json.Marshal of a decoded value: object keys sorted, strings HTML-escaped, float64 in
    encoding/json's spelling (the library's own formatter, helpers.go:14)."""
    import ctypes
    if v is None:
        return "null"
    if v is True:
        return "true"
    if v is False:
        return "false"
    if isinstance(v, float):
        buf = ctypes.create_string_buffer(48)
        one = (ctypes.c_double * 1)(v)
        n = eg.lib().etn_json_format_array(one, 1, buf, 48)
        if n < 3:
            raise ValueError("unsupported value")          # NaN/Inf: Go's encoder refuses them too
        return buf.raw[1:n - 1].decode()
    if isinstance(v, str):
        return _go_json_string(v)
    if isinstance(v, list):
        return "[" + ",".join(_go_encode(x) for x in v) + "]"
    if isinstance(v, dict):
        return "{" + ",".join(_go_json_string(k) + ":" + _go_encode(v[k]) for k in sorted(v)) + "}"
    raise ValueError("unsupported value")


def _reference_path(body: bytes, dims: int):
    """decode -> truncateEmbeddings -> encode, the way the reference does it (handlers.go:2047-2050,
    2062-2078, helpers.go:14), for bodies the packer hands back.  Returns the response text or None
    when the body does not decode into a JSON object (Go: "decode cloud embed response")."""
    try:
        result = _loads(body)
    except (ValueError, RecursionError):
        return None
    if result is None:
        return "null"                                  # a nil map encodes as null
    if not isinstance(result, dict):
        return None
    if dims > 0:
        eg.truncate_embeddings(result, dims, renorm=False)     # the O(1) reslice, no GPU involved
    return _go_encode(result)


def handle_embeddings(method: str, raw_body: bytes, env=None, post: Post | None = None, metrics: Metrics | None = None):
    """POST /v1/embeddings, cloud branch.  Returns (status, headers, body bytes), or LOCAL_BRANCH
    for a model without "/" (handlers.go:1867: that request belongs to the Ollama branch)."""
    import os
    env = os.environ if env is None else env
    post = _default_post if post is None else post
    if method != "POST":                               # handlers.go:1823-1826
        return write_error(405, "method_not_allowed", "Only POST allowed")
    req = _decode_request(raw_body)                    # :1828-1837
    if req is None:
        return write_error(400, "invalid_json", "Invalid JSON body")
    if req["model"] == "":                             # :1838-1841
        return write_error(400, "model_required", "Field 'model' is required")
    inp = req["input"]                                 # :1843-1861: a string, or the strings of an array
    if isinstance(inp, str):
        texts = [inp]
    elif isinstance(inp, list):
        texts = [t for t in inp if isinstance(t, str)]
    else:
        return write_error(400, "invalid_input", "Field 'input' must be a string or array of strings")
    if not texts:
        return write_error(400, "empty_input", "Input is empty")
    model = req["model"]
    if "/" not in model:                               # :1867
        return LOCAL_BRANCH
    api_key = env.get("OPENROUTER_API_KEY", "") or ""  # :1868-1873 (config.Getenv: empty -> default)
    api_url = env.get("OPENROUTER_BASE_URL", "") or "https://openrouter.ai/api/v1"
    if api_key in ("", "not-used"):
        return write_error(503, "no_api_key", "OPENROUTER_API_KEY not configured")
    dims = eg.resolve_dims(req["dimensions"], env)     # :1874-1877
    metrics = metrics if metrics is not None else Metrics()

    # proxyOpenAIEmbed (:2018-2060): json.Marshal of the map sorts the keys; dimensions only when > 0
    fields = {"input": "[" + ",".join(_go_json_string(t) for t in texts) + "]", "model": _go_json_string(model)}
    if dims > 0:
        fields["dimensions"] = str(dims)
    fwd = ("{" + ",".join(f'"{k}":{fields[k]}' for k in sorted(fields)) + "}").encode()
    start = time.perf_counter()
    err = None
    out_text = None
    tokens = 0
    try:
        status, body = post(api_url + "/embeddings", {"Content-Type": "application/json", "Authorization": "Bearer " + api_key}, fwd)
    except OSError as e:
        err = f"cloud embed request failed: {e}"
    else:
        if status != 200:
            err = f"cloud embed returned {status}: {body[:1024].decode('utf-8', errors='replace')}"
        else:
            out_text, tokens, err = _rewrite(body, dims, env)
    elapsed = time.perf_counter() - start
    if err is not None:                                # :1881-1886
        metrics.inc(model, "openrouter", "error")
        return write_error(502, "cloud_embed_failed", err)
    metrics.inc(model, "openrouter", "ok")             # :1887-1889
    metrics.observe(model, "openrouter", elapsed)
    metrics.add_tokens(model, "openrouter", tokens)
    return 200, {"Content-Type": "application/json"}, out_text + b"\n"      # WriteJSON, :1894


def _rewrite(body: bytes, dims: int, env):
    """The provider's 200 body -> (response bytes, prompt tokens, error text).  Stands where the
    reference decodes (:2047-2050), reads usage.prompt_tokens (:2052-2057), cuts (:1891-1893) and
    re-encodes (helpers.go:14)."""
    decode_error = "decode cloud embed response: invalid JSON"
    stripped = body.lstrip(b" \t\r\n")
    if not stripped.startswith(b"{"):                  # Decode(&map): only an object (or null) will do
        text = _reference_path(body, dims)
        return (None, 0, decode_error) if text is None else (text.encode(), 0, None)
    try:
        tokens = eg.prompt_tokens(body)
        out, _ = eg.truncate_response_json(body, dims, renorm=eg.renorm_enabled(env))
        return out.rstrip(b" \t\r\n"), tokens, None
    except eg.EtnError as e:
        if e.code not in (eg.ERR_JSON, eg.ERR_JSON_DUP):
            # the GPU path failed (not bound, CUDA error): the request still succeeds, the reference's way
            eg.lib().etn_note_fallback(1)
    text = _reference_path(body, dims)
    if text is None:
        return None, 0, decode_error
    result = _loads(body)
    usage = result.get("usage") if isinstance(result, dict) else None
    pt = usage.get("prompt_tokens") if isinstance(usage, dict) else None
    tokens = int(pt) if isinstance(pt, (int, float)) and not isinstance(pt, bool) else 0
    return text.encode(), tokens, None
