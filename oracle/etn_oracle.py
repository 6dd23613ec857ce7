"""ctypes binding for oracle/libetn_oracle.so plus the map-level restatement of
``truncateEmbeddings`` (reference: core/internal/api/handlers.go:2062-2078).

TEST INFRASTRUCTURE ONLY — see the header of etn_oracle.c.  Imported by tests/,
``__graft_entry__.smoke()`` and bench.py's cpu_baseline / ``--impl reference`` legs; the
product package (llm-mcp_b200/) never imports it.  PARITY UNPINNED: the reference holds no
golden vectors for this path.
"""
from __future__ import annotations

import ctypes
import json
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libetn_oracle.so")

F32, F64 = 0, 1
TRUNCATE, TRUNCATE_RENORM = 0, 1
_NP = {F32: np.float32, F64: np.float64}


def build(force: bool = False) -> str:
    """Compile the C oracle with gcc (oracle/Makefile)."""
    src = os.path.join(_HERE, "etn_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = ctypes.CDLL(_SO)
        i64, vp, ci = ctypes.c_int64, ctypes.c_void_p, ctypes.c_int
        L.etn_oracle_fill.argtypes = [ctypes.c_uint64, i64, i64, i64, i64, i64, ci, vp]
        L.etn_oracle_truncate.argtypes = [i64, i64, i64, ci, vp, vp]
        L.etn_oracle_renorm_exact.argtypes = [i64, i64, i64, ci, vp, vp]
        L.etn_oracle_ragged.argtypes = [i64, i64, vp, vp, ci, ci, vp, vp]
        L.etn_oracle_renorm_go.argtypes = [i64, i64, i64, ci, vp, vp, ci]
        L.etn_oracle_reslice.argtypes = [i64, i64, vp]
        L.etn_oracle_reslice.restype = i64
        L.etn_oracle_max_ulp.argtypes = [i64, ci, vp, vp, ctypes.POINTER(i64)]
        L.etn_oracle_max_ulp.restype = i64
        _lib = L
    return _lib


def dtype_code(dt) -> int:
    dt = np.dtype(dt)
    if dt == np.float32:
        return F32
    if dt == np.float64:
        return F64
    raise TypeError(f"unsupported dtype {dt}")


def _ptr(a: np.ndarray) -> ctypes.c_void_p:
    return ctypes.c_void_p(a.ctypes.data)


def _check(rc: int, what: str) -> None:
    if rc < 0:
        raise ValueError(f"{what}: oracle error {rc}")


def fill(seed: int, rows: int, D: int, dtype, row0: int = 0, ncols: int | None = None) -> np.ndarray:
    """Synthetic matrix of SURVEY.md §8(d): rows [row0, row0+rows) x cols [0, ncols)."""
    ncols = D if ncols is None else ncols
    out = np.empty((rows, ncols), dtype=dtype)
    _check(lib().etn_oracle_fill(seed, row0, rows, D, ncols, ncols, dtype_code(dtype), _ptr(out)), "fill")
    return out


def fill_numpy(seed: int, rows: int, D: int, dtype, row0: int = 0, ncols: int | None = None) -> np.ndarray:
    """Same generator in numpy integer arithmetic (cross-check of the C one)."""
    ncols = D if ncols is None else ncols
    r = np.arange(row0, row0 + rows, dtype=np.uint64)[:, None]
    c = np.arange(ncols, dtype=np.uint64)[None, :]
    with np.errstate(over="ignore"):
        z = np.uint64(seed) ^ (r * np.uint64(D) + c)
        z = z + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    v = ((z >> np.uint64(32)).astype(np.uint32).view(np.int32) >> 8).astype(np.float64) * 2.0 ** -28
    return v.astype(dtype)


def truncate(x: np.ndarray, k: int) -> np.ndarray:
    x = np.ascontiguousarray(x)
    batch, D = x.shape
    out = np.empty((batch, min(k, D)), dtype=x.dtype)
    _check(lib().etn_oracle_truncate(batch, D, k, dtype_code(x.dtype), _ptr(x), _ptr(out)), "truncate")
    return out


def renorm_exact(x: np.ndarray, k: int) -> np.ndarray:
    x = np.ascontiguousarray(x)
    batch, D = x.shape
    out = np.empty((batch, min(k, D)), dtype=x.dtype)
    _check(lib().etn_oracle_renorm_exact(batch, D, k, dtype_code(x.dtype), _ptr(x), _ptr(out)), "renorm_exact")
    return out


def renorm_go(x: np.ndarray, k: int, nthreads: int = 1, out: np.ndarray | None = None) -> np.ndarray:
    x = np.ascontiguousarray(x)
    batch, D = x.shape
    if out is None:
        out = np.empty((batch, min(k, D)), dtype=x.dtype)
    _check(lib().etn_oracle_renorm_go(batch, D, k, dtype_code(x.dtype), _ptr(x), _ptr(out), nthreads), "renorm_go")
    return out


_fast = None


def renorm_fast(x: np.ndarray, k: int, nthreads: int = 1, out: np.ndarray | None = None) -> np.ndarray:
    """Vectorised "best CPU" baseline (etn_oracle_fast.c); timing only, not the oracle of record."""
    global _fast
    if _fast is None:
        so = os.path.join(_HERE, "libetn_oracle_fast.so")
        if not os.path.exists(so):
            subprocess.check_call(["make", "-C", _HERE, "-s", "libetn_oracle_fast.so"])
        _fast = ctypes.CDLL(so)
        _fast.etn_oracle_renorm_fast.argtypes = [ctypes.c_int64] * 3 + [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    x = np.ascontiguousarray(x)
    batch, D = x.shape
    if out is None:
        out = np.empty((batch, min(k, D)), dtype=x.dtype)
    _check(_fast.etn_oracle_renorm_fast(batch, D, k, dtype_code(x.dtype), _ptr(x), _ptr(out), nthreads), "renorm_fast")
    return out


def ragged_offsets(k_per_row: np.ndarray, D: int) -> np.ndarray:
    """Exclusive prefix sum of min(k, D): CSR element offsets, length batch+1."""
    kk = np.minimum(np.asarray(k_per_row, dtype=np.int64), D)
    off = np.zeros(len(kk) + 1, dtype=np.int64)
    np.cumsum(kk, out=off[1:])
    return off


def ragged(x: np.ndarray, k_per_row: np.ndarray, mode: int) -> tuple[np.ndarray, np.ndarray]:
    x = np.ascontiguousarray(x)
    batch, D = x.shape
    kpr = np.ascontiguousarray(k_per_row, dtype=np.int32)
    off = ragged_offsets(kpr, D)
    out = np.empty(int(off[-1]), dtype=x.dtype)
    _check(lib().etn_oracle_ragged(batch, D, _ptr(kpr), _ptr(off), dtype_code(x.dtype), mode, _ptr(x), _ptr(out)), "ragged")
    return out, off


def shard_ragged_ref(k_per_row, D: int, world: int) -> list[tuple[int, int]]:
    """Contiguous row ranges of a ragged batch balanced by kept elements:
    shard g starts at the first row r with (kept elements before r) * world >= total * g.
    Plain Python integers; the checker for etn_shard_ragged and for the cuts
    embed_truncate_norm_ragged_host makes.  Each vector is independent in the reference
    (handlers.go:2068-2076), so any contiguous cut is legal; this one is the balanced one."""
    kept = [min(max(int(k), 0), D) for k in k_per_row]
    csum = [0]
    for v in kept:
        csum.append(csum[-1] + v)
    n, total = len(kept), csum[-1]
    cuts = [0]
    for g in range(1, world):
        r = cuts[-1]
        while r < n and csum[r] * world < total * g:
            r += 1
        cuts.append(r)
    cuts.append(n)
    return [(cuts[g], cuts[g + 1]) for g in range(world)]


def max_ulp(a: np.ndarray, b: np.ndarray) -> tuple[int, int]:
    """(largest ulp distance, flat index); distance -1 means a NaN was seen."""
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    assert a.dtype == b.dtype and a.size == b.size, (a.dtype, b.dtype, a.shape, b.shape)
    arg = ctypes.c_int64(0)
    d = lib().etn_oracle_max_ulp(a.size, dtype_code(a.dtype), _ptr(a), _ptr(b), ctypes.byref(arg))
    return int(d), int(arg.value)


def reslice(lens: np.ndarray, dims: int) -> int:
    lens = np.ascontiguousarray(lens, dtype=np.int64)
    return int(lib().etn_oracle_reslice(len(lens), dims, _ptr(lens)))


# ---------------------------------------------------------------------------
# Map-level restatement of the reference function (pure Python; small cases only)
# ---------------------------------------------------------------------------

def truncate_embeddings_ref(result, dims: int) -> None:
    """``truncateEmbeddings(result map[string]any, dims int)`` — handlers.go:2062-2078.

    JSON model: object → dict, array → list, number → float.  Mutates ``result`` in place:
      :2064-2067  ``data`` missing or not an array → no-op
      :2069-2072  items that are not objects are skipped
      :2073       ``embedding`` that is not an array (base64 string, null) is left alone
      :2074       only vectors with len > dims are cut; shorter ones are never padded
      :2075       keep the first ``dims`` components, same order, same values
    Nothing is renormalised.
    """
    if not isinstance(result, dict):
        return
    data = result.get("data")
    if not isinstance(data, list):
        return
    for item in data:
        if not isinstance(item, dict):
            continue
        emb = item.get("embedding")
        if isinstance(emb, list) and len(emb) > dims:
            item["embedding"] = emb[:dims]


def resolve_dims(req_dimensions: int, env: dict | None = None) -> int:
    """handlers.go:1874-1877 + config.go:17-24: request value, else CLOUD_EMBED_DIMENSIONS
    (unset, empty or unparsable → 0).  The op runs only when the result is > 0 (:1891)."""
    env = os.environ if env is None else env
    dims = req_dimensions
    if dims == 0:
        v = env.get("CLOUD_EMBED_DIMENSIONS", "")
        if v != "":
            try:
                dims = _go_atoi(v)
            except ValueError:
                dims = 0
    return dims


def _go_atoi(s: str) -> int:
    """strconv.Atoi: optional sign then decimal digits only (no spaces, no underscores)."""
    body = s[1:] if s[:1] in "+-" else s
    if not body or not body.isascii() or not body.isdigit():
        raise ValueError(s)
    n = int(s)
    if not -(1 << 63) <= n < (1 << 63):
        raise ValueError(s)
    return n


# ---------------------------------------------------------------------------
# encoding/json number formatting and the response-level restatement (N1 oracle)
# ---------------------------------------------------------------------------

def go_format_float(x: float) -> str:
    """encoding/json's floatEncoder for float64 (Go 1.22 encode.go): strconv.AppendFloat with
    the shortest round-trip digits, format 'f' unless |x| < 1e-6 or |x| >= 1e21, then 'e' with
    a two-digit negative exponent "e-0d" shortened to "e-d".  Restated from Python's repr,
    which yields the same shortest digit string."""
    import math
    from decimal import Decimal
    if math.isnan(x) or math.isinf(x):
        raise ValueError("json: unsupported value")
    if x == 0:
        return "-0" if math.copysign(1.0, x) < 0 else "0"
    sign, digits, exp = Decimal(repr(x)).as_tuple()
    digits = list(digits)
    while len(digits) > 1 and digits[-1] == 0:
        digits.pop(); exp += 1
    ds = "".join(map(str, digits))
    e10 = exp + len(ds) - 1                      # decimal exponent of the first digit
    neg = "-" if sign else ""
    a = abs(x)
    if a < 1e-6 or a >= 1e21:
        mant = ds[0] + ("." + ds[1:] if len(ds) > 1 else "")
        ax = abs(e10)
        es = ("-" if e10 < 0 else "+") + (str(ax) if (ax >= 10 or e10 < 0) else "0" + str(ax))
        return f"{neg}{mant}e{es}"
    if e10 < 0:
        return neg + "0." + "0" * (-e10 - 1) + ds
    if len(ds) <= e10 + 1:
        return neg + ds + "0" * (e10 + 1 - len(ds))
    return neg + ds[:e10 + 1] + "." + ds[e10 + 1:]


def response_ref(body: bytes, dims: int, renorm: bool):
    """What the handler returns, as a decoded value: json.Decode (handlers.go:2047-2050),
    truncateEmbeddings when dims > 0 (:1891-1893), and — only with renorm — the exact
    renormalisation of every cut vector made of numbers."""
    import json
    result = json.loads(body)
    if dims > 0:
        data = result.get("data") if isinstance(result, dict) else None
        before = {id(it): len(it["embedding"]) for it in (data if isinstance(data, list) else [])
                  if isinstance(it, dict) and isinstance(it.get("embedding"), list)}
        truncate_embeddings_ref(result, dims)
        if renorm and isinstance(result, dict) and isinstance(result.get("data"), list):
            for it in result["data"]:
                if isinstance(it, dict) and isinstance(it.get("embedding"), list) and before.get(id(it), 0) > dims \
                        and all(type(v) in (int, float) for v in it["embedding"]):
                    v = np.array([[float(t) for t in it["embedding"]]], dtype=np.float64)
                    it["embedding"] = renorm_exact(v, dims)[0].tolist()
    return result


def b64_vector_ref(text: str, dims: int, renorm: bool):
    """OpenAI encoding_format "base64": the string carries little-endian float32 (N4).  Returns
    the string the packer must produce with ETN_JSON_BASE64: cut to dims floats when longer,
    exactly renormalised when asked, untouched when it is not a float32 vector or not longer."""
    import base64
    import binascii
    try:
        raw = base64.b64decode(text, validate=True)
    except (binascii.Error, ValueError):
        return text
    if len(raw) == 0 or len(raw) % 4 != 0 or dims <= 0 or len(raw) // 4 <= dims:
        return text
    v = np.frombuffer(raw, dtype="<f4")[:dims].astype(np.float32)
    if renorm:
        v = renorm_exact(v[None, :], dims)[0]
    return base64.b64encode(v.astype("<f4").tobytes()).decode()


# ---------------------------------------------------------------------------
# Handler-level restatement (pure Python): the cloud branch of HandleEmbeddings
# ---------------------------------------------------------------------------

def handle_embeddings_ref(request: dict, env: dict, provider):
    """core/internal/api/handlers.go:1821-1896 + proxyOpenAIEmbed :2018-2060 on an already decoded,
    well-formed request dict {"model", "input", "dimensions"?}: what the reference forwards to the
    provider and what it answers.  `provider(forwarded: dict) -> (status, body bytes)`.
    Returns (status, decoded JSON value of the response body, forwarded dict or None).
    Test infrastructure: the checker for llm-mcp_b200/handler.py."""
    model, inp = request.get("model", ""), request.get("input")
    def err(status, code, message):
        return status, {"error": code, "message": message}, None
    if model == "":
        return err(400, "model_required", "Field 'model' is required")
    if isinstance(inp, str):
        texts = [inp]
    elif isinstance(inp, list):
        texts = [t for t in inp if isinstance(t, str)]                  # :1849-1853: other items silently dropped
    else:
        return err(400, "invalid_input", "Field 'input' must be a string or array of strings")
    if not texts:
        return err(400, "empty_input", "Input is empty")
    if "/" not in model:
        raise ValueError("local branch: not restated here")
    key = env.get("OPENROUTER_API_KEY", "")
    if key in ("", "not-used"):
        return err(503, "no_api_key", "OPENROUTER_API_KEY not configured")
    dims = resolve_dims(int(request.get("dimensions", 0) or 0), env)    # :1874-1877
    forwarded = {"model": model, "input": texts}
    if dims > 0:                                                        # :2023-2025
        forwarded["dimensions"] = dims
    status, body = provider(forwarded)
    if status != 200:                                                   # :2042-2045
        msg = "cloud embed returned %d: %s" % (status, body[:1024].decode("utf-8", errors="replace"))
        return 502, {"error": "cloud_embed_failed", "message": msg}, forwarded
    try:
        result = json.loads(body)                                       # :2047-2050
        if result is not None and not isinstance(result, dict):
            raise ValueError("not an object")
    except ValueError:
        return 502, {"error": "cloud_embed_failed", "message": "decode cloud embed response: ..."}, forwarded
    if dims > 0 and result is not None:                                 # :1891-1893
        truncate_embeddings_ref(result, dims)
    return 200, result, forwarded
