"""embedgpu — host side above the C-ABI of libembed_truncate_norm.so.

This module mirrors, name for name, the Go package ``core/internal/embedgpu`` proposed in
INTEGRATION.md (source under ``llm-mcp_b200/go/embedgpu``; the image has no Go toolchain,
so the Python mirror is what the tests and the bench drive).  Its public surface follows the
reference call site:

    truncate_embeddings(result, dims)   <->  truncateEmbeddings(result, dims)
                                             core/internal/api/handlers.go:2062-2078
    resolve_dims(req_dimensions)        <->  handlers.go:1874-1877 + config.GetenvInt
                                             (core/internal/config/config.go:17-24)

The computation runs only on the GPU.  If the shared library is missing this module raises
on first use; there is no CPU path here (the CPU restatement lives in oracle/ and is test
infrastructure only).
"""
from __future__ import annotations

import ctypes
import os
from typing import Any

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libembed_truncate_norm.so")

F32, F64 = 0, 1
TRUNCATE, TRUNCATE_RENORM = 0, 1
ERR_NULL, ERR_SHAPE, ERR_ENUM, ERR_ALIGN, ERR_NOT_INIT, ERR_DEVICE, ERR_TOO_LARGE = -1, -2, -3, -4, -5, -6, -7
ERR_JSON, ERR_JSON_DUP, ERR_TICKET = -8, -9, -10
STATIC_STAGE_BYTES = 16 << 20

EXPORTS = (
    "etn_version", "etn_strerror", "embed_truncate_norm", "embed_truncate",
    "embed_truncate_norm_ragged", "etn_init", "etn_shutdown", "etn_device_count",
    "etn_alloc_pinned", "etn_free_pinned", "embed_truncate_norm_host", "etn_fill_synthetic",
    "embed_truncate_norm_host_static",
    "etn_json_truncate_response", "etn_json_parse_array", "etn_json_format_array", "etn_json_prompt_tokens",
    "etn_get_counters", "etn_counter_name",
    "embed_truncate_norm_gather", "embed_truncate_norm_ragged_host", "etn_shard_ragged", "etn_submit_host", "etn_wait",
    "etn_note_fallback", "etn_get_histogram", "etn_histogram_name", "etn_histogram_help", "etn_raw_copy_seconds",
    "etn_json_parse_tokens_device", "etn_json_format_numbers_device",
)


class EtnError(RuntimeError):
    def __init__(self, code: int, text: str, where: str):
        super().__init__(f"{where}: {text} (code {code})")
        self.code = code


_lib = None


def lib() -> ctypes.CDLL:
    """Loads libembed_truncate_norm.so (built by __graft_entry__.build / csrc/Makefile)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(
                f"{LIB_PATH} is missing: build it with `make -C llm-mcp_b200/csrc` "
                "(nvcc, sm_100a). There is no CPU fallback.")
        L = ctypes.CDLL(LIB_PATH)
        i64, vp, ci = ctypes.c_int64, ctypes.c_void_p, ctypes.c_int
        L.etn_version.restype = ci
        L.etn_strerror.argtypes = [ci]
        L.etn_strerror.restype = ctypes.c_char_p
        L.embed_truncate_norm.argtypes = [i64, i64, i64, ci, vp, vp, vp]
        L.embed_truncate.argtypes = [i64, i64, i64, ci, vp, vp, vp]
        L.embed_truncate_norm_ragged.argtypes = [i64, i64, vp, vp, ci, ci, vp, vp, vp]
        L.etn_init.argtypes = [ctypes.POINTER(ci), ci]
        L.etn_shutdown.restype = None
        L.etn_device_count.restype = ci
        L.etn_alloc_pinned.argtypes = [ctypes.c_size_t]
        L.etn_alloc_pinned.restype = vp
        L.etn_free_pinned.argtypes = [vp]
        L.etn_free_pinned.restype = None
        L.embed_truncate_norm_host.argtypes = [i64, i64, i64, ci, ci, vp, i64, vp, ci]
        L.etn_fill_synthetic.argtypes = [ctypes.c_uint64, i64, i64, i64, i64, i64, ci, vp, vp]
        L.embed_truncate_norm_host_static.argtypes = [i64, i64, i64, ci, ci, vp, i64, vp, vp]
        L.etn_json_truncate_response.argtypes = [ctypes.c_char_p, ctypes.c_size_t, i64, ci, vp, ctypes.c_size_t, ci, vp]
        L.etn_json_truncate_response.restype = i64
        L.etn_json_parse_array.argtypes = [ctypes.c_char_p, ctypes.c_size_t, vp, i64]
        L.etn_json_parse_array.restype = i64
        L.etn_json_format_array.argtypes = [vp, i64, vp, ctypes.c_size_t]
        L.etn_json_format_array.restype = i64
        L.etn_json_prompt_tokens.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
        L.etn_json_prompt_tokens.restype = i64
        L.etn_get_counters.argtypes = [vp, ci]
        L.etn_counter_name.argtypes = [ci]
        L.etn_counter_name.restype = ctypes.c_char_p
        L.embed_truncate_norm_gather.argtypes = [i64, vp, vp, vp, ci, ci, vp, vp, vp]
        L.embed_truncate_norm_ragged_host.argtypes = [i64, i64, vp, ci, ci, vp, i64, vp, ci]
        L.etn_shard_ragged.argtypes = [i64, i64, vp, ci, vp]
        L.etn_submit_host.argtypes = [i64, i64, i64, vp, ci, ci, vp, i64, vp, ci, ctypes.POINTER(i64)]
        L.etn_wait.argtypes = [i64]
        L.etn_note_fallback.argtypes = [i64]
        L.etn_note_fallback.restype = None
        L.etn_get_histogram.argtypes = [ci, vp, vp, ci, vp]
        L.etn_histogram_name.argtypes = [ci]
        L.etn_histogram_name.restype = ctypes.c_char_p
        L.etn_histogram_help.argtypes = [ci]
        L.etn_histogram_help.restype = ctypes.c_char_p
        L.etn_raw_copy_seconds.argtypes = [vp, vp, ctypes.c_size_t, ci, ci, vp]
        L.etn_json_parse_tokens_device.argtypes = [vp, vp, vp, i64, vp, vp, vp]
        L.etn_json_format_numbers_device.argtypes = [vp, i64, vp, vp, vp]
        _lib = L
    return _lib


def strerror(code: int) -> str:
    return lib().etn_strerror(code).decode()


def check(code: int, where: str) -> None:
    if code != 0:
        raise EtnError(code, strerror(code), where)


def dtype_code(dt) -> int:
    dt = np.dtype(dt)
    if dt == np.float32:
        return F32
    if dt == np.float64:
        return F64
    raise TypeError(f"unsupported element type {dt} (float32 or float64)")


# --------------------------------------------------------------------------- init ----
def init(device_ids: list[int] | None = None) -> int:
    """etn_init; returns the number of bound devices."""
    if device_ids:
        arr = (ctypes.c_int * len(device_ids))(*device_ids)
        check(lib().etn_init(arr, len(device_ids)), "etn_init")
    else:
        check(lib().etn_init(None, 0), "etn_init")
    return lib().etn_device_count()


def shutdown() -> None:
    lib().etn_shutdown()


# ------------------------------------------------------------------ pinned buffers ----
class PinnedBuffer:
    """Page-locked host memory from etn_alloc_pinned, viewed as a numpy array."""

    def __init__(self, nbytes: int):
        self.nbytes = int(nbytes)
        self.ptr = lib().etn_alloc_pinned(self.nbytes) if self.nbytes else None
        if self.nbytes and not self.ptr:
            raise MemoryError(f"etn_alloc_pinned({self.nbytes}) failed")

    def array(self, shape, dtype) -> np.ndarray:
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * dtype.itemsize
        if n > self.nbytes:
            raise ValueError("view larger than the buffer")
        if n == 0:
            return np.empty(shape, dtype=dtype)
        buf = (ctypes.c_char * n).from_address(self.ptr)
        return np.frombuffer(buf, dtype=dtype).reshape(shape)

    def free(self) -> None:
        if self.ptr:
            lib().etn_free_pinned(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


# ------------------------------------------------------------------- host entry ----
def truncate_norm_host(x: np.ndarray, k: int, mode: int = TRUNCATE_RENORM, out: np.ndarray | None = None,
                       n_devices: int = 0) -> np.ndarray:
    """embed_truncate_norm_host on a C-contiguous-row [batch x D] host array (pinned or not)."""
    if x.ndim != 2 or x.strides[1] != x.itemsize:
        raise ValueError("x must be 2-D with unit column stride")
    batch, D = x.shape
    kk = min(int(k), D) if k > 0 else 0
    if out is None:
        out = np.empty((batch, max(kk, 0)), dtype=x.dtype)
    pitch = x.strides[0] if batch > 1 else D * x.itemsize
    check(lib().embed_truncate_norm_host(batch, D, k, dtype_code(x.dtype), mode,
                                         x.ctypes.data if batch else None, pitch,
                                         out.ctypes.data if batch else None, n_devices),
          "embed_truncate_norm_host")
    return out


def truncate_norm_ragged_host(x: np.ndarray, k_per_row, mode: int = TRUNCATE_RENORM, out: np.ndarray | None = None,
                              n_devices: int = 0) -> np.ndarray:
    """embed_truncate_norm_ragged_host: row r keeps min(k_per_row[r], D) components; the result is
    the rows packed back to back (1-D).  One call for responses of requests that resolved
    different `dimensions` (handlers.go:1874-1877)."""
    if x.ndim != 2 or x.strides[1] != x.itemsize:
        raise ValueError("x must be 2-D with unit column stride")
    batch, D = x.shape
    kpr = np.ascontiguousarray(k_per_row, dtype=np.int32)
    if kpr.shape != (batch,):
        raise ValueError("k_per_row must have one entry per row")
    total = int(np.clip(kpr.astype(np.int64), 0, D).sum())
    if out is None:
        out = np.empty(total, dtype=x.dtype)
    pitch = x.strides[0] if batch > 1 else D * x.itemsize
    check(lib().embed_truncate_norm_ragged_host(batch, D, kpr.ctypes.data if batch else None, dtype_code(x.dtype), mode,
                                                x.ctypes.data if batch else None, pitch,
                                                out.ctypes.data if total else (x.ctypes.data if batch else None), n_devices),
          "embed_truncate_norm_ragged_host")
    return out


class Ticket:
    """etn_submit_host / etn_wait: a host call in flight.  Keeps its buffers alive until waited for."""

    def __init__(self, x: np.ndarray, k: int, mode: int = TRUNCATE_RENORM, out: np.ndarray | None = None, n_devices: int = 0):
        if x.ndim != 2 or x.strides[1] != x.itemsize:
            raise ValueError("x must be 2-D with unit column stride")
        batch, D = x.shape
        self.x = x
        self.out = np.empty((batch, min(int(k), D)), dtype=x.dtype) if out is None else out
        t = ctypes.c_int64(0)
        pitch = x.strides[0] if batch > 1 else D * x.itemsize
        check(lib().etn_submit_host(batch, D, k, None, dtype_code(x.dtype), mode, x.ctypes.data if batch else None, pitch,
                                    self.out.ctypes.data if batch else None, n_devices, ctypes.byref(t)), "etn_submit_host")
        self.ticket = t.value

    def wait(self) -> np.ndarray:
        check(lib().etn_wait(self.ticket), "etn_wait")
        return self.out


# --------------------------------------------------------- reference call-site mirror ----
def getenv_int(key: str, default: int, env=None) -> int:
    """config.GetenvInt (config.go:17-24): unset, empty or not an integer -> default."""
    env = os.environ if env is None else env
    v = env.get(key, "")
    if v == "":
        return default
    body = v[1:] if v[:1] in "+-" else v
    if not body or not body.isascii() or not body.isdigit():
        return default
    n = int(v)
    return n if -(1 << 63) <= n < (1 << 63) else default


def resolve_dims(req_dimensions: int, env=None) -> int:
    """handlers.go:1874-1877: the request's `dimensions`, else CLOUD_EMBED_DIMENSIONS."""
    return req_dimensions if req_dimensions != 0 else getenv_int("CLOUD_EMBED_DIMENSIONS", 0, env)


def renorm_enabled(env=None) -> bool:
    """EMBED_RENORMALIZE=1 switches the handler from the reference's truncate-only
    behaviour to truncate+renormalise.  Default off: renormalising changes response values
    relative to the reference."""
    return getenv_int("EMBED_RENORMALIZE", 0, env) == 1


def _packable(emb: list) -> bool:
    return all(type(v) in (float, int) for v in emb)


def truncate_embeddings(result: Any, dims: int, renorm: bool | None = None) -> dict:
    """GPU drop-in for ``truncateEmbeddings(result, dims)`` (handlers.go:2062-2078).

    Same contract as the reference, item by item: ``data`` that is not a list is a no-op;
    non-dict items are skipped; an ``embedding`` that is not a list (base64 string, null) is
    left alone; vectors with ``len <= dims`` are left whole; longer ones keep their first
    ``dims`` components in order.  With ``renorm`` (default: EMBED_RENORMALIZE) the kept
    prefix is L2-renormalised as well.  Vectors are packed per source length into pinned
    buffers and each group goes through one ``embed_truncate_norm_host`` call; a vector
    holding anything but JSON numbers cannot be packed and gets the reference reslice.
    Without ``renorm`` nothing is computed, so — exactly like the Go binding
    (go/embedgpu/embedgpu.go) — every vector gets the reference's O(1) reslice and the GPU is
    not involved (no PCIe round trip to take a slice; integer elements stay integers).
    Returns counters {"gpu_vectors", "resliced_vectors", "launch_groups"}.
    """
    stats = {"gpu_vectors": 0, "resliced_vectors": 0, "launch_groups": 0}
    if renorm is None:
        renorm = renorm_enabled()
    if not isinstance(result, dict) or dims <= 0:
        return stats
    data = result.get("data")
    if not isinstance(data, list):
        return stats
    groups: dict[int, list[dict]] = {}
    for item in data:
        if not isinstance(item, dict):
            continue
        emb = item.get("embedding")
        if not isinstance(emb, list) or len(emb) <= dims:
            continue
        if renorm and _packable(emb):
            groups.setdefault(len(emb), []).append(item)
        else:
            item["embedding"] = emb[:dims]
            stats["resliced_vectors"] += 1
    if renorm and stats["resliced_vectors"]:
        lib().etn_note_fallback(stats["resliced_vectors"])      # renorm was asked for and could not be done
    mode = TRUNCATE_RENORM
    for D, items in groups.items():
        n = len(items)
        pin_in, pin_out = PinnedBuffer(n * D * 8), PinnedBuffer(n * dims * 8)
        try:
            x = pin_in.array((n, D), np.float64)
            for r, item in enumerate(items):
                x[r] = item["embedding"]
            y = truncate_norm_host(x, dims, mode, out=pin_out.array((n, dims), np.float64))
            for r, item in enumerate(items):
                item["embedding"] = y[r].tolist()
        finally:
            pin_in.free()
            pin_out.free()
        stats["gpu_vectors"] += n
        stats["launch_groups"] += 1
    return stats


# ----------------------------------------------------------------- row sharding ----
def shard_rows(batch: int, world: int) -> list[tuple[int, int]]:
    """Contiguous row blocks, the split embed_truncate_norm_host uses: shard g gets rows
    [g*ceil(B/G), min(B, (g+1)*ceil(B/G))).  Rows are independent (each vector is handled on
    its own at handlers.go:2068-2076), so shards never exchange data."""
    if world < 1 or batch < 0:
        raise ValueError("world >= 1 and batch >= 0")
    per = -(-batch // world) if batch else 0
    return [(min(batch, g * per), min(batch, (g + 1) * per)) for g in range(world)]


def shard_ragged(k_per_row, D: int, world: int) -> list[tuple[int, int]]:
    """Contiguous row ranges of a ragged batch balanced by bytes moved (sum of min(k, D)),
    not by row count: the cost of a row is proportional to its k.  The cut
    points are etn_shard_ragged's, the ones embed_truncate_norm_ragged_host uses itself:
    shard g starts at the first row r with (kept elements before r) * world >= total * g."""
    if world < 1:
        raise ValueError("world >= 1")
    kpr = np.ascontiguousarray(k_per_row, dtype=np.int32)
    cuts = np.zeros(world + 1, dtype=np.int64)
    check(lib().etn_shard_ragged(len(kpr), D, kpr.ctypes.data if len(kpr) else None, world, cuts.ctypes.data), "etn_shard_ragged")
    return [(int(cuts[g]), int(cuts[g + 1])) for g in range(world)]


# ------------------------------------------------------------------ JSON packer (N1) ----
JSON_BASE64 = 0x10


def truncate_response_json(body: bytes, dims: int, renorm: bool | None = None, n_devices: int = 0, base64_vectors: bool = False):
    """etn_json_truncate_response: the provider's response text in, the handler's response
    text out, with the embedding arrays cut (and, with renorm, renormalised on the GPU)
    without ever building the decoded map (handlers.go:2047-2050 / helpers.go:11-15).
    Returns (bytes, stats).  Raises EtnError with ERR_JSON / ERR_JSON_DUP when the caller
    should take the reference path instead."""
    if renorm is None:
        renorm = renorm_enabled()
    mode = (TRUNCATE_RENORM if renorm else TRUNCATE) | (JSON_BASE64 if base64_vectors else 0)
    cap = len(body) + 2 if not renorm else len(body) + 2 + 26 * max(dims, 0) * (body.count(b'"embedding"') + 1)
    out = ctypes.create_string_buffer(cap)
    stats = (ctypes.c_int64 * 3)()
    n = lib().etn_json_truncate_response(body, len(body), dims, mode, out, cap, n_devices, stats)
    if n < 0:
        code = int(n) if n > -1000 else int(-1000 - n)
        raise EtnError(code, strerror(code), "etn_json_truncate_response")
    return out.raw[:n], {"gpu_vectors": stats[0], "resliced_vectors": stats[1], "arrays": stats[2]}


def prompt_tokens(body: bytes) -> int:
    """usage.prompt_tokens the way proxyOpenAIEmbed reads it (handlers.go:2052-2057)."""
    n = lib().etn_json_prompt_tokens(body, len(body))
    if n < 0 and n in (ERR_JSON, ERR_NULL):
        raise EtnError(int(n), strerror(int(n)), "etn_json_prompt_tokens")
    return int(n)


# ------------------------------------------------------------------- counters (N3) ----
def counters() -> dict[str, int]:
    """etn_get_counters as {metric name: value} — what the Go side would register beside the
    embedding collectors of core/internal/metrics/metrics.go:9-26."""
    buf = (ctypes.c_int64 * 32)()
    n = lib().etn_get_counters(buf, 32)
    return {lib().etn_counter_name(i).decode(): int(buf[i]) for i in range(n)}


HIST_BATCH_ROWS, HIST_KERNEL_SECONDS, HIST_H2D_SECONDS, HIST_D2H_SECONDS, HIST_CALL_SECONDS, HIST_REQS_PER_LAUNCH = range(6)
HIST_COUNT = 6


def histogram(which: int) -> dict:
    """etn_get_histogram as {"name", "help", "le": [...], "counts": [...per bucket...], "sum", "count"}."""
    L = lib()
    n = L.etn_get_histogram(which, None, None, 0, None)
    if n < 0:
        raise EtnError(n, strerror(n), "etn_get_histogram")
    le, cnt, total = (ctypes.c_double * n)(), (ctypes.c_int64 * n)(), ctypes.c_double(0)
    check(min(0, L.etn_get_histogram(which, le, cnt, n, ctypes.byref(total))), "etn_get_histogram")
    counts = [int(c) for c in cnt]
    return {"name": L.etn_histogram_name(which).decode(), "help": L.etn_histogram_help(which).decode(),
            "le": [float(v) for v in le], "counts": counts, "sum": float(total.value), "count": sum(counts)}


def _fmt_le(v: float) -> str:
    return "+Inf" if v == float("inf") else repr(float(v)) if v != int(v) else str(int(v))


def metrics_text() -> str:
    """Prometheus text exposition of the counters and histograms (the /metrics route,
    server.go:61): what promauto.NewCounter / NewHistogram collectors registered beside
    metrics.go:9-26 would write — cumulative `le` buckets, `_sum`, `_count`."""
    out = [f"# TYPE {k} counter\n{k} {v}\n" for k, v in counters().items()]
    for w in range(HIST_COUNT):
        h = histogram(w)
        out.append(f"# HELP {h['name']} {h['help']}\n# TYPE {h['name']} histogram\n")
        run = 0
        for le, c in zip(h["le"], h["counts"]):
            run += c
            out.append(f'{h["name"]}_bucket{{le="{_fmt_le(le)}"}} {run}\n')
        out.append(f"{h['name']}_sum {h['sum']!r}\n{h['name']}_count {h['count']}\n")
    return "".join(out)


# ------------------------------------------------------------ typed vectors (N2) ----
def truncate_typed(embeddings, dims: int, renorm: bool | None = None) -> list:
    """The op on the local Ollama branch: there the vectors arrive typed,
    ``Embeddings [][]float64`` (handlers.go:1976-1984), and ``dimensions`` is ignored today
    (handlers.go:1898-1939 never truncates).  Same rules as truncateEmbeddings: vectors with
    len <= dims stay whole, longer ones keep their first ``dims`` components; dims <= 0 is a
    no-op.  Typed rows pack with one copy, so every cut vector goes through the GPU host path
    in one call per source length.  Returns a new list; the input is not modified."""
    if renorm is None:
        renorm = renorm_enabled()
    out = [list(v) if not isinstance(v, list) else v for v in embeddings]
    if dims <= 0:
        return out
    if not renorm:                                   # a slice needs no GPU (as in the Go TruncateTyped)
        return [v[:dims] if len(v) > dims else v for v in out]
    groups: dict[int, list[int]] = {}
    for i, v in enumerate(out):
        if len(v) > dims:
            groups.setdefault(len(v), []).append(i)
    for D, idx in groups.items():
        x = np.array([out[i] for i in idx], dtype=np.float64)
        y = truncate_norm_host(x, dims, TRUNCATE_RENORM)
        for r, i in enumerate(idx):
            out[i] = y[r].tolist()
    return out


# ---------------------------------------------------- async job results (N4, worker side) ----
def truncate_job_results(results: list, dims: int, renorm: bool | None = None) -> dict:
    """Post-hoc cut of ``ollama.embed`` job results: the worker returns
    ``{"status", "ms", "data": {"embedding": [...]}}`` per job
    (worker/llm_worker/main.py:246-261, doc/README.md:302-307: ``result.data.embedding``), one
    vector each, so a batch of finished jobs is packed into one host-path call per source
    length.  Same rules as truncateEmbeddings (handlers.go:2068-2076): anything that is not a
    list of numbers longer than ``dims`` is left alone.  Mutates ``results`` in place."""
    stats = {"gpu_vectors": 0, "resliced_vectors": 0, "launch_groups": 0}
    if renorm is None:
        renorm = renorm_enabled()
    if dims <= 0:
        return stats
    groups: dict[int, list[dict]] = {}
    for res in results:
        data = res.get("data") if isinstance(res, dict) else None
        emb = data.get("embedding") if isinstance(data, dict) else None
        if not isinstance(emb, list) or len(emb) <= dims:
            continue
        if renorm and _packable(emb):
            groups.setdefault(len(emb), []).append(data)
        else:
            data["embedding"] = emb[:dims]
            stats["resliced_vectors"] += 1
    if renorm and stats["resliced_vectors"]:
        lib().etn_note_fallback(stats["resliced_vectors"])
    for D, items in groups.items():
        x = np.array([it["embedding"] for it in items], dtype=np.float64)
        y = truncate_norm_host(x, dims, TRUNCATE_RENORM)
        for r, it in enumerate(items):
            it["embedding"] = y[r].tolist()
        stats["gpu_vectors"] += len(items)
        stats["launch_groups"] += 1
    return stats
