"""Helpers for the -m gpu tests: torch is used for device memory and streams only; every
computation under test goes through the C-ABI of libembed_truncate_norm.so."""
import numpy as np
import torch

from llm_mcp_b200 import embedgpu as eg

TORCH_DT = {np.dtype(np.float32): torch.float32, np.dtype(np.float64): torch.float64}


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def to_dev(a: np.ndarray) -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def dev_fill(seed, rows, D, dtype, row0=0, ncols=None):
    """Synthetic matrix generated on the device by etn_fill_synthetic."""
    ncols = D if ncols is None else ncols
    t = torch.empty((rows, ncols), dtype=TORCH_DT[np.dtype(dtype)], device="cuda")
    eg.check(eg.lib().etn_fill_synthetic(seed, row0, rows, D, ncols, ncols, eg.dtype_code(dtype),
                                         t.data_ptr(), stream_ptr()), "etn_fill_synthetic")
    return t


def run_dense(x: torch.Tensor, k: int, mode: int, out: torch.Tensor | None = None) -> torch.Tensor:
    """embed_truncate_norm / embed_truncate on a contiguous [batch x D] device tensor."""
    batch, D = x.shape
    npdt = np.float32 if x.dtype == torch.float32 else np.float64
    if out is None:
        out = torch.empty((batch, min(k, D)), dtype=x.dtype, device=x.device)
    fn = eg.lib().embed_truncate_norm if mode == eg.TRUNCATE_RENORM else eg.lib().embed_truncate
    rc = fn(batch, D, k, eg.dtype_code(npdt), x.data_ptr() if batch else None,
            out.data_ptr() if batch else None, stream_ptr())
    eg.check(rc, "embed_truncate_norm" if mode else "embed_truncate")
    return out


def run_ragged(x: torch.Tensor, kpr: np.ndarray, off: np.ndarray, mode: int) -> torch.Tensor:
    batch, D = x.shape
    npdt = np.float32 if x.dtype == torch.float32 else np.float64
    out = torch.empty(int(off[-1]), dtype=x.dtype, device=x.device)
    dk = to_dev(np.ascontiguousarray(kpr, dtype=np.int32))
    do = to_dev(np.ascontiguousarray(off[:-1], dtype=np.int64))
    rc = eg.lib().embed_truncate_norm_ragged(batch, D, dk.data_ptr(), do.data_ptr(), eg.dtype_code(npdt),
                                             mode, x.data_ptr(), out.data_ptr(), stream_ptr())
    eg.check(rc, "embed_truncate_norm_ragged")
    torch.cuda.synchronize()
    return out


def bits(a: np.ndarray) -> np.ndarray:
    a = np.ascontiguousarray(a)
    return a.view(np.uint32 if a.dtype == np.float32 else np.uint64)
