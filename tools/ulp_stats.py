"""Distribution of the ulp distance between the CUDA path and the exact oracle on sampled rows
of the BASELINE shapes (written to profiles/ as evidence for the <= 2 ulp claim)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from llm_mcp_b200 import embedgpu as eg
from oracle import etn_oracle as oracle
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import tests.gpu_util as gu

torch.cuda.set_device(0)
for name, rows, D, k, dt, seed in (("C1 f64", 1024, 4096, 1024, np.float64, 1), ("C2 f32", 65536, 768, 256, np.float32, 2),
                                   ("C3 f32 (64Ki rows)", 65536, 4096, 1024, np.float32, 3), ("C3 f64 (32Ki rows)", 32768, 4096, 1024, np.float64, 3),
                                   ("C4 f32 (64Ki rows)", 65536, 3072, 256, np.float32, 4), ("k=64 f64", 65536, 4096, 64, np.float64, 5),
                                   ("random mantissas f64", 16384, 1024, 1024, np.float64, -1), ("random mantissas f32", 16384, 1024, 1024, np.float32, -1)):
    if seed >= 0:
        xh = oracle.fill(seed, rows, D, dt, ncols=k)
        x = gu.to_dev(xh)
        D = k
    else:
        xh = (np.random.default_rng(9).standard_normal((rows, D)) * np.exp2(np.random.default_rng(8).integers(-60, 60, (rows, 1)))).astype(dt)
        x = gu.to_dev(xh)
    got = gu.run_dense(x, k, eg.TRUNCATE_RENORM).cpu().numpy()
    want = oracle.renorm_exact(xh, k)
    iv = np.int32 if dt == np.float32 else np.int64
    d = np.abs(got.view(iv).astype(np.int64) - want.view(iv).astype(np.int64))
    hist = np.bincount(np.minimum(d.ravel(), 5), minlength=6)
    print("%-24s %10d elements  ulp distance 0: %.4f%%  1: %.4f%%  2: %.6f%%  >2: %d   max %d" %
          (name, d.size, 100 * hist[0] / d.size, 100 * hist[1] / d.size, 100 * hist[2] / d.size, hist[3:].sum(), d.max()), flush=True)
