#!/usr/bin/env python3
"""Authors the golden vectors under tests/golden/.

The reference (plagness/LLM-MCP) has no test, fixture or golden vector for the
/v1/embeddings Matryoshka path and its Go code cannot be run here (F4), so
these vectors are authored from the semantics of core/internal/api/handlers.go:2062-2078
(truncation) and from BASELINE.json's definition of the renormalise step.  They are
computed with Python's ``decimal`` (120 significant digits) and ``fractions`` only —
no numpy arithmetic, no C oracle, no CUDA — so they pin oracle/etn_oracle.c and the
kernels independently.

    python tests/golden/make_golden.py        # rewrites kat.json and synthetic.npz

Outputs are correctly rounded (round-to-nearest-even) to the element type.
"""
from __future__ import annotations

import json
import os
import struct
from decimal import Decimal, getcontext
from fractions import Fraction

import numpy as np  # container + float32 neighbours only

getcontext().prec = 120
HERE = os.path.dirname(os.path.abspath(__file__))
M64 = (1 << 64) - 1


# ----------------------------------------------------------------- rounding ----
def f32_bits(x: float) -> int:
    return struct.unpack("<I", struct.pack("<f", x))[0]


def f32_from_bits(b: int) -> float:
    return struct.unpack("<f", struct.pack("<I", b))[0]


def round_f64(q: Fraction) -> float:
    # int/int true division in CPython is correctly rounded (round-half-even),
    # including subnormal results.
    return q.numerator / q.denominator


def round_f32(q: Fraction) -> float:
    """Nearest float32 to the exact rational q (ties to even)."""
    if q == 0:
        return 0.0
    sign = -1 if q < 0 else 1
    a = abs(q)
    guess = float(np.float32(a.numerator / a.denominator))  # may be double-rounded: fix below
    b = f32_bits(guess)
    best = None
    for cand_bits in (b - 1, b, b + 1):
        if cand_bits < 0 or cand_bits >= 0x7F800000:
            continue
        c = f32_from_bits(cand_bits)
        err = abs(Fraction(c) - a)
        key = (err, cand_bits & 1)  # smaller error first, then even mantissa
        if best is None or key < best[0]:
            best = (key, c)
    return sign * best[1]


def dec(q: Fraction) -> Decimal:
    return Decimal(q.numerator) / Decimal(q.denominator)


def frac(d: Decimal) -> Fraction:
    return Fraction(d)


# -------------------------------------------------------------- definitions ----
def renorm_row(xs: list[float], k: int, dtype: str) -> list[float]:
    """L2-renormalise the first k components; an all-zero prefix is returned unchanged."""
    pre = xs[:k]
    ss = sum(Fraction(v) * Fraction(v) for v in pre)
    if ss == 0:
        return list(pre)
    norm = dec(ss).sqrt()
    rnd = round_f32 if dtype == "f32" else round_f64
    out = []
    for v in pre:
        if v == 0:
            out.append(v)  # keeps the sign of zero
        else:
            out.append(rnd(frac(Decimal(Fraction(v).numerator) / Decimal(Fraction(v).denominator) / norm)))
    return out


def truncate_row(xs: list[float], k: int) -> list[float]:
    """handlers.go:2074-2075 — emb[:dims] only when len(emb) > dims."""
    return list(xs[:k]) if len(xs) > k else list(xs)


# ---------------------------------------------------------------- generator ----
def splitmix64(z: int) -> int:
    z = (z + 0x9E3779B97F4A7C15) & M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
    return z ^ (z >> 31)


def synth_int(seed: int, row: int, D: int, col: int) -> int:
    h = splitmix64((seed ^ (row * D + col)) & M64)
    v = h >> 32
    v = v - (1 << 32) if v >= (1 << 31) else v
    return v >> 8  # arithmetic shift: signed 24-bit, exact in float32


def synth(seed: int, row: int, D: int, col: int) -> float:
    return synth_int(seed, row, D, col) * 2.0 ** -28  # exact in f32 and f64


# --------------------------------------------------------------------- KATs ----
def hexes(vs):
    return [float(v).hex() for v in vs]


def kat_cases():
    f32 = lambda v: float(np.float32(v))
    cases = []

    def add(name, dtype, rows, k):
        if dtype == "f32":
            rows = [[f32(v) for v in r] for r in rows]
        D = len(rows[0])
        kk = min(k, D)
        cases.append({
            "name": name, "dtype": dtype, "D": D, "k": k,
            "in": [hexes(r) for r in rows],
            "truncate": [hexes(truncate_row(r, k)) for r in rows],
            "renorm": [hexes(renorm_row(r, kk, dtype)) for r in rows],
        })

    for dt in ("f64", "f32"):
        add(f"pythagorean_k2_{dt}", dt, [[3.0, 4.0, 0.0, 0.0], [-3.0, 4.0, 12.0, 1.0]], 2)
        add(f"pythagorean_k4_{dt}", dt, [[3.0, 4.0, 0.0, 0.0], [-3.0, 4.0, 12.0, 84.0]], 4)
        add(f"k_ge_D_{dt}", dt, [[1.0, 2.0, 2.0]], 7)
        add(f"zero_prefix_{dt}", dt, [[0.0, -0.0, 5.0, 1.0], [0.0, 0.0, 0.0, 0.0]], 2)
        add(f"single_nonzero_{dt}", dt, [[0.0, -7.25, 0.0, 9.0], [1e-3, 0.0, 0.0, 2.0]], 3)
        add(f"k1_{dt}", dt, [[-0.125, 3.0], [5.5, -1.0]], 1)
        add(f"unit_idempotent_{dt}", dt, [[0.6, 0.8, 0.0], [0.28, -0.96, 1.0]], 2)
        add(f"mixed_signs_{dt}", dt, [[0.1, -0.2, 0.3, -0.4, 0.5, -0.6, 0.7]], 5)
    # float64 range: squares overflow above ~1.3e154 and underflow below ~1.5e-162
    add("huge_f64", "f64", [[1e200, -3e200, 2e199, 1.0], [1.7e308, 1.7e308, 1e308, 0.0]], 3)
    add("tiny_f64", "f64", [[1e-200, -3e-200, 2e-199, 1.0], [5e-324, 1e-323, -5e-324, 0.0]], 3)
    add("subnormal_mix_f64", "f64", [[2.2e-308, 1e-310, -3e-309, 4.0], [1e-170, 1e-165, -1e-160, 0.0]], 3)
    add("wide_dynamic_f64", "f64", [[1e150, 1.0, 1e-150, 3.0], [1e-150, 1e-151, 2e-150, 0.0]], 3)
    # float32 range: squares overflow above ~1.8e19, underflow below ~1e-23 in f32 arithmetic
    add("huge_f32", "f32", [[3e38, -1e38, 2e37, 1.0], [3.4e38, 3.4e38, 3.4e38, 0.0]], 3)
    add("tiny_f32", "f32", [[1e-30, -3e-30, 2e-31, 1.0], [1e-45, 3e-45, -1e-45, 0.0]], 3)
    add("subnormal_mix_f32", "f32", [[1e-38, 1e-40, -3e-39, 4.0], [1e-25, 1e-22, -1e-20, 0.0]], 3)
    return cases


# ---------------------------------------------------------------- synthetic ----
SYNTH = [
    # name, seed, rows, D, k, dtype   (seeds as in SURVEY.md §8(d): C1=1 … C5=5)
    ("c1_qwen3_f64", 1, 4, 4096, 1024, "f64"),
    ("c2_nomic_f32", 2, 8, 768, 256, "f32"),
    ("c3_qwen3_f32", 3, 4, 4096, 1024, "f32"),
    ("c4_te3large_f32", 4, 8, 3072, 256, "f32"),
    ("odd_D770_k257_f32", 7, 5, 770, 257, "f32"),
    ("odd_D771_k100_f64", 8, 5, 771, 100, "f64"),
    ("odd_D33_k32_f32", 9, 6, 33, 32, "f32"),
    ("odd_D40_k40_f64", 10, 3, 40, 40, "f64"),
    ("odd_D40_k39_f64", 11, 3, 40, 39, "f64"),
    ("odd_D4096_k1_f32", 12, 4, 4096, 1, "f32"),
    ("big_k_D8192_k6000_f64", 13, 2, 8192, 6000, "f64"),
    ("big_k_D12288_k9000_f32", 14, 2, 12288, 9000, "f32"),
]
RAGGED = ("c5_ragged_f64", 5, 16, 4096, [64, 128, 256, 512, 1024], "f64")


def main():
    with open(os.path.join(HERE, "kat.json"), "w") as f:
        json.dump({"comment": "authored by make_golden.py; values are C99 hex floats",
                   "cases": kat_cases()}, f, indent=0)

    arrays, manifest = {}, []
    for name, seed, rows, D, k, dt in SYNTH:
        npdt = np.float32 if dt == "f32" else np.float64
        out = np.empty((rows, k), dtype=npdt)
        for r in range(rows):
            xs = [synth(seed, r, D, c) for c in range(k)]
            out[r] = renorm_row(xs, k, dt)
        arrays[name] = out
        manifest.append({"name": name, "seed": seed, "rows": rows, "D": D, "k": k, "dtype": dt})
    name, seed, rows, D, ks, dt = RAGGED
    kpr = [ks[splitmix64(seed * 1000 + r) % len(ks)] for r in range(rows)]
    packed = []
    for r in range(rows):
        xs = [synth(seed, r, D, c) for c in range(kpr[r])]
        packed.extend(renorm_row(xs, kpr[r], dt))
    arrays[name] = np.array(packed, dtype=np.float64)
    arrays[name + "_k"] = np.array(kpr, dtype=np.int32)
    manifest.append({"name": name, "seed": seed, "rows": rows, "D": D, "k": "ragged", "dtype": dt})
    # raw generator outputs, as signed 24-bit integers before scaling (pins splitmix64 and the index formula)
    pins = [(s, r, D, c) for s in (1, 2, 5) for (r, D, c) in
            ((0, 4096, 0), (0, 4096, 1), (3, 768, 255), (1048575, 4096, 1023), (8388607, 3072, 255))]
    arrays["generator_pins"] = np.array([[s, r, D, c, synth_int(s, r, D, c)] for s, r, D, c in pins],
                                        dtype=np.int64)
    arrays["manifest"] = np.array(json.dumps(manifest))
    np.savez_compressed(os.path.join(HERE, "synthetic.npz"), **arrays)
    print("wrote kat.json (%d cases) and synthetic.npz (%d arrays)" % (len(kat_cases()), len(arrays)))


if __name__ == "__main__":
    main()
