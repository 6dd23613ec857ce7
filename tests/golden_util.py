"""Loaders for tests/golden/ (authored by tests/golden/make_golden.py)."""
import json
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NP = {"f32": np.float32, "f64": np.float64}


def kat_cases():
    with open(os.path.join(GOLDEN, "kat.json")) as f:
        cases = json.load(f)["cases"]
    out = []
    for c in cases:
        dt = NP[c["dtype"]]
        conv = lambda rows: [np.array([float.fromhex(h) for h in r], dtype=dt) for r in rows]
        out.append({"name": c["name"], "dtype": dt, "D": c["D"], "k": c["k"],
                    "in": np.stack(conv(c["in"])), "truncate": conv(c["truncate"]),
                    "renorm": np.stack(conv(c["renorm"]))})
    return out


def synthetic():
    z = np.load(os.path.join(GOLDEN, "synthetic.npz"))
    manifest = json.loads(str(z["manifest"]))
    return z, manifest
