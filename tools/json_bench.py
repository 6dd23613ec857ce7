"""N1 timing: C1-sized provider response (1,024 vectors x 4096 float64 as JSON text) through
the packer vs. a decode-to-objects / re-encode round trip (Python's C json here; Go's
encoding/json is the real comparison and is not available in this image)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from llm_mcp_b200 import embedgpu as eg
from oracle import etn_oracle as oracle

n, D, k = int(os.environ.get("N", 1024)), 4096, 1024
x = oracle.fill(1, n, D, np.float64)
body = json.dumps({"object": "list", "data": [{"object": "embedding", "index": i, "embedding": r.tolist()} for i, r in enumerate(x)],
                   "usage": {"prompt_tokens": 9}}).encode()
print("response: %d vectors, %.1f MB of JSON" % (n, len(body) / 1e6))
t0 = time.perf_counter(); d = json.loads(body); t1 = time.perf_counter()
oracle.truncate_embeddings_ref(d, k); t2 = time.perf_counter()
s = json.dumps(d); t3 = time.perf_counter()
print("decode %.0f ms + reslice %.3f ms + encode %.0f ms = %.0f ms (map round trip, Python C json)" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t3 - t0) * 1e3))
for rep in range(2):
    t0 = time.perf_counter(); out, st = eg.truncate_response_json(body, k, renorm=False); t1 = time.perf_counter()
print("packer TRUNCATE: %.0f ms (%.0f MB/s of input text), %s" % ((t1 - t0) * 1e3, len(body) / (t1 - t0) / 1e6, st))
assert json.loads(out) == d
if len(sys.argv) > 1 and sys.argv[1] == "gpu":
    eg.init()
    for rep in range(3):
        t0 = time.perf_counter(); out, st = eg.truncate_response_json(body, k, renorm=True); t1 = time.perf_counter()
    print("packer TRUNCATE_RENORM: %.0f ms, %s" % ((t1 - t0) * 1e3, st))
    eg.shutdown()
