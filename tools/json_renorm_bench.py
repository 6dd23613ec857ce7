"""C1 response text -> response text with renormalisation (etn_json_truncate_response), number
work on the GPU (default) and on the host cores (ETN_JSON_GPU=0), same process."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from llm_mcp_b200 import embedgpu as eg
for flag in ("1", "0", "1"):
    os.environ["ETN_JSON_GPU"] = flag
    r = bench.json_path_measure(eg)
    print("ETN_JSON_GPU=%s  truncate %.2f ms   truncate+renorm %.2f ms   (%s)" % (flag, r["truncate"]["ms_per_response"], r["truncate_renorm"]["ms_per_response"], r["workload"]), flush=True)
