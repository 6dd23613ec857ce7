"""Writes profiles/r02_sass_evidence.txt from the built library: python tools/sass_evidence.py
(cuobjdump -sass; counts mnemonics, checks that the in-place instantiations carry no .CONSTANT loads)."""
import collections, hashlib, os, re, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "llm-mcp_b200", "libembed_truncate_norm.so")
rev = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
out = []
out.append("# SASS evidence, round-2 build: cuobjdump -sass llm-mcp_b200/libembed_truncate_norm.so (sm_100a), source %s, library sha256 %s" % (rev, hashlib.sha256(open(so, "rb").read()).hexdigest()[:16]))
cur = None
per = collections.defaultdict(collections.Counter)
allc = collections.Counter()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        continue
    m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", line)
    if m and cur:
        per[cur][m.group(1)] += 1
        allc[m.group(1)] += 1
alias = [k for k in per if "rows_reg" in k and k.endswith("Li1EEEvPKcPclli")]
plain = [k for k in per if "rows_reg" in k and k.endswith("Li0EEEvPKcPclli")]
tot = lambda keys, pred: sum(v for k in keys for op, v in per[k].items() if pred(op))
out += ["#",
        "# 1. In-place instantiations (out == in, template ALIAS = 1) load through the coherent path: no ld.global.nc",
        "#    rows_reg_kernel<..., ALIAS=1>: %d instantiations, %d LDG, of which .CONSTANT: %d" % (len(alias), tot(alias, lambda o: o.startswith("LDG")), tot(alias, lambda o: o.startswith("LDG") and "CONSTANT" in o)),
        "#    rows_reg_kernel<..., ALIAS=0>: %d instantiations, %d LDG, of which .CONSTANT: %d" % (len(plain), tot(plain, lambda o: o.startswith("LDG")), tot(plain, lambda o: o.startswith("LDG") and "CONSTANT" in o)),
        "#    (the host path no longer runs anything in place: staging in and out are separate buffers; the only in-place",
        "#     callers are embed_truncate_norm(.., in == out) with k >= D and embed_truncate_norm_host_static)",
        "#",
        "# 2. Mnemonics over all kernels: 256-bit global accesses with L2 evict-first (sm_100 only), bulk-async copy (TMA",
        "#    engine, UBLKCP) with mbarrier transaction counts (SYNCS.*), programmatic dependent launch (PREEXIT/ACQBULK),",
        "#    64x64->128 integer multiplies of the JSON number kernels (IMAD.WIDE.U32 chains)"]
for op, c in allc.most_common():
    if any(t in op for t in ("LDG", "STG", "UBLKCP", "SYNCS", "PREEXIT", "ACQBULK", "LDS", "STS", "DFMA", "DADD", "DMUL", "MUFU", "SHFL", "IMAD.WIDE", "IMAD.HI", "BAR", "F2F", "ATOM")):
        out.append("%8d %s" % (c, op))
out += ["#", "# 3. Kernels in the library"]
for k in sorted(per):
    d = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip()
    out.append("#   %s" % d[:150])
open(os.path.join(ROOT, "profiles", "r02_sass_evidence.txt"), "w").write("\n".join(out) + "\n")
print(out[0]); print(out[3]); print(out[4])
