"""CPU test of the per-number code of the JSON packer's GPU path (llm-mcp_b200/csrc/etn_numfmt.cuh):
the same host+device source is compiled for the CPU and checked against the library's host
formatter / parser (std::to_chars / std::from_chars: the guarantees of Go's strconv, which is
what encoding/json calls at handlers.go:2048 and helpers.go:14).  format: byte-identical for
every value; parse: bit-identical whenever the fast algorithm answers (it may say "ask the
host", it may never answer wrongly)."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_device_number_code_matches_host_formatter_and_parser(tmp_path):
    exe = tmp_path / "numfmt_check"
    libdir = os.path.join(ROOT, "llm-mcp_b200")
    build = subprocess.run(["g++", "-O2", "-std=c++17", "-o", str(exe), os.path.join(ROOT, "tests", "native", "numfmt_check.cpp"),
                            "-L" + libdir, "-lembed_truncate_norm", "-Wl,-rpath," + libdir], capture_output=True, text=True)
    assert build.returncode == 0, build.stderr[-2000:]
    run = subprocess.run([str(exe), os.environ.get("ETN_NUMFMT_MILLIONS", "3")], capture_output=True, text=True, timeout=280)
    assert run.returncode == 0, (run.stdout[-500:], run.stderr[-3000:])
    m = re.search(r"parse answered (\d+), fell back (\d+), failures 0", run.stdout)
    assert m, run.stdout
    answered, fell_back = int(m.group(1)), int(m.group(2))
    assert answered > 1_000_000 and fell_back < 0.01 * answered      # the fall-back is the rare case


def test_pow10_table_is_what_the_generator_writes(tmp_path):
    """The committed table equals tools/gen_pow10.py's output (exact integer arithmetic)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("gen_pow10", os.path.join(ROOT, "tools", "gen_pow10.py"))
    g = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(g)
    rows = [l for l in open(os.path.join(ROOT, "llm-mcp_b200", "csrc", "etn_pow10_table.inc")) if l.startswith("{")]
    assert len(rows) == g.QMAX - g.QMIN + 1 == 667
    for q, line in zip(range(g.QMIN, g.QMAX + 1), rows):
        hi, lo = (int(x, 16) for x in re.findall(r"0x([0-9A-F]{16})ULL", line))
        assert (hi << 64) | lo == g.entry(q) and hi >> 63 == 1, q
    # spot values: exact small powers, and 1/10 rounded up in the last place
    assert g.entry(0) == 1 << 127 and g.entry(1) == 0xA << 124 and g.entry(-1) == 0xCCCCCCCCCCCCCCCCCCCCCCCCCCCCCCCD
