"""-m gpu: the C-ABI driven from plain C (gcc, only include/embed_truncate_norm.h) — the
view the cgo package core/internal/embedgpu would have; no Python, torch or CUDA headers on
the calling side."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_plain_c_caller(tmp_path):
    exe = tmp_path / "abi_c_caller"
    libdir = os.path.join(ROOT, "llm-mcp_b200")
    build = subprocess.run(["/usr/bin/gcc", "-std=c11", "-O1", "-Wall", "-Wextra", "-o", str(exe),
                            os.path.join(ROOT, "tests", "native", "abi_c_caller.c"),
                            "-L" + libdir, "-lembed_truncate_norm", "-Wl,-rpath," + libdir, "-lm"],
                           capture_output=True, text=True)
    assert build.returncode == 0, build.stderr[-2000:]
    run = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    assert run.returncode == 0, (run.stdout[-500:], run.stderr[-2000:])
    assert "c caller ok" in run.stdout
