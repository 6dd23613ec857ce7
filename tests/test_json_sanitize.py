"""Builds the JSON packer with AddressSanitizer + UBSan and runs the mutation harness
(tests/native/json_fuzz.cpp).  The reference relies on Go's memory safety for this code
(encoding/json, handlers.go:2047-2050); the C++ replacement has to earn it."""
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.timeout(280)
def test_packer_under_asan_ubsan(tmp_path):
    exe = tmp_path / "json_fuzz"
    cmd = ["/usr/bin/g++", "-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=all",
           "-o", str(exe), os.path.join(HERE, "native", "json_fuzz.cpp"), "-pthread"]
    build = subprocess.run(cmd, capture_output=True, text=True)
    if build.returncode != 0 and "asan" in (build.stderr + build.stdout).lower():
        pytest.skip("no sanitizer runtime in this image")
    assert build.returncode == 0, build.stderr[-2000:]
    run = subprocess.run([str(exe)], capture_output=True, text=True, timeout=240)
    assert run.returncode == 0, (run.stdout[-500:], run.stderr[-3000:])
    assert "ok=" in run.stdout


@pytest.mark.timeout(120)
def test_oracle_under_asan_ubsan(tmp_path):
    exe = tmp_path / "oracle_sanitize"
    cmd = ["/usr/bin/gcc", "-O1", "-g", "-fopenmp", "-fsanitize=address,undefined", "-fno-sanitize-recover=all",
           "-o", str(exe), os.path.join(HERE, "native", "oracle_sanitize.c"), "-lquadmath", "-lm"]
    build = subprocess.run(cmd, capture_output=True, text=True)
    if build.returncode != 0 and "asan" in (build.stderr + build.stdout).lower():
        pytest.skip("no sanitizer runtime in this image")
    assert build.returncode == 0, build.stderr[-2000:]
    run = subprocess.run([str(exe)], capture_output=True, text=True, timeout=100, env=dict(os.environ, ASAN_OPTIONS="detect_leaks=0"))
    assert run.returncode == 0, (run.stdout[-500:], run.stderr[-3000:])
    assert "oracle sanitize ok" in run.stdout
