"""Race detection for the C++ host runtime: build ``tests/host/tsan_stress.cpp`` against ``csrc/host/*.cpp``
with ThreadSanitizer and run it (the reference has no race detector; its pools are `synchronized`)."""
import shutil
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def test_host_runtime_is_clean_under_thread_sanitizer(tmp_path):
    gxx = shutil.which("g++")
    if gxx is None:
        pytest.skip("g++ not available")
    exe = tmp_path / "tsan_stress"
    srcs = [ROOT / "tests/host/tsan_stress.cpp", ROOT / "csrc/host/pool.cpp", ROOT / "csrc/host/parse.cpp"]
    build = subprocess.run([gxx, "-O1", "-g", "-std=c++17", "-pthread", "-fsanitize=thread", "-I", str(ROOT / "include")]
                           + [str(s) for s in srcs] + ["-o", str(exe)], capture_output=True, text=True)
    if build.returncode != 0 and ("tsan" in build.stderr.lower() or "sanitize" in build.stderr.lower()):
        pytest.skip("ThreadSanitizer runtime not installed: " + build.stderr[-200:])
    assert build.returncode == 0, build.stderr
    run = subprocess.run([str(exe)], capture_output=True, text=True, timeout=300)
    if "FATAL: ThreadSanitizer" in run.stderr and "unexpected memory mapping" in run.stderr:
        pytest.skip("ThreadSanitizer cannot run under this kernel's address-space layout")
    assert "WARNING: ThreadSanitizer" not in run.stderr, run.stderr[-3000:]
    assert run.returncode == 0 and "errors=0" in run.stdout, (run.stdout, run.stderr[-2000:])
