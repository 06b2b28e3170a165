"""AddressSanitizer + UndefinedBehaviorSanitizer build of the C++ host runtime (csrc/host/*.cpp)
driven by its self-test (csrc/host/selftest/host_selftest.cpp): dataset index builders, bucket planner,
shared-memory mailbox with a producer thread.  The reference has no sanitizer target."""
import os
import shutil
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


@pytest.mark.skipif(shutil.which("g++") is None, reason="no g++")
def test_host_runtime_under_asan_ubsan(tmp_path):
    exe = tmp_path / "host_selftest"
    srcs = sorted(str(p) for p in (ROOT / "csrc/host").glob("*.cpp")) + [str(ROOT / "csrc/host/selftest/host_selftest.cpp")]
    cmd = ["g++", "-std=c++17", "-g", "-O1", "-fsanitize=address,undefined", "-fno-omit-frame-pointer", "-pthread",
           "-I", str(ROOT / "include"), *srcs, "-o", str(exe), "-lrt"]
    build = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    if build.returncode != 0 and "sanitize" in build.stderr.lower() and "cannot find" in build.stderr.lower():
        pytest.skip("sanitizer runtime libraries not installed")
    assert build.returncode == 0, build.stderr[-3000:]
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=1:abort_on_error=0",
               UBSAN_OPTIONS="halt_on_error=1:print_stacktrace=1")
    run = subprocess.run([str(exe)], capture_output=True, text=True, timeout=300, env=env)
    assert run.returncode == 0, (run.stdout + run.stderr)[-3000:]
    assert "host_selftest: ok" in run.stdout
    assert "runtime error" not in run.stderr and "AddressSanitizer" not in run.stderr
