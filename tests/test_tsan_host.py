"""ThreadSanitizer run of the native host runtime ("race detection / sanitizers"): the library's
host code is rebuilt with -fsanitize=thread and hammered from 8 threads by tests/native/tsan_host.cpp.  No GPU needed -
every entry point is driven up to its first driver call, which fails cleanly on a CPU box."""
import os
import shutil
import subprocess

import pytest

from distributedvolunteercomputing_b200.ops import build as B


def _tsan_toolchain_ok(tmp_path) -> bool:
    if shutil.which("g++") is None or not os.path.exists(B.nvcc_path()):
        return False
    src = tmp_path / "probe.cpp"
    src.write_text("int main() { return 0; }\n")
    return subprocess.run(["g++", "-fsanitize=thread", str(src), "-o", str(tmp_path / "probe")], capture_output=True).returncode == 0


def test_host_runtime_is_race_free_under_tsan(tmp_path):
    if not _tsan_toolchain_ok(tmp_path):
        pytest.skip("no ThreadSanitizer-capable toolchain on this box")
    exe = B.build_tsan()
    env = {**os.environ, "TSAN_OPTIONS": "halt_on_error=0 exitcode=66"}
    # the sanitizer is live in this build: a deliberate race in the harness is reported
    st = subprocess.run([exe, "--selftest"], capture_output=True, text=True, env=env, timeout=300)
    assert st.returncode == 66 and "ThreadSanitizer: data race" in st.stderr, (st.returncode, st.stderr[-400:])
    # and the library itself is instrumented
    nm = subprocess.run(["nm", "-D", "--undefined-only", os.path.join(os.path.dirname(exe), "libdvc_b200_tsan.so")],
                        capture_output=True, text=True).stdout
    assert "__tsan_func_entry" in nm and "__tsan_read" in nm
    r = subprocess.run([exe, "8", "300"], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, (r.returncode, r.stdout[-400:], r.stderr[-2000:])
    assert "tsan_host ok: 8 threads x 300 iterations" in r.stdout
    assert "ThreadSanitizer" not in r.stderr, r.stderr[-2000:]
