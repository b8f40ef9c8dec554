"""The public entrypoints: `python server.py [ip]` / `python worker.py [server_ip own_ip]` with the reference's REPL
commands, and wire compatibility of the new client with the UNMODIFIED reference server."""
import os
import re
import socket
import subprocess
import sys
import time

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")
SHIMS = os.path.join(ROOT, "baseline", "ref_harness", "shims")


def _udp_port():
    s = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_server_and_worker_cli_repl(tmp_path):
    pytest.importorskip("zmq")
    port = _udp_port()
    env = dict(os.environ, PYTHONUNBUFFERED="1", PYTHONPATH=ROOT)
    srv = subprocess.Popen([sys.executable, os.path.join(ROOT, "server.py"), "127.0.0.1", "--control-port", str(port)],
                           stdin=subprocess.PIPE, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=tmp_path, env=env)
    try:
        time.sleep(1.0)
        wrk = subprocess.Popen([sys.executable, os.path.join(ROOT, "worker.py"), "127.0.0.1", "127.0.0.1", "--control-port", str(port),
                                "--my-port", "0", "--control-retry-s", "1", "--device", "cpu"],
                               stdin=subprocess.PIPE, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=tmp_path, env=env)
        out, _ = wrk.communicate("request\nend\nquit\n", timeout=60)     # bare `request` prints usage instead of crashing
        assert "usage: request live | request <path-to-video>" in out
        assert "done." in out and wrk.returncode == 0
        sout, _ = srv.communicate("quit\n", timeout=30)
        assert "done." in sout and "manager terminated." in sout and srv.returncode == 0
    finally:
        for p in (srv,):
            if p.poll() is None:
                p.kill()


@pytest.mark.skipif(not os.path.exists(os.path.join(REF, "server.py")), reason="reference not installed in baseline/_ref")
def test_new_clients_against_unmodified_reference_server(tmp_path):
    """Wire compatibility: our worker + requester talk to the reference's own server.py (UDP grammar + imagezmq wire)."""
    cv2 = pytest.importorskip("cv2")
    pytest.importorskip("zmq")
    probe = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
    try:
        probe.bind(("", 9999))           # the reference server's control port is fixed (server.py:11)
    except OSError:
        pytest.skip("UDP port 9999 is busy")
    finally:
        probe.close()
    from distributedvolunteercomputing_b200.client import client
    from distributedvolunteercomputing_b200.config import Config
    from distributedvolunteercomputing_b200.session import write_synthetic_video
    from tests.test_compat_e2e import StubEngine, _count_frames
    env = dict(os.environ, PYTHONUNBUFFERED="1", PYTHONPATH=os.pathsep.join([SHIMS, REF]))
    srv = subprocess.Popen([sys.executable, os.path.join(REF, "server.py"), "127.0.0.1"], stdin=subprocess.PIPE,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=tmp_path, env=env, start_new_session=True)
    try:
        time.sleep(1.5)
        cfg = lambda: Config(control_port=9999, my_port=0, number_of_frames_in_chunk=8, control_retry_s=2.0)
        w = client("127.0.0.1", "127.0.0.1", cfg(), engine=StubEngine(), out_dir=str(tmp_path))
        r = client("127.0.0.1", "127.0.0.1", cfg(), engine=StubEngine(), out_dir=str(tmp_path))
        assert w.ready.wait(5) and r.ready.wait(5)
        video = write_synthetic_video(str(tmp_path / "in.mp4"), 16, 96, 128)
        r.become_requester(video)
        assert r.wait_job(60) is not None
        assert r.final_sent_frame == 16 and _count_frames(str(tmp_path / "video0.mp4")) == 16
        w.exit_threads(); r.exit_threads()
    finally:
        import signal
        try:
            os.killpg(os.getpgid(srv.pid), signal.SIGKILL)     # the reference server never exits on its own
            srv.wait(timeout=10)                               # its fixed ports are free again for the next test
        except Exception:
            pass


def test_bench_contract_without_a_gpu():
    """bench.py on a CPU box: the native arm refuses with one clear line (no CPU fallback posing as a number), the
    reference arm prints exactly one JSON line with the contract's keys (or the documented `unavailable` line)."""
    import json
    import torch
    if torch.cuda.is_available():
        pytest.skip("CPU-box behaviour")
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "0"], cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=300)
    assert p.returncode != 0 and "needs a CUDA device" in p.stderr and p.stdout.strip() == ""
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference"
    if "unavailable" not in d:
        for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                    "vs_baseline", "dtype", "data", "config"):
            assert key in d, key
        assert d["value"] > 0 and d["unit"] == "frames/s" and d["config"]["global_batch"] == 64 and d["data"] == "synthetic"
