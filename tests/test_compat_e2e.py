"""Loopback integration of the compat surface: coordinator + clients over UDP/ZeroMQ, CPU engine.

Covers SURVEY.md §3.1-3.4: join handshake, `request <path>`, chunk relay, in-order mp4 output, the
job-time line - and the fix the reference lacks: a worker quitting mid-stream does not strand the job."""
import os
import socket
import threading
import time

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")
pytest.importorskip("zmq")

from distributedvolunteercomputing_b200 import protocol as proto
from distributedvolunteercomputing_b200.client import client
from distributedvolunteercomputing_b200.config import Config
from distributedvolunteercomputing_b200.coordinator import coordinator
from distributedvolunteercomputing_b200.engine import BatchResult
from distributedvolunteercomputing_b200.session import write_synthetic_video


def _udp_port():
    s = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class StubEngine:
    """Cheap stand-in for the model: 400-wide resize + frame number burned in; optional delay."""

    def __init__(self, delay=0.0):
        self.delay = delay
        self.frames = 0

    def process(self, frames, requester_id=None, annotate=True):
        time.sleep(self.delay)
        frames = np.asarray(frames)
        B, h, w = frames.shape[:3]
        out = np.stack([cv2.resize(f, (400, int(h * 400 / w)), interpolation=cv2.INTER_AREA) for f in frames])
        self.frames += B
        z = np.zeros
        return BatchResult(z((B, 100, 7), np.float32), z((B,), np.int32), z((B, 100, 4), np.int32), z((B,), np.int32), out)


def _cfg(control_port, **kw):
    base = dict(control_port=control_port, my_port=0, number_of_frames_in_chunk=8, control_retry_s=1.0,
                data_port_lo=0, data_port_hi=-1, worker_timeout_s=3.0, data_timeout_s=2.0)
    base.update(kw)
    return Config(**base)


def _count_frames(path):
    cap = cv2.VideoCapture(path)
    n = 0
    while True:
        ok, f = cap.read()
        if not ok:
            break
        n += 1
    cap.release()
    return n


def test_request_video_roundtrip_with_real_cpu_engine(tmp_path, graph, capsys):
    from distributedvolunteercomputing_b200.engine import CpuEngine
    port = _udp_port()
    srv = coordinator("127.0.0.1", _cfg(port))
    video = write_synthetic_video(str(tmp_path / "in.mp4"), 11, 120, 160)
    try:
        eng = CpuEngine(graph, _cfg(port))
        w = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=eng, out_dir=str(tmp_path))
        r = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(), out_dir=str(tmp_path))
        assert w.ready.wait(5) and r.ready.wait(5)
        r.become_requester(video)
        secs = r.wait_job(90)
        assert secs is not None, srv.sched.snapshot()
        assert r.final_sent_frame == 11                         # frame 0 kept, partial tail chunk (3 frames) flushed
        out = str(tmp_path / "video0.mp4")
        assert os.path.exists(out) and _count_frames(out) == 11
        assert eng.model is not None and w.metrics.counters.get("frames_processed") == 11
        text = capsys.readouterr().out
        assert "final frame sent : 11" in text and "final frame time taken for the job = " in text
        snap = srv.sched.snapshot()
        assert snap["inflight"] == 0 and snap["pending"] == 0
        w.exit_threads(); r.exit_threads()
        assert all(not t.is_alive() for t in w.threads + r.threads)      # quit really terminates
    finally:
        srv.exit_threads()
    assert all(not t.is_alive() for t in srv.threads)


def test_worker_leaving_midstream_does_not_strand_the_job(tmp_path):
    port = _udp_port()
    srv = coordinator("127.0.0.1", _cfg(port))
    video = write_synthetic_video(str(tmp_path / "in.mp4"), 64, 96, 128)
    try:
        slow = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(delay=0.4), out_dir=str(tmp_path))
        fast = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(delay=0.05), out_dir=str(tmp_path))
        req = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(), out_dir=str(tmp_path))
        for c in (slow, fast, req):
            assert c.ready.wait(5)
        req.become_requester(video)
        time.sleep(0.6)                                   # chunks are now in flight on both workers
        slow.exit_threads()                               # volunteer quits (`end`) while it owns chunks
        secs = req.wait_job(60)
        assert secs is not None, srv.sched.snapshot()     # the reference hangs here forever
        assert _count_frames(str(tmp_path / "video0.mp4")) == 64
        assert srv.sched.stats["redispatched"] >= 1
        fast.exit_threads(); req.exit_threads()
    finally:
        srv.exit_threads()


@pytest.mark.parametrize("fault", ["inject_stall_after", "inject_leave_after"])
def test_fault_injection_flags_job_still_completes(tmp_path, fault):
    """`--inject-stall-after N` (accepts chunks, never answers) and `--inject-leave-after N` (dies without `end`):
    the coordinator's in-flight timeout re-dispatches the lost chunks and the output is complete and in order."""
    port = _udp_port()
    srv = coordinator("127.0.0.1", _cfg(port, worker_timeout_s=1.5))
    video = write_synthetic_video(str(tmp_path / "in.mp4"), 48, 96, 128)
    try:
        bad = client("127.0.0.1", "127.0.0.1", _cfg(port, **{fault: 1}), engine=StubEngine(delay=0.02), out_dir=str(tmp_path))
        good = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(delay=0.02), out_dir=str(tmp_path))
        req = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(), out_dir=str(tmp_path))
        for c in (bad, good, req):
            assert c.ready.wait(5)
        req.become_requester(video)
        secs = req.wait_job(60)
        assert secs is not None, srv.sched.snapshot()
        assert _count_frames(str(tmp_path / "video0.mp4")) == 48
        assert srv.sched.stats["redispatched"] >= 1 and bad.chunks_taken >= 2
        good.exit_threads(); req.exit_threads()
        if fault == "inject_stall_after":
            bad.exit_threads()
    finally:
        srv.exit_threads()


def test_slow_only_worker_is_revived_and_the_job_finishes(tmp_path):
    """The only worker needs longer per chunk than the in-flight timeout: it is presumed dead, its late results are
    still accepted (cancelling the re-dispatch) and it is taken back, so the job completes instead of stalling."""
    port = _udp_port()
    srv = coordinator("127.0.0.1", _cfg(port, worker_timeout_s=0.6))
    video = write_synthetic_video(str(tmp_path / "in.mp4"), 24, 96, 128)
    try:
        slow = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(delay=1.3), out_dir=str(tmp_path))
        req = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(), out_dir=str(tmp_path))
        assert slow.ready.wait(5) and req.ready.wait(5)
        req.become_requester(video)
        secs = req.wait_job(90)
        assert secs is not None, srv.sched.snapshot()
        assert _count_frames(str(tmp_path / "video0.mp4")) == 24
        st = srv.sched.stats
        assert st["timeouts"] >= 1 and st["revived"] >= 1 and st["late_results"] >= 1, dict(st)
        # a chunk that was re-dispatched to the revived worker before its first result arrived may still be in flight as a
        # harmless duplicate when the job completes: it must drain, not linger
        deadline = time.time() + 15
        while srv.sched.outstanding() and time.time() < deadline:
            time.sleep(0.1)
        snap = srv.sched.snapshot()
        assert snap["inflight"] == 0 and snap["pending"] == 0, snap
        slow.exit_threads(); req.exit_threads()
    finally:
        srv.exit_threads()


def test_idempotent_join_and_compat_drop_modes(tmp_path):
    port = _udp_port()
    srv = coordinator("127.0.0.1", _cfg(port))
    try:
        s = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
        p = proto.free_tcp_port()
        r1 = proto.control_roundtrip(s, ("127.0.0.1", port), proto.make_control("join", f"127.0.0.1:{p}"), 1.0, 3)
        r2 = proto.control_roundtrip(s, ("127.0.0.1", port), proto.make_control("join", f"127.0.0.1:{p}"), 1.0, 3)
        assert r1 == r2 and r1.startswith("ok||") and len(srv.sched.members.alive()) == 1
        assert proto.control_roundtrip(s, ("127.0.0.1", port), proto.make_control("end", f"127.0.0.1:{p}"), 1.0, 3) == "ok"
        # reference-compat mode: frame 0 dropped and the partial tail never sent (worker.py:107,112,125-126)
        video = write_synthetic_video(str(tmp_path / "in.mp4"), 20, 96, 128)
        w = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(), out_dir=str(tmp_path))
        r = client("127.0.0.1", "127.0.0.1", _cfg(port, compat_drop_first=True, compat_drop_tail=True),
                   engine=StubEngine(), out_dir=str(tmp_path))
        assert w.ready.wait(5) and r.ready.wait(5)
        r.become_requester(video)
        assert r.wait_job(30) is not None
        assert r.final_sent_frame == 16 and _count_frames(str(tmp_path / "video0.mp4")) == 16
        w.exit_threads(); r.exit_threads()
    finally:
        srv.exit_threads()


def test_job_journal_and_resume(tmp_path):
    from distributedvolunteercomputing_b200.session import JobJournal, JobTracker
    j = JobJournal(str(tmp_path / "video0.mp4.journal"), every=4)
    tr = JobTracker(None, first_frame=1, out=lambda m: None, journal=j)
    for n in (2, 1, 3, 4, 5):
        tr.push(n, np.zeros((2, 2, 3), np.uint8))
    assert JobJournal.load(j.path) in (4, 5)          # flushed every 4 contiguous frames
    tr.set_final(5)
    assert tr.finished and JobJournal.load(j.path) == 5
    tr2 = JobTracker(None, first_frame=JobJournal.load(j.path) + 1, out=lambda m: None)
    assert tr2.reorder.next == 6
    assert JobJournal.load(str(tmp_path / "missing.journal")) == 0


def test_trace_span_records_metrics():
    from distributedvolunteercomputing_b200.utils.metrics import Metrics
    from distributedvolunteercomputing_b200.utils.trace import span
    m = Metrics()
    with span("batch", m, n=3):
        time.sleep(0.01)
    assert m.percentile("batch_ms", 50) >= 5.0
