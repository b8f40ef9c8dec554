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
    tr2 = JobTracker(None, first_frame=JobJournal.load(j.path) + 1, out=lambda m: None)
    assert tr2.reorder.next in (5, 6)
    tr.set_final(5)
    st = JobJournal.load_state(j.path)
    assert tr.finished and st["finished"] and st["last_contiguous_frame"] == 5
    assert JobJournal.load(j.path) == 0, "the journal of a finished job must not make a later job skip frames"
    assert JobJournal.load(str(tmp_path / "missing.journal")) == 0
    # a frame the pool gave up is passed over by the in-order writer
    tr3 = JobTracker(None, first_frame=1, out=lambda m: None)
    tr3.push(1, np.zeros((2, 2, 3), np.uint8)); tr3.push(3, np.zeros((2, 2, 3), np.uint8)); tr3.skip(2)
    tr3.set_final(3)
    assert tr3.finished and tr3.lost == 1


def test_resume_keeps_the_head_and_writes_the_tail_to_a_new_segment(tmp_path):
    """become_requester(resume=True) end to end with a real sink: the frames written before the interruption stay in
    video0.mp4, the tail goes to video0.part1.mp4; a journal of a finished job or of another source is ignored."""
    import json
    port = _udp_port()
    srv = coordinator("127.0.0.1", _cfg(port))
    try:
        video = write_synthetic_video(str(tmp_path / "in.mp4"), 20, 96, 128)
        w = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(), out_dir=str(tmp_path))
        r = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(), out_dir=str(tmp_path))
        assert w.ready.wait(5) and r.ready.wait(5)
        r.become_requester(video)
        assert r.wait_job(30) is not None and _count_frames(str(tmp_path / "video0.mp4")) == 20
        r.exit_threads()
        jpath = str(tmp_path / "video0.mp4.journal")
        assert json.load(open(jpath))["finished"] is True
        # (a) finished journal: a resume of the same source starts over (nothing skipped)
        r2 = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(), out_dir=str(tmp_path))
        assert r2.ready.wait(5)
        r2.become_requester(video, resume=True)
        assert r2.wait_job(30) is not None and _count_frames(str(tmp_path / "video0.mp4")) == 20
        assert not (tmp_path / "video0.part1.mp4").exists()
        r2.exit_threads()
        # (b) interrupted after frame 8 of this source: the tail (12 frames) goes to a new segment, the head file is kept
        json.dump({"last_contiguous_frame": 8, "source": str(tmp_path / "in.mp4"), "segment": 0, "finished": False}, open(jpath, "w"))
        r3 = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(), out_dir=str(tmp_path))
        assert r3.ready.wait(5)
        r3.become_requester(video, resume=True)
        assert r3.wait_job(30) is not None
        assert _count_frames(str(tmp_path / "video0.part1.mp4")) == 12 and _count_frames(str(tmp_path / "video0.mp4")) == 20
        r3.exit_threads()
        # (c) journal of another source: ignored
        json.dump({"last_contiguous_frame": 8, "source": "/somewhere/else.mp4", "segment": 0, "finished": False}, open(jpath, "w"))
        r4 = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(), out_dir=str(tmp_path))
        assert r4.ready.wait(5)
        r4.become_requester(video, resume=True)
        assert r4.wait_job(30) is not None and _count_frames(str(tmp_path / "video0.mp4")) == 20
        r4.exit_threads(); w.exit_threads()
    finally:
        srv.exit_threads()


def test_malformed_chunk_does_not_kill_the_ingest_thread(tmp_path):
    """A bad header / undecodable message on a client's data port is answered and skipped; the next good job completes."""
    port = _udp_port()
    srv = coordinator("127.0.0.1", _cfg(port))
    try:
        video = write_synthetic_video(str(tmp_path / "in.mp4"), 12, 96, 128)
        w = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(), out_dir=str(tmp_path))
        r = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(), out_dir=str(tmp_path))
        assert w.ready.wait(5) and r.ready.wait(5)
        junk = np.zeros((2, 2, 3), np.uint8)
        r.sender.send_image("not-a-header", junk)                       # malformed header -> coordinator ingest thread
        r.sender.send_image("a||request||x-y||2||2", junk)              # unparsable frame numbers
        r.sender.sock.send_multipart([b"{this is not json", b"\x00"])     # undecodable imagezmq message
        assert r.sender.sock.recv() == b"OK"
        # straight at the worker's own hub as well (client.recv_image_thread)
        direct = proto.ImageSender("tcp://127.0.0.1:" + w.my_ip.split(":")[1])
        direct.send_image("garbage", junk)
        direct.close()
        r.become_requester(video)
        assert r.wait_job(30) is not None and _count_frames(str(tmp_path / "video0.mp4")) == 12
        assert srv.metrics.counters["bad_messages"] >= 3 and w.metrics.counters["bad_messages"] >= 1
        w.exit_threads(); r.exit_threads()
    finally:
        srv.exit_threads()


def test_trace_span_records_metrics():
    from distributedvolunteercomputing_b200.utils.metrics import Metrics
    from distributedvolunteercomputing_b200.utils.trace import span
    m = Metrics()
    with span("batch", m, n=3):
        time.sleep(0.01)
    assert m.percentile("batch_ms", 50) >= 5.0


def test_chunk_given_up_by_the_pool_reaches_the_requester_as_a_gap(tmp_path):
    """Two volunteers that swallow every chunk: each chunk times out on its volunteer and is dropped (max_attempts = 1) - and the
    requester is TOLD (a `dropped` marker per chunk), so its in-order writer finishes instead of waiting for ever."""
    port = _udp_port()
    srv = coordinator("127.0.0.1", _cfg(port, worker_timeout_s=0.6))
    srv.sched.max_attempts = 1           # the first time-out already gives the chunk up (both volunteers are then presumed dead)
    try:
        video = write_synthetic_video(str(tmp_path / "in.mp4"), 16, 96, 128)
        a = client("127.0.0.1", "127.0.0.1", _cfg(port, inject_stall_after=1), engine=StubEngine(), out_dir=str(tmp_path))
        b = client("127.0.0.1", "127.0.0.1", _cfg(port, inject_stall_after=1), engine=StubEngine(), out_dir=str(tmp_path))
        for c in (a, b):
            c.chunks_taken = 10 ** 6          # every chunk counts as "after the limit": swallowed
        r = client("127.0.0.1", "127.0.0.1", _cfg(port), engine=StubEngine(), out_dir=str(tmp_path))
        assert a.ready.wait(5) and b.ready.wait(5) and r.ready.wait(5)
        r.become_requester(video)
        assert r.wait_job(30) is not None, "the requester was never told about the dropped chunks"
        assert r.tracker.lost == 16 and srv.sched.stats["dropped_after_max_attempts"] == 2
        assert srv.metrics.counters["dropped_chunks"] == 2 and not srv.missing
        a.exit_threads(); b.exit_threads(); r.exit_threads()
    finally:
        srv.exit_threads()


def test_long_live_stream_keeps_the_servers_tables_bounded():
    """`request live` never ends (worker.py:96,114): after 20 000 chunks (160 000 frames) through the coordinator's
    ingest / dispatch / completion bookkeeping every table is as small as after the first chunk and the Python heap
    did not grow with the stream - only counters did."""
    import queue
    import tracemalloc
    srv = coordinator("127.0.0.1", _cfg(_udp_port()), start=False)
    srv.running = True
    req, wrk = "10.0.0.1:5554", "10.0.0.2:5554"
    for a in (req, wrk):
        srv.sched.join(a)
    srv.sched.become_requester(req)
    srv.results[req] = queue.Queue()
    img = np.zeros((2, 8 * 2, 3), np.uint8)

    def one_chunk(i):
        frames = list(range(i * 8 + 1, i * 8 + 9))
        srv._on_chunk(proto.Header(req, "request", frames, 2, 2).encode(), img, req)
        batch, worker, _ = srv.sched.next_dispatch()
        assert worker.addr == wrk and list(batch.frame_ids) == frames
        # results come back in two half-chunks, the second one twice (a re-sent reply)
        srv._on_chunk(proto.Header(req, "processed", frames[:4], 2, 2).encode(), img, wrk)
        srv._on_chunk(proto.Header(req, "processed", frames[4:], 2, 2).encode(), img, wrk)
        srv._on_chunk(proto.Header(req, "processed", frames[4:], 2, 2).encode(), img, wrk)
        assert srv.results[req].qsize() == 2
        srv.results[req].get_nowait(), srv.results[req].get_nowait()

    def sizes():
        return (len(srv.frame_owner[req]), len(srv.missing), len(srv.sched.inflight), srv.sched.pending(), len(srv.sched.dropped),
                len(srv.sched.done), len(srv.sched.members.workers))

    for i in range(200):
        one_chunk(i)
    base = sizes()
    tracemalloc.start()
    snap0 = tracemalloc.take_snapshot()
    n = 20000
    for i in range(200, 200 + n):
        one_chunk(i)
    snap1 = tracemalloc.take_snapshot()
    tracemalloc.stop()
    assert sizes() == base == (0, 0, 0, 0, 0, 1, 2)
    assert srv.sched.done[req] == 200 + n and srv.sched.stats["completed"] == 200 + n
    assert srv.metrics.counters["frames_in"] == 8 * (200 + n) and srv.metrics.counters["duplicate_chunks"] == 200 + n
    grown = sum(s.size_diff for s in snap1.compare_to(snap0, "filename") if s.size_diff > 0)
    assert grown < 256 * 1024, f"{grown} bytes retained over {n} chunks"
