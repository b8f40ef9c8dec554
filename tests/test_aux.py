"""Auxiliary subsystems: thread-safety of the scheduler under real threads (the reference shares plain
lists between threads without locks, server.py:20-24 / worker.py:33-38), property tests for the reorder buffer and the
wire header, metrics / trace / clock-sampler helpers.  CPU only."""
import json
import random
import threading
import time

import numpy as np
from hypothesis import given, settings, strategies as st

from distributedvolunteercomputing_b200 import protocol as proto
from distributedvolunteercomputing_b200.parallel.scheduler import ReorderBuffer, Scheduler
from distributedvolunteercomputing_b200.utils.clocks import ClockSampler
from distributedvolunteercomputing_b200.utils.metrics import Logger, Metrics
from distributedvolunteercomputing_b200.utils.trace import span


def test_scheduler_under_real_threads_exactly_once():
    """Six worker threads pull, 'compute' and complete concurrently while a chaos thread makes workers leave and
    re-join and a submitter keeps adding batches: every frame is delivered exactly once, nothing is left in flight."""
    s = Scheduler(worker_timeout_s=30.0)
    workers = [f"w{i}" for i in range(6)]
    for w in workers:
        s.join(w, capacity=2)
    s.join("client"); s.become_requester("client")
    n_batches, bs = 300, 8
    delivered = []                    # frame ids, appended under lock by whoever completes
    dl = threading.Lock()
    stop = threading.Event()
    errors = []

    def dispatcher_and_workers():
        # one dispatcher hands work to per-worker inboxes, like coordinator.send_request
        inbox = {w: [] for w in workers}
        il = threading.Lock()

        def dispatch():
            while not stop.is_set():
                d = s.wait_dispatch(timeout=0.01)
                if d is None:
                    continue
                b, w, _ = d
                with il:
                    inbox[w.addr].append(b)

        def work(name, seed):
            rng = random.Random(seed)
            while not stop.is_set():
                with il:
                    b = inbox[name].pop(0) if inbox[name] else None
                if b is None:
                    time.sleep(0.0005)
                    continue
                time.sleep(rng.random() * 0.002)
                try:
                    done = s.complete(b.batch_id, name)     # None when the batch was re-dispatched meanwhile
                except Exception as e:                       # pragma: no cover
                    errors.append(e)
                    return
                if done is not None:
                    with dl:
                        delivered.extend(done.frame_ids)

        ts = [threading.Thread(target=dispatch, daemon=True)]
        ts += [threading.Thread(target=work, args=(w, i), daemon=True) for i, w in enumerate(workers)]
        return ts

    def chaos():
        rng = random.Random(99)
        while not stop.is_set():
            w = rng.choice(workers[2:])
            s.leave(w)
            time.sleep(0.003)
            s.join(w, capacity=2)
            time.sleep(0.003)

    def submitter():
        for b in range(n_batches):
            s.submit("client", range(b * bs + 1, (b + 1) * bs + 1))
            if b % 50 == 0:
                time.sleep(0.002)

    threads = dispatcher_and_workers() + [threading.Thread(target=chaos, daemon=True), threading.Thread(target=submitter, daemon=True)]
    for t in threads:
        t.start()
    deadline = time.time() + 60
    while time.time() < deadline:
        with dl:
            if len(delivered) >= n_batches * bs:
                break
        time.sleep(0.01)
    stop.set()
    for t in threads:
        t.join(timeout=2)
    assert not errors
    assert sorted(delivered) == list(range(1, n_batches * bs + 1))      # exactly once, none lost
    assert s.outstanding() == 0 and s.pending() == 0


@settings(max_examples=60, deadline=None)
@given(st.permutations(list(range(1, 41))), st.lists(st.integers(1, 40), max_size=15))
def test_reorder_buffer_any_arrival_order_with_duplicates(order, dups):
    """In-order writer (worker.py:214-239): any arrival order plus late duplicates yields 1..n exactly once, in order."""
    rb = ReorderBuffer(1)
    out = []
    arrivals = list(order)
    rng = random.Random(len(dups))
    for d in dups:
        arrivals.insert(rng.randrange(len(arrivals) + 1), d)
    for n in arrivals:
        out.extend(k for k, _ in rb.push(n, None))
    assert out == list(range(1, 41)) and len(rb) == 0 and rb.next == 41


@settings(max_examples=60, deadline=None)
@given(st.sampled_from(["request", "processed"]), st.lists(st.integers(1, 10**6), min_size=1, max_size=100, unique=True),
       st.integers(1, 2160), st.integers(1, 3840), st.integers(1, 255), st.integers(1024, 65535))
def test_header_roundtrip(cmd, frames, h, w, ip_tail, port):
    """Wire header `id||cmd||n1-...-nk||H||W` (worker.py:153-155) survives encode/parse for any contents."""
    hdr = proto.Header(f"10.0.0.{ip_tail}:{port}", cmd, frames, h, w)
    back = proto.parse_header(hdr.encode())
    assert (back.requester, back.cmd, list(back.frames), back.height, back.width) == (hdr.requester, cmd, frames, h, w)


@settings(max_examples=25, deadline=None)
@given(st.integers(1, 7), st.integers(1, 9), st.integers(1, 11))
def test_chunk_pack_unpack_roundtrip(k, h, w):
    """Side-by-side chunk canvas (worker.py:141-156, 172-177) is lossless for any frame count and size."""
    rng = np.random.default_rng(k * 100 + h * 10 + w)
    frames = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for _ in range(k)]
    canvas = proto.pack_chunk(frames)
    assert canvas.shape == (h, k * w, 3)
    back = proto.unpack_chunk(canvas, k, h, w)
    assert all(np.array_equal(a, b) for a, b in zip(frames, back))


def test_metrics_counters_percentiles_and_jsonl(tmp_path):
    path = tmp_path / "m.jsonl"
    m = Metrics(str(path))

    def bump():
        for i in range(1000):
            m.inc("frames", 2)
            m.observe("lat_ms", i % 100)

    ts = [threading.Thread(target=bump) for _ in range(4)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert m.counters["frames"] == 8000
    assert m.percentile("lat_ms", 50) in (49.0, 50.0) and m.percentile("lat_ms", 99) in (98.0, 99.0)
    assert m.percentile("missing", 50) is None
    with span("batch", m, frames=64):
        time.sleep(0.001)
    s = m.summary()
    assert s["batch_ms"]["n"] == 1 and s["batch_ms"]["p50"] >= 1.0 and s["counters"]["frames"] == 8000
    recs = [json.loads(l) for l in path.read_text().splitlines()]
    assert recs and recs[-1]["event"] == "batch" and recs[-1]["frames"] == 64 and recs[-1]["ms"] >= 1.0


def test_metrics_http_endpoint_on_the_coordinator():
    """`--metrics-port` serves Prometheus text and a JSON summary next to the reference's stdout lines."""
    import socket
    import urllib.request
    from distributedvolunteercomputing_b200.config import Config
    from distributedvolunteercomputing_b200.coordinator import coordinator
    s = socket.socket(socket.AF_INET, socket.SOCK_DGRAM); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    srv = coordinator("127.0.0.1", Config(control_port=port, metrics_port=0, data_port_lo=0, data_port_hi=-1))
    try:
        srv.metrics.inc("frames_in", 128)
        srv.metrics.observe("batch_ms", 1.5)
        http_port = srv.metrics_http.server_address[1]
        text = urllib.request.urlopen(f"http://127.0.0.1:{http_port}/metrics", timeout=5).read().decode()
        assert "dvc_frames_in_total 128" in text and 'dvc_batch_ms{quantile="0.5"} 1.5' in text
        assert "dvc_workers_alive 0" in text and "dvc_chunks_pending 0" in text
        summ = json.loads(urllib.request.urlopen(f"http://127.0.0.1:{http_port}/summary", timeout=5).read())
        assert summ["counters"]["frames_in"] == 128
    finally:
        srv.exit_threads()


def test_worker_declared_dead_can_rejoin_cleanly():
    """A worker the coordinator gave up on (send timeout) that joins again gets a fresh port, sender and threads; a
    repeated join of a live worker stays idempotent (worker.py:50-58 retries the datagram)."""
    import socket
    from distributedvolunteercomputing_b200.config import Config
    from distributedvolunteercomputing_b200.coordinator import coordinator
    s = socket.socket(socket.AF_INET, socket.SOCK_DGRAM); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    srv = coordinator("127.0.0.1", Config(control_port=port, data_port_lo=0, data_port_hi=-1))
    try:
        addr = "127.0.0.1:%d" % proto.free_tcp_port()
        r1 = srv.handle_control("join", addr)
        assert r1.startswith("ok||") and srv.handle_control("join", addr) == r1          # idempotent while alive
        old_stop = srv.ingest_stop[addr]
        srv.sched.leave(addr)                                                            # what a send timeout does
        r2 = srv.handle_control("join", addr)
        assert r2.startswith("ok||") and old_stop.is_set() and not srv.ingest_stop[addr].is_set()
        assert [w.addr for w in srv.sched.members.alive()] == [addr] and addr in srv.outbox and addr in srv.senders
        assert srv.handle_control("end", addr) == "ok" and addr not in srv.port_of and addr not in srv.outbox
    finally:
        srv.exit_threads()


def test_lru_cache_evicts_oldest_unpinned_and_calls_hook():
    from distributedvolunteercomputing_b200.utils.lru import LruCache
    gone = []
    c = LruCache(3, on_evict=lambda k, v: gone.append(k), can_evict=lambda v: not v.get("pinned"))
    c.put("a", {}); c.put("p", {"pinned": True}); c.put("b", {})
    assert c.get("a") is not None                 # refreshes a: order is now p, b, a
    c.put("c", {})                                # over capacity: p is pinned, so b goes
    assert gone == ["b"] and "a" in c and "p" in c and "c" in c and len(c) == 3
    c.put("d", {}); c.put("e", {})
    assert gone == ["b", "a", "c"] and "p" in c and c.evictions == 3
    for k in "xyz":                               # everything else pinned: the cache may exceed its capacity, never drops the newcomer
        c.put(k, {"pinned": True})
    assert "z" in c and "p" in c and len(c) >= 4
    c.clear(); assert len(c) == 0


def test_engine_state_cache_eviction_and_close(monkeypatch):
    """NativeEngine.state() bookkeeping without a GPU (device work stubbed out): same key -> same state, the oldest
    unpinned state is evicted and its native plan destroyed, stream / lane states (slot > 0) survive, close() frees all."""
    import contextlib
    import types
    import torch
    from distributedvolunteercomputing_b200.config import Config
    from distributedvolunteercomputing_b200.engine import NativeEngine, _ShapeState
    from distributedvolunteercomputing_b200.utils.lru import LruCache
    monkeypatch.setattr(torch.cuda, "device", lambda d: contextlib.nullcontext())
    monkeypatch.setattr(torch.cuda, "synchronize", lambda d=None: None)
    e = NativeEngine.__new__(NativeEngine)
    e.cfg, e.device, e.requester_id = Config(max_cached_states=3), torch.device("cpu"), ""
    destroyed = []
    e.lib = types.SimpleNamespace(dvc_plan_destroy=lambda p: destroyed.append(p))
    e.L = types.SimpleNamespace(set_device=lambda d: None)
    e._states = LruCache(3, on_evict=e._release_state, can_evict=lambda st: getattr(st, "evictable", False))

    def make(B, H, W, rid):
        st = _ShapeState()
        st.plan, st.graph_ann = ("plan", B, rid), object()
        return st

    e._make_state = make
    a = e.state(1, 2, 3)
    assert e.state(1, 2, 3) is a and e.state(1, 2, 3, requester_id="other") is not a
    lane = e.state(4, 2, 3, slot=5)                       # pinned: belongs to a stream / lane
    e.state(2, 2, 3)
    assert len(e._states) == 3 and destroyed == [("plan", 1, "")] and not hasattr(a, "graph_ann") and a.plan is None
    e.state(3, 2, 3); e.state(5, 2, 3)
    assert (4, 2, 3, "", 5) in e._states and lane.plan is not None and len(e._states) == 3
    n = len(destroyed)
    e.close()
    assert len(e._states) == 0 and len(destroyed) == n + 3
    e.close()                                             # idempotent (also what __del__ calls)


def _recorded_plan(monkeypatch, fuse, implicit=True):
    """NativeEngine._build_plan against a recording stand-in for the native library (no GPU needed)."""
    import ctypes as C
    import types
    import torch
    from distributedvolunteercomputing_b200.config import Config
    from distributedvolunteercomputing_b200.engine import NativeEngine, _ShapeState
    from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
    from distributedvolunteercomputing_b200.ops import lib as Lmod
    monkeypatch.setenv("DVC_FUSE_DWPW", "1" if fuse else "0")
    monkeypatch.setenv("DVC_IMPLICIT_CONV", "1" if implicit else "0")
    e = NativeEngine.__new__(NativeEngine)
    e.graph, e.cfg, e.device = SSDGraph.random(seed=1), Config(), torch.device("cpu")
    calls = []

    class FakeLib:
        def __getattr__(self, name):
            def f(*a):
                if name == "dvc_plan_create":
                    return 1234
                if name == "dvc_plan_num_ops":
                    return len(calls)
                calls.append((name.replace("dvc_plan_add_", ""),) + tuple(x for x in a[1:] if isinstance(x, int) and x < 10 ** 6))
                return 0
            return f

    e.lib = FakeLib()
    e.L = types.SimpleNamespace(ptr=lambda t: 0 if t is None else t.data_ptr(), check=lambda rc, what="": None, HeadOut=Lmod.HeadOut)
    e._pack_weights()
    e.priors, e.variances = torch.zeros(4), torch.zeros(4)
    st = _ShapeState()
    st.B, st.h4, st.w4, st.stride400 = 64, 400, 400, 480000
    st.frames400_flat = torch.zeros(16)
    st.t_lin = types.SimpleNamespace(xs=torch.zeros(1), xq=torch.zeros(1), ys=torch.zeros(1), yq=torch.zeros(1))
    st.mean = (C.c_float * 3)(127.5, 0, 0)
    st.loc, st.conf = torch.zeros((64, 1917 * 4)), torch.zeros((64, 1917 * 21))
    st.det = st.det_count = st.boxes_i = st.box_count = st.det_scratch = torch.zeros(1)
    st.det_params = Lmod.DetParams()
    e._build_plan(st)
    assert calls[-1][0] == "dvc_plan_set_head_branch" and calls[-1][1:] == (1,), "the six heads run on the plan's side branch"
    assert set(st.taps) == {"conv11", "conv13", "conv14_2", "conv15_2", "conv16_2", "conv17_2"}, "every tap has a buffer of its own"
    return calls[:-1], st.n_ops


def test_plan_construction_default_and_fused(monkeypatch):
    """The launch sequence the engine asks the native executor for: 42 ops by default (stem, 13 dw + 13 pw, 4 x (1x1,
    implicit 3x3) extras, 6 heads, detection), 46 with DVC_IMPLICIT_CONV=0 (an im2col op in front of each 3x3 extra);
    with the experimental switch the three stride-1 pairs whose pointwise has <= 256 outputs (conv1, conv3, conv5)
    become one fused op each and nothing else moves."""
    calls, n = _recorded_plan(monkeypatch, fuse=False)
    kinds = [c[0] for c in calls]
    assert n == 42 and kinds[0] == "preproc_conv0" and kinds[-1] == "detection"
    assert kinds.count("dwconv") == 13 and kinds.count("im2col") == 0 and kinds.count("conv3x3s2") == 4
    assert kinds.count("gemm") == 13 + 4 + 6 and "dwpw_fused" not in kinds
    assert [c[1:5] for c in calls if c[0] == "conv3x3s2"] == [(64, 10, 10, 256), (64, 5, 5, 128), (64, 3, 3, 128), (64, 2, 2, 64)]
    explicit, ne = _recorded_plan(monkeypatch, fuse=False, implicit=False)
    ek = [c[0] for c in explicit]
    assert ne == 46 and ek.count("im2col") == 4 and ek.count("gemm") == 13 + 8 + 6 and "conv3x3s2" not in ek
    assert [k for k in ek if k != "im2col"] == [("gemm" if k == "conv3x3s2" else k) for k in kinds]
    fused, nf = _recorded_plan(monkeypatch, fuse=True)
    fk = [c[0] for c in fused]
    assert nf == 39 and fk.count("dwpw_fused") == 3 and fk.count("dwconv") == 10 and fk.count("gemm") == 13 + 4 + 6 - 3
    assert [c[1:] for c in fused if c[0] == "dwpw_fused"] == [(64, 150, 150, 32, 64), (64, 75, 75, 128, 128), (64, 38, 38, 256, 256)]
    assert [k for k in fk if k != "dwpw_fused"] == [k for i, k in enumerate(kinds) if not (
        (k == "dwconv" and calls[i][1:] in {(64, 150, 150, 32, 1, 1), (64, 75, 75, 128, 1, 1), (64, 38, 38, 256, 1, 1)})
        or (k == "gemm" and i > 0 and kinds[i - 1] == "dwconv" and calls[i - 1][1:] in {(64, 150, 150, 32, 1, 1), (64, 75, 75, 128, 1, 1), (64, 38, 38, 256, 1, 1)}))]


def test_logger_is_gated_by_verbose(capsys):
    Logger(False, "[x] ").log("hidden")
    Logger(True, "[x] ").log("shown")
    out = capsys.readouterr().out
    assert "hidden" not in out and "[x] shown" in out


def test_clock_sampler_degrades_without_a_gpu():
    """No nvidia-smi / NVML on the CPU box: start/stop still work and report zero samples (bench.py relies on it)."""
    for backend in ("smi", "nvml"):
        r = ClockSampler(0, period_ms=50, backend=backend).start().stop()
        assert set(r) >= {"sm_mhz", "sm_max_mhz", "reasons", "samples"} and r["samples"] >= 0
