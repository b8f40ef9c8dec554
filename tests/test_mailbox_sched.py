"""ElasticBox (parallel/mailbox.py) host logic against a CPU stand-in for the mailbox client: dispatch by credit, results
exactly once, re-dispatch after a time-out / leave with the epoch bumped first, re-join, no attempts burnt while waiting
for a lane.  The device side of the same protocol is covered by tests/test_gpu_mailbox.py."""
import collections

from distributedvolunteercomputing_b200.parallel.mailbox import BoxLayout, ElasticBox
from distributedvolunteercomputing_b200.parallel.scheduler import Scheduler


class FakeClient:
    """Same surface as ClientBox.  A dispatch completes after `latency` polls; workers in `dead` never answer (the poll
    reports "timeout" after `timeout_polls`), like a wait kernel running into its deadline."""

    def __init__(self, workers=4, lanes=2, B=4, slots=32, latency=2, timeout_polls=6):
        self.layout = BoxLayout(workers, lanes, B, 8, 8, slots)
        self.latency, self.timeout_polls = latency, timeout_polls
        self.dead = set()
        self.flight = {}
        self.epochs = collections.defaultdict(list)
        self.stopped, self.resets, self.log = [], [], []
        self.free = list(range(slots))
        self.uploads = 0

    def dispatch(self, w, lane, slot, n_valid, batch_id, kind=0, host_src=None, fetch=True):
        assert (w, lane) not in self.flight, "lane dispatched twice"
        assert slot is not None
        self.uploads += host_src is not None
        self.flight[(w, lane)] = [0, batch_id, slot]
        self.log.append((w, lane, batch_id))

    def poll(self, w, lane):
        f = self.flight.get((w, lane))
        if f is None:
            return None
        f[0] += 1
        if f[0] >= 10 ** 9:
            del self.flight[(w, lane)]
            return "ok"
        if w in self.dead:
            if f[0] >= self.timeout_polls:
                del self.flight[(w, lane)]
                return "timeout"
            return None
        if f[0] >= self.latency:
            self.done = getattr(self, "done", {})
            self.done[(w, lane)] = f
            del self.flight[(w, lane)]
            return "ok"
        return None

    def result(self, w, lane):
        return {"batch_id": self.done[(w, lane)][1]}

    def cancel(self, w, lane):
        f = self.flight.get((w, lane))
        if f is not None:
            f[0] = 10 ** 9                # the client released the wait itself: the poll returns at once

    def set_epoch(self, w, e):
        self.epochs[w].append(e)

    def stop_worker(self, w):
        self.stopped.append(w)

    def reset_worker(self, w):
        self.resets.append(w)

    def alloc_slot(self):
        return self.free.pop() if self.free else None

    def free_slot(self, s):
        assert s not in self.free
        self.free.append(s)


def run(box, max_pumps=10000):
    for _ in range(max_pumps):
        box.pump()
        if not box.outstanding() and not box.flights:
            return True
    return False


def test_all_batches_complete_once_and_balanced():
    cl = FakeClient()
    got = []
    box = ElasticBox(cl, on_result=lambda b, v, w: got.append((b.batch_id, v["batch_id"], w)))
    for w in range(4):
        box.add_worker(w)
    ids = [box.submit_host(object(), range(i * 4, i * 4 + 4)).batch_id for i in range(40)]
    assert run(box)
    assert sorted(g[0] for g in got) == sorted(ids) and all(a == b for a, b, _ in got)
    per = collections.Counter(w for _, _, w in got)
    assert set(per) == {0, 1, 2, 3} and max(per.values()) - min(per.values()) <= 2
    assert len(cl.free) == 32, "every arena slot came back"
    assert cl.uploads == 40


def test_timeout_redispatches_and_bumps_epoch_before_requeue():
    cl = FakeClient()
    got = []
    box = ElasticBox(cl, on_result=lambda b, v, w: got.append((b.batch_id, w)))
    for w in range(3):
        box.add_worker(w)
    ids = [box.submit_host(object(), range(4)).batch_id for _ in range(12)]
    box.pump()
    cl.dead.add(1)                       # worker 1 goes silent with two batches in flight
    assert run(box)
    assert sorted(b for b, _ in got) == sorted(ids), "nothing lost, nothing twice"
    assert all(w != 1 for b, w in got if b in {bid for (ww, _, bid) in cl.log if ww == 1})
    assert box.failed_workers == [1] and 1 in cl.stopped
    assert len(cl.epochs[1]) == 2 and cl.epochs[1][1] > cl.epochs[1][0], "K14: stale deliveries of the dead worker are gated"
    assert box.stats["redispatched"] >= 1 and box.sched.stats["dropped_after_max_attempts"] == 0
    assert cl.uploads > 12, "a re-dispatched host batch is uploaded again on its new lane's stream"


def test_leave_and_rejoin_mid_stream():
    cl = FakeClient(latency=3)
    got = []
    box = ElasticBox(cl, on_result=lambda b, v, w: got.append((b.batch_id, w)))
    for w in range(4):
        box.add_worker(w)
    slots = [cl.alloc_slot() for _ in range(8)]
    ids = [box.submit_slot(slots[i % 8], range(4)).batch_id for i in range(60)]
    for i in range(5):
        box.pump()
    assert box.remove_worker(2) >= 1 and box.remove_worker(3) >= 1
    cl.dead.update({2, 3})               # what they still had in flight is rejected by the K9 epoch gate: never answers
    n_before = len(got)
    for i in range(30):
        box.pump()
    assert all(w in (0, 1) for _, w in got[n_before:])
    cl.dead.discard(2)
    box.add_worker(2)
    assert run(box)
    assert sorted(b for b, _ in got) == sorted(ids)
    assert any(w == 2 for _, w in got[-20:]), "the re-joined worker takes work again"
    assert cl.resets.count(2) == 2 and len(cl.free) == 32 - 8, "kept slots stay with their owner"


def test_waiting_for_a_lane_does_not_burn_attempts():
    cl = FakeClient(workers=1, lanes=1, latency=50)
    box = ElasticBox(cl, scheduler=Scheduler(max_attempts=3, requester_works=True))
    box.add_worker(0)
    box.submit_host(object(), range(4))
    box.pump()
    # the worker is declared gone and comes straight back while its old flight still occupies the lane
    box.remove_worker(0, reason="timeout")
    box.add_worker(0)
    for _ in range(40):
        box.pump()
    assert box.sched.stats["dropped_after_max_attempts"] == 0
    assert run(box) and box.stats["completed"] == 1


def test_undo_dispatch_returns_credit_and_attempt():
    s = Scheduler(max_attempts=2, requester_works=True)
    s.join("a", capacity=1)
    b = s.submit("r", [1])
    for _ in range(10):                       # far more hand-over failures than max_attempts
        d = s.next_dispatch()
        assert d is not None and d[0] is b
        assert s.undo_dispatch(b.batch_id) is b and b.attempts == 0
    assert s.stats["dropped_after_max_attempts"] == 0 and s.members.workers["a"].inflight == 0


def test_dropped_batches_are_reported():
    s = Scheduler(max_attempts=2, requester_works=True)
    dropped = []
    s.on_drop = dropped.append
    s.join("a")
    b = s.submit("r", [1, 2])
    for _ in range(2):
        s.next_dispatch()
        s.fail(b.batch_id)
    assert [d.batch_id for d in dropped] == [b.batch_id] and list(s.dropped)[0] is b


def test_box_layout_regions_do_not_overlap_and_are_aligned():
    """Every descriptor, done word, epoch word and result slot has its own bytes; descriptors are 8-byte aligned (the post
    kernel checks that), result slots 256-byte, arena slots 16-byte (K1a's vector loads)."""
    for (W_, Ln, B, H, Wd) in [(8, 4, 64, 300, 300), (2, 2, 4, 120, 160), (8, 3, 16, 720, 1280), (3, 1, 1, 1080, 1920), (5, 2, 7, 97, 399)]:
        lay = BoxLayout(W_, Ln, B, H, Wd, arena_slots=W_ * Ln + 4, weights_bytes=12345)
        spans = []
        for w in range(W_):
            spans.append((lay.epoch_addr(0, w), 4))
            for l in range(Ln):
                spans.append((lay.desc_addr(0, w, l), 32))
                spans.append((lay.done_addr(0, w, l), 4))
                assert lay.desc_addr(0, w, l) % 8 == 0
        spans.append((lay.weights_off, lay.weights_bytes))
        spans.sort()
        for (a, n), (b, _) in zip(spans, spans[1:]):
            assert a + n <= b, (spans,)
        assert spans[-1][0] + spans[-1][1] <= lay.ctl_bytes
        res = sorted(lay.res_addr(0, w, l) for w in range(W_) for l in range(Ln))
        assert all(b - a >= lay.res_slot_bytes for a, b in zip(res, res[1:])) and res[-1] + lay.res_slot_bytes <= lay.res_bytes
        assert all(r % 256 == 0 for r in res) and lay.flags_off + 4 * B <= lay.res_slot_bytes and lay.copy_bytes == lay.flags_off
        assert lay.raw_stride % 16 == 0 and lay.raw_stride >= H * Wd * 3 and lay.arena_off(3) == 3 * B * lay.raw_stride
        assert lay.arena_bytes == lay.arena_slots * lay.slot_bytes
