import random

from distributedvolunteercomputing_b200.parallel.scheduler import ReorderBuffer, Scheduler


class FakeClock:
    def __init__(self):
        self.t = 0.0

    def __call__(self):
        return self.t


def test_round_robin_excludes_requester_like_reference():
    s = Scheduler(policy="round_robin")
    for a in ("A", "B", "C"):
        s.join(a, capacity=100)
    s.become_requester("A")
    for i in range(6):
        s.submit("A", [i])
    owners = [s.next_dispatch()[1].addr for _ in range(6)]
    assert owners == ["B", "C", "B", "C", "B", "C"]
    assert s.next_dispatch() is None


def test_join_is_idempotent_and_work_waits_for_a_worker():
    s = Scheduler()
    s.join("R")
    assert s.join("R")[1] is False and len(s.members.alive()) == 1
    s.become_requester("R")
    b = s.submit("R", [1, 2, 3])
    assert s.next_dispatch() is None          # the reference would have dropped this chunk (server.py:82,88)
    s.join("W")
    d = s.next_dispatch()
    assert d is not None and d[0].batch_id == b.batch_id and d[1].addr == "W"


def test_capacity_credits_and_least_loaded():
    s = Scheduler()
    s.join("W1", capacity=1); s.join("W2", capacity=2); s.join("R"); s.become_requester("R")
    for i in range(5):
        s.submit("R", [i])
    got = [s.next_dispatch() for _ in range(5)]
    disp = [g for g in got if g is not None]
    assert len(disp) == 3                      # 1 + 2 credits
    assert sorted(w.addr for _, w, _ in disp) == ["W1", "W2", "W2"]
    s.complete(disp[0][0].batch_id)
    assert s.next_dispatch() is not None


def test_leave_requeues_inflight_in_order_and_dedups_results():
    s = Scheduler()
    s.join("W1", capacity=4); s.join("W2", capacity=4); s.join("R"); s.become_requester("R")
    ids = [s.submit("R", [i]).batch_id for i in range(6)]
    disp = [s.next_dispatch() for _ in range(6)]
    on_w1 = [b.batch_id for b, w, _ in disp if w.addr == "W1"]
    assert on_w1
    e0 = s.members.epoch
    re = s.leave("W1")
    assert s.members.epoch > e0
    assert [b.batch_id for b in re] == sorted(on_w1)
    # re-dispatched batches go out first, to the survivor
    nxt = s.next_dispatch()
    assert nxt[0].batch_id == sorted(on_w1)[0] and nxt[1].addr == "W2"
    # the old owner's late result is accepted once, the second copy is a harmless duplicate
    assert s.complete(nxt[0].batch_id, "W1") is not None
    assert s.complete(nxt[0].batch_id, "W2") is None
    assert s.stats["duplicate_results"] == 1


def test_timeout_redispatches_and_marks_worker_dead():
    clk = FakeClock()
    s = Scheduler(worker_timeout_s=5.0, clock=clk)
    s.join("W1"); s.join("W2"); s.join("R"); s.become_requester("R")
    s.submit("R", [1]); s.submit("R", [2])
    d1, d2 = s.next_dispatch(), s.next_dispatch()
    assert {d1[1].addr, d2[1].addr} == {"W1", "W2"}
    clk.t = 3.0
    s.complete(d2[0].batch_id, d2[1].addr)       # W? is alive and answering
    clk.t = 9.0
    re = s.check_timeouts()
    assert [b.batch_id for b in re] == [d1[0].batch_id]
    assert not s.members.workers[d1[1].addr].alive
    nd = s.next_dispatch()
    assert nd[1].addr == d2[1].addr and nd[0].attempts == 2


def test_multi_requester_fair_share():
    s = Scheduler()
    s.join("W", capacity=1000)
    for r in ("R1", "R2", "R3"):
        s.join(r); s.become_requester(r)
    for i in range(30):
        s.submit("R1", [i])
    for i in range(3):
        s.submit("R2", [i]); s.submit("R3", [i])
    first9 = [s.next_dispatch()[0].requester for _ in range(9)]
    assert first9.count("R1") == 3 and first9.count("R2") == 3 and first9.count("R3") == 3


def test_elastic_8_4_8_no_lost_frames_randomized():
    """Fake workers with random latencies; 4 of 8 leave mid-stream (with work in flight) and later
    re-join.  Every frame must be delivered exactly once through the reorder buffer."""
    rng = random.Random(0)
    s = Scheduler()
    workers = [f"gpu{i}" for i in range(8)]
    for w in workers:
        s.join(w, capacity=2)
    s.join("client"); s.become_requester("client")
    n_batches, bs = 200, 64
    for b in range(n_batches):
        s.submit("client", range(b * bs + 1, (b + 1) * bs + 1))
    rb = ReorderBuffer(1)
    delivered = 0
    running = {}            # batch_id -> (worker, finish_step, frames)
    zombies = []            # results from workers that left, arriving late
    step = 0
    left = set()
    while delivered < n_batches * bs:
        step += 1
        assert step < 100000
        while True:
            d = s.next_dispatch()
            if d is None:
                break
            b, w, ep = d
            running[b.batch_id] = (w.addr, step + rng.randint(1, 6), b.frame_ids)
        if step == 40:
            for w in workers[4:]:
                held = [bid for bid, (ww, _, _) in running.items() if ww == w]
                for bid in held:
                    zombies.append((step + 20, bid, w, running.pop(bid)[2]))
                s.leave(w); left.add(w)
        if step == 120:
            for w in workers[4:]:
                s.join(w, capacity=2); left.discard(w)
        for bid in [bid for bid, (_, t, _) in running.items() if t <= step]:
            w, _, frames = running.pop(bid)
            if s.complete(bid, w) is not None:
                for n in frames:
                    delivered += len(rb.push(n, None))
        for z in [z for z in zombies if z[0] <= step]:
            zombies.remove(z)
            if s.complete(z[1], z[2]) is not None:
                for n in z[3]:
                    delivered += len(rb.push(n, None))
    assert delivered == n_batches * bs and len(rb) == 0 and rb.next == n_batches * bs + 1
    assert s.stats["redispatched"] > 0 and s.outstanding() == 0


def test_late_result_cancels_redispatch_and_presumed_dead_worker_is_revived():
    """A slow worker is declared dead by the timeout and its batch re-queued; when its result arrives after all the
    result counts (once), the queued copy is dropped, and the worker is taken back.  A worker that said `end` is not."""
    t = [0.0]
    s = Scheduler(worker_timeout_s=5.0, clock=lambda: t[0])
    s.join("slow"); s.join("client"); s.become_requester("client")
    b = s.submit("client", [1, 2, 3])
    d = s.next_dispatch()
    assert d is not None and d[1].addr == "slow"
    t[0] = 6.0
    assert [x.batch_id for x in s.check_timeouts()] == [b.batch_id]
    assert s.pending() == 1 and not s.members.workers["slow"].alive and s.next_dispatch() is None
    done = s.complete(b.batch_id, "slow")                      # the late result
    assert done is not None and done.batch_id == b.batch_id and s.pending() == 0 and s.outstanding() == 0
    assert s.complete(b.batch_id, "slow") is None              # and only once
    assert s.stats["late_results"] == 1 and s.stats["duplicate_results"] == 1
    assert s.revive("slow") and s.members.workers["slow"].alive and not s.revive("slow")
    b2 = s.submit("client", [4])
    d2 = s.next_dispatch()
    assert d2 is not None and d2[1].addr == "slow" and s.complete(b2.batch_id, "slow") is not None
    s.leave("slow")                                            # voluntary `end`: must join again
    assert not s.revive("slow") and not s.members.workers["slow"].alive


def test_reorder_buffer():
    rb = ReorderBuffer(1)
    assert rb.push(2, "b") == []
    assert rb.push(1, "a") == [(1, "a"), (2, "b")]
    assert rb.push(2, "dup") == [] and rb.duplicates == 1
    assert rb.push(3, "c") == [(3, "c")]
