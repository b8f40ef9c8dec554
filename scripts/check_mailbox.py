"""python scripts/check_mailbox.py [--gpus N]: the elastic multi-GPU path on N GPUs of one box (no torchrun: the client
process owns GPU 0 and spawns one worker process per other GPU).

  1. scatter -> infer -> gather through the mailboxes equals the single-GPU engine bit for bit, on every worker
  2. a worker that leaves with a batch in flight has that batch rejected on its own GPU by K9's epoch gate (stale > 0),
     and the batch is delivered by someone else
  3. a worker that goes silent (fault injection) is found by the bounded device-side wait and its batches move on
  3b. a worker that pauses 5 ms after every batch only loses its own share: the others keep their rate (no lock-step)
  4. a worker process killed with SIGKILL mid-stream costs no frame: the job completes, bit-identical
Prints one "mailbox check ok" line per stage."""
import argparse
import os
import signal
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedvolunteercomputing_b200.config import Config
from distributedvolunteercomputing_b200.engine import NativeEngine
from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
from distributedvolunteercomputing_b200.parallel.mailbox import ClientBox, ElasticBox, WorkerLanes


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=torch.cuda.device_count())
    args = ap.parse_args()
    n = min(args.gpus, torch.cuda.device_count())
    assert n >= 2, "needs at least 2 GPUs"
    B, H, W = 8, 240, 320
    torch.cuda.set_device(0)
    graph = SSDGraph.random(seed=1, conf_gain=2.5)
    cfg = Config(batch=B, use_cuda_graph=True)
    eng = NativeEngine(graph, "cuda:0", cfg, requester_id="client")
    client = ClientBox("cuda:0", B, H, W, max_workers=n, lanes=2, engine=eng, done_timeout_s=0.5)
    client.attach_local(0, WorkerLanes(eng, client.layout, *client.bases(), worker=0))
    got = {}
    box = ElasticBox(client, on_result=lambda b, v, w: got.__setitem__(b.batch_id, ({k: a.copy() for k, a in v.items()}, w)),
                     log=lambda m: print("  [sched]", m, flush=True))
    box.add_worker(0)
    for w in range(1, n):
        box.spawn_worker(w, w, graph, cfg)
    rng = np.random.default_rng(7)
    pool = [torch.from_numpy(rng.integers(0, 256, (B, H, W, 3), dtype=np.uint8)).pin_memory() for _ in range(6)]
    refs = [eng.process(f).numpy() for f in pool]

    def retire(w):
        """Reap worker process w (it may already have returned from serve() and be gone): its last message, if any."""
        p, conn = box.procs.pop(w)
        try:
            conn.send("stop")
        except (BrokenPipeError, OSError):
            pass
        msg = None
        try:
            while conn.poll(20.0 if msg is None else 0.0):
                msg = conn.recv()
        except (EOFError, OSError):
            pass
        p.join(30.0)
        if p.is_alive():
            p.kill()
        return msg

    def check(ids, what):
        per = {}
        for k, bid in enumerate(ids):
            g, w = got[bid]
            r = refs[k % len(pool)]
            assert np.array_equal(g["det"], r.det), (what, bid, w, "det")
            assert np.array_equal(g["det_count"], r.det_count) and np.array_equal(g["box_count"], r.box_count), (what, bid, w)
            for i in range(B):
                assert np.array_equal(g["boxes_i"][i, :r.box_count[i]], r.boxes_i[i, :r.box_count[i]]), (what, bid, w, "boxes")
            per[w] = per.get(w, 0) + 1
        return per

    # ---- 1. bit identity on every worker
    ids = [box.submit_host(pool[k % len(pool)], range(B)).batch_id for k in range(12 * n)]
    assert box.drain(60.0), box.stats
    per = check(ids, "identity")
    assert set(per) == set(range(n)), f"some worker never served: {per}"
    assert sum(int(r.det_count.sum()) for r in refs) > 0
    print(f"mailbox check ok: identity on {n} GPUs, batches per worker {dict(sorted(per.items()))}, stats {dict(box.stats)}", flush=True)

    # ---- 2. leave with a batch in flight: K9 on the leaver rejects it (stale), someone else delivers it
    stale_seen = 0
    for attempt in range(5):
        victim = n - 1
        ids = [box.submit_host(pool[k % len(pool)], range(B)).batch_id for k in range(4 * n)]
        box.pump()                                    # everything that fits is now in flight, also on the victim
        box.remove_worker(victim, reason="leave", stop=True)
        assert box.drain(60.0), box.stats
        check(ids, "leave")
        msg = retire(victim)
        stale_seen = msg[2] if msg else -1
        box.spawn_worker(victim, victim, graph, cfg)
        if stale_seen > 0:
            break
    assert stale_seen > 0, "the leaver's in-flight batch was never observed as stale"
    print(f"mailbox check ok: leave mid-batch, {stale_seen} stale frames rejected on gpu{victim}, all batches delivered once", flush=True)

    # ---- 3. silent worker (stalls after 3 batches): bounded wait -> time-out -> re-dispatch
    victim = 1
    box.remove_worker(victim)
    retire(victim)
    box.spawn_worker(victim, victim, graph, cfg, stall_after=3)
    f0 = len(box.failed_workers)
    ids = [box.submit_host(pool[k % len(pool)], range(B)).batch_id for k in range(16 * n)]
    t0 = time.monotonic()
    assert box.drain(60.0), box.stats
    check(ids, "stall")
    assert box.failed_workers[f0:] == [victim], box.failed_workers
    print(f"mailbox check ok: silent worker gpu{victim} found by the device-side time-out after {time.monotonic() - t0:.2f} s, "
          f"{box.stats['redispatched']} batches re-dispatched so far", flush=True)
    retire(victim)
    box.spawn_worker(victim, victim, graph, cfg)

    # ---- 3b. one slow worker (pauses 5 ms after every batch) must not slow the others down
    def window(seconds):
        counts = {}
        box.on_result = lambda b, v, w: counts.__setitem__(w, counts.get(w, 0) + 1)
        t_end = time.monotonic() + seconds
        while time.monotonic() < t_end:
            while box.sched.pending() < 4 * n:
                box.submit_host(pool[0], range(B))
            box.pump()
        assert box.drain(60.0)
        return counts

    window(0.3)
    normal = window(1.0)
    victim = n - 1
    box.remove_worker(victim)
    retire(victim)
    box.spawn_workers([(victim, victim)], graph, cfg, delay_ms=5.0)
    slow = window(1.0)
    others_normal = sum(c for w, c in normal.items() if w != victim)
    others_slow = sum(c for w, c in slow.items() if w != victim)
    assert slow.get(victim, 0) < 0.6 * normal[victim], (normal, slow)
    assert others_slow >= 0.9 * others_normal, f"the other workers stalled behind the slow one: {normal} -> {slow}"
    print(f"mailbox check ok: gpu{victim} delayed 5 ms per batch serves {slow.get(victim, 0)} instead of {normal[victim]} batches/s; "
          f"the other workers {others_normal} -> {others_slow} batches/s (no lock-step)", flush=True)
    box.on_result = lambda b, v, w: got.__setitem__(b.batch_id, ({k: a.copy() for k, a in v.items()}, w))
    box.remove_worker(victim)
    retire(victim)
    box.spawn_worker(victim, victim, graph, cfg)

    # ---- 4. SIGKILL mid-stream
    victim = n - 1
    ids = [box.submit_host(pool[k % len(pool)], range(B)).batch_id for k in range(20 * n)]
    for _ in range(200):
        box.pump()
    p, conn = box.procs.pop(victim)
    os.kill(p.pid, signal.SIGKILL)
    f0 = len(box.failed_workers)
    k0 = len(ids)
    while victim in box.workers_alive() and len(ids) < k0 + 400 * n:      # the scheduler still believes in the victim:
        ids.append(box.submit_host(pool[len(ids) % len(pool)], range(B)).batch_id)    # keep the stream going until it is found
        box.pump()
    assert box.drain(60.0), box.stats
    per = check(ids, "kill")
    assert victim in box.failed_workers[f0:], "the killed worker was never declared dead"
    print(f"mailbox check ok: gpu{victim} killed with SIGKILL mid-stream, {len(ids)} batches all delivered once and bit-identical; "
          f"per worker {dict(sorted(per.items()))}", flush=True)
    box.shutdown()
    client.close()
    print("mailbox check done", flush=True)


if __name__ == "__main__":
    main()
