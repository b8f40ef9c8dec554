"""Elastic in-process cluster: a set of engines (one per GPU, or CPU engines) owned by the
scheduler; workers join and leave at will while jobs are running.

This is the in-box replacement of the reference's server loop + worker processes
(server.py:77-158, worker.py:202-280) for the case where all workers live in one box: no
sockets, no chunk packing; the scheduler hands whole batches to per-worker threads, keeps every
batch in the in-flight table until its result is in, and re-dispatches the batches of a worker
that leaves (or stops answering) - the behaviour the reference lacks (a job
hangs forever when a worker quits mid-stream).

Results are delivered per requester through a `ReorderBuffer`, so duplicates produced by a
re-dispatch race are dropped and frames come out in order.
"""
from __future__ import annotations

import queue
import threading
import time
from dataclasses import dataclass
from typing import Callable, Dict, List, Optional

import numpy as np

from ..config import Config
from ..engine import BatchResult
from ..utils.metrics import Metrics
from .scheduler import Batch, ReorderBuffer, Scheduler


@dataclass
class FrameResult:
    frame_id: int
    det: np.ndarray          # [n,7]
    boxes: np.ndarray        # [m,4] int (app-filtered)
    count: int
    frame400: Optional[np.ndarray]
    worker: str
    latency_s: float


class _Worker:
    def __init__(self, name: str, engine, cluster: "LocalCluster"):
        self.name = name
        self.engine = engine
        self.cluster = cluster
        self.q: "queue.Queue[Optional[Batch]]" = queue.Queue()
        self.alive = True
        self.hang = False            # fault injection: accept work but never answer
        self.thread = threading.Thread(target=self.loop, daemon=True, name=f"worker-{name}")
        self.thread.start()

    def loop(self):
        if hasattr(self.engine, "process_stream") and self.cluster.cfg.pipelined_worker:
            return self._loop_pipelined()
        return self._loop_blocking()

    def _loop_pipelined(self):
        """GPU engines: batches go through `NativeEngine.process_stream`, so the upload of batch i+1, the kernels of
        batch i and the read-back of batch i-1 overlap; an empty queue flushes the pipeline (a `None` item), so the last
        batch of a stream is answered at once.  Same bookkeeping as the blocking loop."""
        cl = self.cluster
        flying = []                       # batches handed to the engine, in order
        stop = [False]

        def batches():
            while True:
                try:
                    b = self.q.get(timeout=0.002 if flying else 0.1)
                except queue.Empty:
                    yield None
                    continue
                if b is None:
                    stop[0] = True
                    return
                if self.hang:
                    continue              # fault injection: the batch stays in flight until the watchdog re-dispatches it
                frames, t_submit, fetch = b.payload
                flying.append((b, t_submit))
                yield (frames, b.requester)

        try:
            for res in self.engine.process_stream(batches(), depth=2, annotate=cl.annotate, fetch_frames=True):
                b, t_submit = flying.pop(0)
                if not self.alive and cl.drop_results_of_departed:
                    continue
                if cl.sched.complete(b.batch_id, self.name) is None:
                    cl.metrics.inc("duplicate_batches")
                    continue
                # the yielded result aliases the slot's pinned buffers (valid until the next item): take copies
                f = lambda t: None if t is None else np.array(t.numpy() if hasattr(t, "numpy") else t)
                cl.deliver(b, BatchResult(f(res.det), f(res.det_count), f(res.boxes_i), f(res.box_count), f(res.frames400)),
                           self.name, t_submit)
        except BaseException as e:        # a failing worker must not lose work: it is retired, its batches move on
            cl.metrics.inc("worker_errors")
            cl.log(f"worker {self.name} failed: {e!r}")
            cl.last_error = repr(e)
            self.alive = False
            cl.sched.leave(self.name)

    def _loop_blocking(self):
        cl = self.cluster
        while True:
            b = self.q.get()
            if b is None:
                return
            if self.hang:
                continue              # the batch stays in flight until the watchdog re-dispatches it
            frames, t_submit, fetch = b.payload
            try:
                res = self.engine.process(frames, requester_id=b.requester, annotate=cl.annotate).numpy()
            except BaseException as e:    # a failing worker must not lose work: it is retired, its batches move on
                cl.metrics.inc("worker_errors")
                cl.log(f"worker {self.name} failed on batch {b.batch_id}: {e!r}")
                cl.last_error = repr(e)
                self.alive = False
                cl.sched.leave(self.name)
                return
            if not self.alive and cl.drop_results_of_departed:
                continue              # it left before finishing: the re-dispatched copy will deliver
            first = cl.sched.complete(b.batch_id, self.name)
            if first is None:
                cl.metrics.inc("duplicate_batches")
                continue
            cl.deliver(b, res, self.name, t_submit)


class LocalCluster:
    def __init__(self, cfg: Optional[Config] = None, scheduler: Optional[Scheduler] = None, annotate: bool = True,
                 on_frame: Optional[Callable[[str, FrameResult], None]] = None, log=None):
        self.cfg = cfg or Config()
        self.sched = scheduler or Scheduler(policy=self.cfg.dispatch_policy, requester_works=True,
                                            worker_timeout_s=self.cfg.worker_timeout_s)
        self.annotate = annotate
        self.metrics = Metrics(self.cfg.metrics_jsonl)
        self.workers: Dict[str, _Worker] = {}
        self.reorder: Dict[str, ReorderBuffer] = {}
        self.delivered: Dict[str, int] = {}
        self.on_frame = on_frame
        self.log = log or (lambda m: None)
        self.lock = threading.RLock()
        self.done_cv = threading.Condition(self.lock)
        self.drop_results_of_departed = False
        self.last_error: Optional[str] = None
        self.running = True
        self.dispatcher = threading.Thread(target=self._dispatch_loop, daemon=True, name="dispatcher")
        self.dispatcher.start()
        self.watchdog = threading.Thread(target=self._watchdog_loop, daemon=True, name="watchdog")
        self.watchdog.start()

    # ------------------------------------------------------------------ membership (at will, mid-stream)
    def join(self, name: str, engine, capacity: int = 2):
        with self.lock:
            old = self.workers.get(name)
            if old is not None and old.alive:
                return                     # idempotent
            self.workers[name] = _Worker(name, engine, self)
        self.sched.join(name, capacity=capacity)
        self.metrics.emit("join", worker=name, epoch=self.sched.members.epoch)

    def leave(self, name: str) -> int:
        """Worker `name` leaves now. Batches it owns are re-dispatched; returns how many."""
        with self.lock:
            w = self.workers.get(name)
            if w is None or not w.alive:
                return 0
            w.alive = False
        re = self.sched.leave(name)
        w.q.put(None)
        self.metrics.inc("redispatched_batches", len(re))
        self.metrics.emit("leave", worker=name, epoch=self.sched.members.epoch, redispatched=len(re))
        return len(re)

    def inject_stall(self, name: str, stalled: bool = True):
        """Fault injection: the worker keeps accepting batches but never answers."""
        self.workers[name].hang = stalled

    # ------------------------------------------------------------------ requesters
    def open(self, requester: str, first_frame: int = 1):
        """Start a job for `requester` (any name; several requesters share the pool with fair share)."""
        with self.lock:
            self.reorder[requester] = ReorderBuffer(first_frame)
            self.delivered[requester] = 0

    def submit(self, requester: str, frames, first_frame_id: int, fetch_frames: bool = True) -> List[Batch]:
        """Split `frames` ([n,H,W,3] uint8) into batches of cfg.batch and queue them."""
        if requester not in self.reorder:
            self.open(requester, first_frame_id)
        out = []
        n = len(frames)
        bs = self.cfg.batch
        for i in range(0, n, bs):
            ids = range(first_frame_id + i, first_frame_id + min(n, i + bs))
            out.append(self.sched.submit(requester, ids, payload=(frames[i:i + bs], time.monotonic(), fetch_frames)))
        self.metrics.inc("frames_in", n)
        return out

    def deliver(self, b: Batch, res, worker: str, t_submit: float):
        lat = time.monotonic() - t_submit
        self.metrics.observe("batch_latency_ms", lat * 1e3)
        with self.done_cv:
            rb = self.reorder.get(b.requester)
            if rb is None:
                return
            for k, fid in enumerate(b.frame_ids):
                nd, nb = int(res.det_count[k]), int(res.box_count[k])
                fr = FrameResult(fid, res.det[k, :nd].copy(), res.boxes_i[k, :nb].copy(), nb,
                                 res.frames400[k] if res.frames400 is not None else None, worker, lat)
                for n_, item in rb.push(fid, fr):
                    self.delivered[b.requester] += 1
                    if self.on_frame:
                        self.on_frame(b.requester, item)
            self.metrics.inc("frames_out", len(b.frame_ids))
            self.done_cv.notify_all()

    def wait(self, requester: str, n_frames: int, timeout: float = 60.0) -> bool:
        deadline = time.monotonic() + timeout
        with self.done_cv:
            while self.delivered.get(requester, 0) < n_frames:
                left = deadline - time.monotonic()
                if left <= 0:
                    return False
                self.done_cv.wait(min(left, 0.1))
        return True

    # ------------------------------------------------------------------ loops
    def _dispatch_loop(self):
        while self.running:
            d = self.sched.wait_dispatch(timeout=0.1)
            if d is None:
                continue
            batch, winfo, epoch = d
            w = self.workers.get(winfo.addr)
            if w is None or not w.alive:
                self.sched.fail(batch.batch_id, reason="gone")
                continue
            w.q.put(batch)

    def _watchdog_loop(self):
        while self.running:
            time.sleep(min(0.25, self.cfg.worker_timeout_s / 4))
            re = self.sched.check_timeouts()
            if re:
                self.metrics.inc("redispatched_batches", len(re))
                self.metrics.inc("timeouts")
                for name, w in list(self.workers.items()):
                    if not self.sched.members.workers.get(name, None) or not self.sched.members.workers[name].alive:
                        w.alive = False

    def close(self):
        self.running = False
        for w in self.workers.values():
            w.q.put(None)
        with self.sched.cv:
            self.sched.cv.notify_all()
