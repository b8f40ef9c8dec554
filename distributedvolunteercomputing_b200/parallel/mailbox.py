"""Mailbox data plane: one requesting client GPU, any number of worker GPUs that come and go.

This is the in-box replacement of the reference's server relay (server.py:40-91: per-client ingest thread, FIFO
`requests`, round-robin `send_request`, `send_reply`) and of its lock-step `b'OK'` flow control (server.py:62,
worker.py:180).  The protocol is written down in csrc/mailbox.cu and DESIGN.md §4; in short:

  * The CLIENT owns and exports three device buffers: a control block (per-worker epoch words, one 32-byte
    descriptor and one done word per (worker, lane), the packed weights), the result slots (one per (worker, lane):
    detection rows, integer boxes, counts) and the frame arena (slots of B raw frames).
  * A dispatch = the client posts a descriptor (st.release.sys).  The worker lane's wait kernel polls it over NVLink,
    K1a pulls the frames with peer loads (scatter), the network runs, K9 pushes the rows with peer stores and its
    last CTA publishes the done word (gather).  The client's wait kernel polls the done word and the stream goes on
    to the device->host copy of the slot.  No collective, no host round trip, every wait bounded by a time-out.
  * Workers are independent processes: started with torchrun or spawned at any time (the client's own GPU serves
    inline, `ClientBox.attach_local`).  None of them takes part in any collective, so one that stalls, leaves or is killed
    only costs its own in-flight batches, which `ElasticBox` re-dispatches (scheduler.py).

`ClientBox`     client-side memory + dispatch / poll primitives (thin over the native calls)
`WorkerLanes`   worker-side lanes (engine states bound to the mailboxes) + the serving loop
`ElasticBox`    the PyTorch-level scheduler driving a ClientBox: join / leave / fail / re-dispatch / results
"""
from __future__ import annotations

import collections
import ctypes as C
import dataclasses
import time
from typing import Any, Callable, Deque, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..ops import lib as L
from .scheduler import Batch, Scheduler

HS_DTYPE = np.dtype([("status", "<u4"), ("seq", "<u4"), ("batch_id", "<u4"), ("n_valid", "<u4"), ("wait_ns", "<u8")])
DESC_BYTES = C.sizeof(L.MbxDesc)
CTL_STRIDE = 64                     # one lane control block per 64 bytes (sizeof(dvc_mbx_lane_ctl) = 40)
CTL_NEXT_SEQ_OFF = L.MbxLaneCtl.next_seq.offset
assert DESC_BYTES == 32 and C.sizeof(L.MbxLaneCtl) <= CTL_STRIDE and HS_DTYPE.itemsize == C.sizeof(L.MbxHostStatus)
assert L.MbxLaneCtl.gate_now.offset == 8 and L.MbxLaneCtl.gate_expect.offset == 12


def _up(n: int, a: int = 256) -> int:
    return (n + a - 1) // a * a


@dataclasses.dataclass
class BoxLayout:
    """Where everything lives inside the client's three exported buffers.  Plain integers: travels to workers as is."""
    max_workers: int
    lanes: int
    B: int
    H: int
    W: int
    arena_slots: int
    weights_bytes: int = 0

    def __post_init__(self):
        self.raw_stride = (self.H * self.W * 3 + 15) & ~15
        self.slot_bytes = self.B * self.raw_stride
        self.arena_bytes = self.arena_slots * self.slot_bytes
        n = self.max_workers * self.lanes
        self.epoch_off = 0
        self.desc_off = _up(4 * self.max_workers)
        self.done_off = _up(self.desc_off + n * DESC_BYTES)
        self.weights_off = _up(self.done_off + 4 * n)
        self.ctl_bytes = _up(self.weights_off + self.weights_bytes)
        B = self.B
        self.det_off = 0
        self.boxes_off = self.det_off + B * 700 * 4
        self.det_count_off = self.boxes_off + B * 400 * 4
        self.box_count_off = self.det_count_off + B * 4
        self.copy_bytes = self.box_count_off + B * 4          # what the client reads back per completed batch
        self.flags_off = self.copy_bytes
        self.res_slot_bytes = _up(self.flags_off + B * 4)
        self.res_bytes = n * self.res_slot_bytes

    def epoch_addr(self, base: int, w: int) -> int:
        return base + self.epoch_off + 4 * w

    def desc_addr(self, base: int, w: int, l: int) -> int:
        return base + self.desc_off + (w * self.lanes + l) * DESC_BYTES

    def done_addr(self, base: int, w: int, l: int) -> int:
        return base + self.done_off + (w * self.lanes + l) * 4

    def res_addr(self, base: int, w: int, l: int) -> int:
        return base + (w * self.lanes + l) * self.res_slot_bytes

    def arena_off(self, slot: int) -> int:
        return slot * self.slot_bytes


def pinned_status(n: int):
    """n dvc_mbx_host_status entries in pinned host memory: (keep-alive tensor, structured numpy view, address).
    Pinned memory is device-accessible at the same address (unified addressing), so kernels write it directly."""
    t = torch.zeros(n * HS_DTYPE.itemsize, dtype=torch.uint8).pin_memory()
    return t, t.numpy().view(HS_DTYPE), t.data_ptr()


# ====================================================================================== worker side
class WorkerLanes:
    """The lanes of ONE worker GPU: per lane an engine state whose plan is bound to the mailbox (K1b gate, K9 sinks in
    the client's result slot, done word), a control block, a stream.  `serve()` is the whole life of a worker."""

    def __init__(self, engine, layout: BoxLayout, ctl_base: int, res_base: int, arena_base: int, worker: int,
                 annotate: bool = True, adopt_weights: bool = False):
        from . import p2p
        self.engine, self.layout, self.worker = engine, layout, worker
        self.ctl_base, self.res_base, self.arena_base = ctl_base, res_base, arena_base
        self.annotate = annotate
        self.dev = engine.device
        lay = layout
        nl = lay.lanes
        with torch.cuda.device(self.dev):
            L.set_device(self.dev)
            if adopt_weights and lay.weights_bytes:
                # K13: the weights come from the client's exported blob (reference: every client reads the caffemodel
                # from its own disk, worker.py:194); one peer read at join, off the hot path
                blob = p2p.tensor_from_ptr(ctl_base + lay.weights_off, (lay.weights_bytes,), torch.uint8, self.dev)
                engine.import_weights(blob)
                torch.cuda.synchronize(self.dev)
            init = L.MbxLaneCtl(0, 0, 0xFFFFFFFF, 0, L.MBX_NOTHING, 0, 0, 1, 0)       # nothing seen yet: gates closed
            raw = np.zeros((nl, CTL_STRIDE), np.uint8)
            raw[:, :C.sizeof(init)] = np.frombuffer(bytes(init), np.uint8)
            self.ctl = torch.from_numpy(raw).to(self.dev)
            self.done_counter = torch.zeros(nl, dtype=torch.int32, device=self.dev)
            self.stale = torch.zeros(1, dtype=torch.int32, device=self.dev)
            self.local_raw = torch.zeros((nl, lay.slot_bytes), dtype=torch.uint8, device=self.dev)   # kind-1 frames
            self.streams = [torch.cuda.Stream(device=self.dev) for _ in range(nl)]
            self.depth = 2
            self._hs_keep, self.hs, self.hs_addr = pinned_status(nl * 4)
            self.states = []
            self.graphs: Dict[int, Any] = {}
            self.my_boxes, self.my_box_count = [], []
            for l in range(nl):
                st = engine.state(lay.B, lay.H, lay.W, slot=300 + worker * 16 + l)
                res = lay.res_addr(res_base, worker, l)
                engine.bind_remote(st, frames_ptr=L.ptr(st.frames400_flat), frame_stride=st.stride400,
                                   det_ptr=res + lay.det_off, det_count_ptr=res + lay.det_count_off,
                                   boxes_ptr=res + lay.boxes_off, box_count_ptr=res + lay.box_count_off,
                                   flags_ptr=res + lay.flags_off, flag_value=0)
                engine.bind_mailbox(st, self.ctl_ptr(l), lay.epoch_addr(ctl_base, worker), lay.done_addr(ctl_base, worker, l),
                                    self.done_counter[l:l + 1], self.stale)
                self.states.append(st)
                self.my_boxes.append(p2p.tensor_from_ptr(res + lay.boxes_off, (lay.B, 100, 4), torch.int32, self.dev))
                self.my_box_count.append(p2p.tensor_from_ptr(res + lay.box_count_off, (lay.B,), torch.int32, self.dev))
            # build every lane's CUDA graph now (gates closed: nothing is pulled, nothing is delivered)
            for l in range(nl):
                with torch.cuda.stream(self.streams[l]):
                    self.enqueue_body(l)
            torch.cuda.synchronize(self.dev)
        self.served = 0
        self.idle_polls = 0
        self.busy_ms = 0.0

    # ------------------------------------------------------------------ addresses
    def ctl_ptr(self, l: int) -> int:
        return self.ctl.data_ptr() + l * CTL_STRIDE

    def frames400(self, l: int) -> torch.Tensor:
        """The annotated 400-wide frames of the lane's last batch (they stay on the GPU that produced them)."""
        return self.states[l].frames400

    def stale_count(self) -> int:
        return int(self.stale.item())

    # ------------------------------------------------------------------ enqueue (current stream)
    def enqueue_wait(self, l: int, hs_index: Optional[int], expect_seq: int = 0, timeout_s: float = 0.002):
        lay = self.layout
        hs = self.hs_addr + hs_index * HS_DTYPE.itemsize if hs_index is not None else None
        L.check(self.engine.lib.dvc_mbx_wait_desc(lay.desc_addr(self.ctl_base, self.worker, l), expect_seq, self.arena_base,
                                                  self.local_raw[l].data_ptr(), self.ctl_ptr(l), hs, int(timeout_s * 1e9),
                                                  L.stream_ptr()), "mbx_wait_desc")

    def enqueue_body(self, l: int):
        """K1a (peer pull of the batch named by the lane's control block) -> network -> K9 (peer push + done word) ->
        K10 overlay of this GPU's own 400-wide frames.  With CUDA graphs the whole chain is ONE replay: every pointer in
        it is fixed (the source of K1a is read from the lane's control block at run time)."""
        if not self.engine.cfg.use_cuda_graph:
            return self._enqueue_body_eager(l)
        g = self.graphs.get(l)
        if g is None:
            from ..engine import _CAPTURE_LOCK
            with _CAPTURE_LOCK:
                self._enqueue_body_eager(l)                  # warm-up outside capture (first-launch attribute set-up)
                torch.cuda.synchronize(self.dev)
                g = torch.cuda.CUDAGraph()
                cap = torch.cuda.Stream(device=self.dev)
                with torch.cuda.graph(g, stream=cap, capture_error_mode="thread_local"):
                    self._enqueue_body_eager(l)
            self.graphs[l] = g
        g.replay()

    def graph_exec(self, l: int) -> Optional[int]:
        """The lane's instantiated CUDA graph (cudaGraphExec_t) for native launches; None when graphs are off."""
        if not self.engine.cfg.use_cuda_graph:
            return None
        if l not in self.graphs:
            with torch.cuda.device(self.dev), torch.cuda.stream(self.streams[l]):
                self.enqueue_body(l)
            torch.cuda.synchronize(self.dev)
        return int(self.graphs[l].raw_cuda_graph_exec())

    def _enqueue_body_eager(self, l: int):
        from ..ops import kernels as K
        st, lay, cp = self.states[l], self.layout, self.ctl_ptr(l)
        K.resize_u8_indirect(cp, lay.raw_stride, lay.B, st.t_area, st.frames400_flat, st.stride400, cp + 8, cp + 12)
        self.engine.run(st, annotate=False, resize=False, use_graph=False)
        if self.annotate:
            m = st.masks
            nl_, lh, lw = m.label_masks.shape
            idh, idw = m.id_mask.shape
            L.check(self.engine.lib.dvc_annotate(L.ptr(st.frames400_flat), st.stride400, None, lay.B, st.h4, st.w4,
                                                 L.ptr(self.my_boxes[l]), L.ptr(self.my_box_count[l]), None, L.ptr(m.id_mask),
                                                 m.id_xy[0], m.id_xy[1], idw, idh, L.ptr(m.label_masks), nl_, m.label_xy[0],
                                                 m.label_xy[1], lw, lh, L.stream_ptr()), "annotate")

    def launches_per_batch(self) -> int:
        return self.states[0].n_ops + 2 + 1 + 1 + (1 if self.annotate else 0)      # plan (detection = 3) + wait + K1a + K10

    # ------------------------------------------------------------------ the serving loop
    def sync_position(self) -> bool:
        """Adopt the lanes' current sequence numbers as the baseline (a worker may join long after the client started).
        Returns False when the client has already told this worker to stop."""
        from . import p2p
        lay = self.layout
        with torch.cuda.device(self.dev):
            for l in range(lay.lanes):
                d = p2p.tensor_from_ptr(lay.desc_addr(self.ctl_base, self.worker, l), (DESC_BYTES // 4,), torch.int32, self.dev)
                seq = int(d[0].item()) & 0xFFFFFFFF
                if seq == L.MBX_SEQ_STOP:
                    return False
                nxt = torch.tensor([seq + 1 if seq + 1 < L.MBX_SEQ_STOP else 1], dtype=torch.int64).to(torch.int32)
                self.ctl[l, CTL_NEXT_SEQ_OFF:CTL_NEXT_SEQ_OFF + 4].copy_(nxt.view(torch.uint8))
            torch.cuda.synchronize(self.dev)
        return True

    def serve(self, should_stop: Callable[[], bool] = lambda: False, max_batches: Optional[int] = None,
              stream_timeout_s: float = 0.003, probe_timeout_s: float = 0.0005, idle_sleep_s: float = 0.0002,
              local_feed: Optional[Callable[[int], torch.Tensor]] = None, expect_work: bool = False,
              on_ready: Optional[Callable[[], None]] = None, synced: bool = False, throttle_s: float = 0.0) -> int:
        """Serve dispatches until the client posts stop, `should_stop()` turns true or `max_batches` were served.

        While batches keep arriving every lane has `depth` (wait + network) pairs queued on its stream, so the GPU never
        waits for this loop; the sequence position lives in the lane's control block on the device.  When a wait times
        out the lane falls back to short probes (a wait kernel alone) so an idle worker does no useless work.
        `local_feed(i)` (optional, sharded ingest) returns the pinned host batch this worker uploads itself before the
        i-th queued batch; the client's descriptor (kind 1) is then only the permission to deliver.
        `expect_work` starts the lanes in streaming mode (a phase the client opens with a full queue).
        `on_ready()` is called once the lanes' sequence position has been read: only then may the client add this worker."""
        # the lanes' position must be read BEFORE the client may dispatch to this worker (`on_ready` tells it so, or the
        # caller did sync_position() itself ahead of a barrier): a descriptor posted in between would be taken for the
        # baseline and its batch never served
        if not synced and not self.sync_position():
            return 0
        if on_ready is not None:
            on_ready()
        lay, dev = self.layout, self.dev
        nl = lay.lanes
        queues: List[Deque[Tuple[str, int, torch.cuda.Event]]] = [collections.deque() for _ in range(nl)]
        free_hs: List[List[int]] = [list(range(l * 4, l * 4 + 4)) for l in range(nl)]
        mode = ["stream" if expect_work else "idle"] * nl      # expect_work: the client dispatches at once, skip the probes
        need_body = [False] * nl
        stopping = False
        fed = 0
        served0 = self.served
        self._first_ev = self._last_ev = None

        gexec = [self.graph_exec(l) for l in range(nl)]
        lib = self.engine.lib
        desc_addr = [lay.desc_addr(self.ctl_base, self.worker, l) for l in range(nl)]
        local_ptr = [self.local_raw[l].data_ptr() for l in range(nl)]
        sptr = [st_.cuda_stream for st_ in self.streams]
        hs_sz = HS_DTYPE.itemsize

        def push(l: int, kind: str):
            nonlocal fed
            idx = free_hs[l].pop() if kind != "body" else -1
            if gexec[l] is not None and local_feed is None and (kind != "body" or self._first_ev is not None):
                # hot path: [wait] -> [graph] -> event in ONE native call on the lane's stream
                ev = torch.cuda.Event(enable_timing=True)
                ev.record(self.streams[l])                       # materialise the event handle (cheap), re-recorded below
                if kind != "body":
                    self.hs[idx]["status"] = 0xFFFF
                    if kind == "full" and self._first_ev is None:
                        # first batch of this serve(): the busy span starts after its wait (two calls, once)
                        L.check(lib.dvc_mbx_worker_enqueue(desc_addr[l], 0, self.arena_base, local_ptr[l], self.ctl_ptr(l),
                                                           self.hs_addr + idx * hs_sz, int(stream_timeout_s * 1e9), None, None,
                                                           sptr[l]), "mbx_worker_enqueue")
                        self._first_ev = torch.cuda.Event(enable_timing=True)
                        self._first_ev.record(self.streams[l])
                        L.check(lib.dvc_mbx_worker_enqueue(None, 0, None, None, None, None, 0, gexec[l], ev.cuda_event, sptr[l]),
                                "mbx_worker_enqueue")
                    else:
                        tmo = int((stream_timeout_s if kind == "full" else probe_timeout_s) * 1e9)
                        L.check(lib.dvc_mbx_worker_enqueue(desc_addr[l], 0, self.arena_base, local_ptr[l], self.ctl_ptr(l),
                                                           self.hs_addr + idx * hs_sz, tmo, gexec[l] if kind == "full" else None,
                                                           ev.cuda_event, sptr[l]), "mbx_worker_enqueue")
                else:
                    L.check(lib.dvc_mbx_worker_enqueue(None, 0, None, None, None, None, 0, gexec[l], ev.cuda_event, sptr[l]),
                            "mbx_worker_enqueue")
                queues[l].append((kind, idx, ev))
                return
            with torch.cuda.stream(self.streams[l]):
                if kind != "body":
                    if local_feed is not None:
                        src = local_feed(fed)
                        fed += 1
                        self.local_raw[l].view(lay.B, lay.raw_stride)[:, : lay.H * lay.W * 3].copy_(
                            src.reshape(lay.B, -1), non_blocking=True)
                    self.hs[idx]["status"] = 0xFFFF
                    self.enqueue_wait(l, idx, 0, stream_timeout_s if kind == "full" else probe_timeout_s)
                if kind != "probe" and self._first_ev is None:      # after the wait, before the batch's first kernel
                    self._first_ev = torch.cuda.Event(enable_timing=True)      # the first real batch of this serve() starts here
                    self._first_ev.record()
                if kind != "probe":
                    self.enqueue_body(l)
                ev = torch.cuda.Event(enable_timing=True)
                ev.record()
            queues[l].append((kind, idx, ev))

        def budget_left() -> bool:
            if max_batches is None:
                return True
            queued = sum(1 for q in queues for e in q if e[0] != "probe") + sum(need_body)
            return self.served - served0 + queued < max_batches

        with torch.cuda.device(dev):
            L.set_device(dev)
            while True:
                progressed = False
                for l in range(nl):
                    q = queues[l]
                    while q and q[0][2].query():
                        kind, idx, ev_done = q.popleft()
                        progressed = True
                        if kind == "body":
                            self.served += 1
                            self._last_ev = ev_done
                            continue
                        status = int(self.hs[idx]["status"])
                        free_hs[l].append(idx)
                        if status == L.MBX_STOP:
                            stopping = True
                        elif status == L.MBX_GO:
                            mode[l] = "stream"
                            if kind == "probe":
                                need_body[l] = True          # the probe consumed the descriptor: run its batch now
                            else:
                                self.served += 1
                                self._last_ev = ev_done
                                if throttle_s:
                                    time.sleep(throttle_s)       # fault injection: a slow volunteer
                        else:
                            mode[l] = "idle"
                            self.idle_polls += 1
                            if self._last_ev is None:
                                self._first_ev = None        # nothing served yet: the busy span has not started
                    if need_body[l]:
                        push(l, "body")
                        need_body[l] = False
                        continue
                    if stopping or should_stop() or not budget_left():
                        continue
                    if mode[l] == "stream":
                        while len(q) < self.depth and budget_left():
                            push(l, "full")
                    elif not q:
                        push(l, "probe")
                done = stopping or should_stop() or not budget_left()
                if done and not any(queues) and not any(need_body):
                    break
                if not progressed:
                    time.sleep(idle_sleep_s if all(m == "idle" for m in mode) else 0.00002)
            torch.cuda.synchronize(dev)
            # device time from the start of the first batch served to the end of the last one (CUDA events on this GPU)
            self.busy_ms = self._first_ev.elapsed_time(self._last_ev) if self._first_ev is not None and self._last_ev is not None else 0.0
        return self.served - served0


# ====================================================================================== client side
class ClientBox:
    """Client-side memory and primitives.  One instance per requesting client GPU."""

    def __init__(self, device, B: int, H: int, W: int, max_workers: int = 8, lanes: int = 2, arena_slots: Optional[int] = None,
                 engine=None, done_timeout_s: float = 1.0):
        from . import p2p
        self.dev = torch.device(device)
        self.lib = L.load()
        wbytes = 0
        if engine is not None:
            _, wbytes = engine.weight_layout()
        slots = arena_slots if arena_slots is not None else max_workers * lanes + 4
        self.layout = lay = BoxLayout(max_workers, lanes, B, H, W, slots, wbytes)
        self.done_timeout_s = done_timeout_s
        with torch.cuda.device(self.dev):
            L.set_device(self.dev)
            self.ctl = p2p.DevBuffer(lay.ctl_bytes, self.dev)
            self.res = p2p.DevBuffer(lay.res_bytes, self.dev)
            self.arena = p2p.DevBuffer(lay.arena_bytes, self.dev)
            if engine is not None and wbytes:
                engine.export_weights(self.ctl.tensor(torch.uint8, (wbytes,), offset_bytes=lay.weights_off))
            n = max_workers * lanes
            self._hs_keep, self.hs, self.hs_addr = pinned_status(n)
            self.h_res = torch.empty((n, lay.copy_bytes), dtype=torch.uint8).pin_memory()
            self.streams = [torch.cuda.Stream(device=self.dev) for _ in range(n)]
            self.events = [torch.cuda.Event() for _ in range(n)]
            for k in range(n):                 # materialise the lazily created CUDA events
                self.events[k].record(self.streams[k])
            self.ctrl_stream = torch.cuda.Stream(device=self.dev)
            torch.cuda.synchronize(self.dev)
        self.seq = [[0] * lanes for _ in range(max_workers)]
        self.epoch = [0] * max_workers
        self.busy = [[False] * lanes for _ in range(max_workers)]
        self.free_slots = list(range(slots - 1, -1, -1))
        self._hs_status = self.hs["status"]
        self._cache = []
        for k in range(max_workers * lanes):
            w, l = divmod(k, lanes)
            d = L.MbxDesc()
            self._cache.append({"desc": d, "desc_ref": C.byref(d), "desc_addr": lay.desc_addr(self.ctl.ptr, w, l),
                                "done_addr": lay.done_addr(self.ctl.ptr, w, l), "hs": self.hs_addr + k * HS_DTYPE.itemsize,
                                "res": lay.res_addr(self.res.ptr, w, l), "h_res": self.h_res[k].data_ptr(),
                                "event": self.events[k].cuda_event, "stream": self.streams[k].cuda_stream,
                                "desc_addr_arr": (C.c_void_p * 1)(lay.desc_addr(self.ctl.ptr, w, l)),
                                "done_addr_arr": (C.c_void_p * 1)(lay.done_addr(self.ctl.ptr, w, l)),
                                "seq_arr": (C.c_uint32 * 1)(0)})
        self._local_gexec: Dict[Tuple[int, int], int] = {}
        n_ev = max_workers * lanes
        self._ev_arr = (C.c_void_p * n_ev)(*[self.events[k].cuda_event for k in range(n_ev)])
        self._busy_arr = (C.c_ubyte * n_ev)()
        self._done_arr = (C.c_ubyte * n_ev)()
        self.local: Dict[int, WorkerLanes] = {}        # workers living on the client's own GPU (same process)
        self._views: Dict[int, Dict[str, np.ndarray]] = {}
        self.dispatches = 0

    # ------------------------------------------------------------------ export / arena
    def export(self) -> Dict[str, Any]:
        """What a worker process needs to map this box (CUDA IPC handles + the layout)."""
        return {"ctl": self.ctl.handle(), "res": self.res.handle(), "arena": self.arena.handle(),
                "layout": dataclasses.asdict(self.layout), "client_device": self.dev.index or 0}

    def bases(self) -> Tuple[int, int, int]:
        return self.ctl.ptr, self.res.ptr, self.arena.ptr

    def alloc_slot(self) -> Optional[int]:
        return self.free_slots.pop() if self.free_slots else None

    def free_slot(self, slot: int):
        self.free_slots.append(slot)

    def slot_tensor(self, slot: int) -> torch.Tensor:
        """The arena slot as a uint8 [B, H, W, 3] device tensor (frame pitch = layout.raw_stride)."""
        lay = self.layout
        flat = self.arena.tensor(torch.uint8, (lay.B, lay.raw_stride), offset_bytes=lay.arena_off(slot))
        return torch.as_strided(flat, (lay.B, lay.H, lay.W, 3), (lay.raw_stride, lay.W * 3, 3, 1))

    def attach_local(self, worker: int, lanes: WorkerLanes):
        self.local[worker] = lanes

    # ------------------------------------------------------------------ membership words
    def set_epoch(self, worker: int, epoch: int):
        """K14: the worker's epoch word (client-local, release.sys): batches dispatched under an older value that
        reach their K9 after this are rejected there (and counted)."""
        self.epoch[worker] = int(epoch) & 0xFFFFFFFF
        with torch.cuda.device(self.dev):
            L.set_device(self.dev)
            L.check(self.lib.dvc_mbx_signal(self.layout.epoch_addr(self.ctl.ptr, worker), self.epoch[worker],
                                            self.ctrl_stream.cuda_stream), "mbx_signal")
            self.ctrl_stream.synchronize()

    def _post_all_lanes(self, worker: int, seqs: Sequence[int]):
        lay = self.layout
        n = lay.lanes
        ptrs = (C.c_void_p * n)(*[lay.desc_addr(self.ctl.ptr, worker, l) for l in range(n)])
        descs = (L.MbxDesc * n)(*[L.MbxDesc(int(seqs[l]) & 0xFFFFFFFF, self.epoch[worker], 0, 0, 0, 0, 0) for l in range(n)])
        with torch.cuda.device(self.dev):
            L.set_device(self.dev)
            L.check(self.lib.dvc_mbx_post(ptrs, descs, n, self.ctrl_stream.cuda_stream), "mbx_post")
            self.ctrl_stream.synchronize()

    def cancel(self, worker: int, lane: int):
        """Release the lane's pending done-wait now (the worker is gone, nobody will publish the word): the client writes
        the awaited sequence number itself.  The caller ignores what the lane's poll then reports."""
        if not self.busy[worker][lane]:
            return
        with torch.cuda.device(self.dev):
            L.set_device(self.dev)
            L.check(self.lib.dvc_mbx_signal(self.layout.done_addr(self.ctl.ptr, worker, lane), self.seq[worker][lane],
                                            self.ctrl_stream.cuda_stream), "mbx_signal")

    def stop_worker(self, worker: int):
        """Tell the worker's serving loop to return (every lane's descriptor gets the stop sequence number)."""
        self._post_all_lanes(worker, [L.MBX_SEQ_STOP] * self.layout.lanes)

    def reset_worker(self, worker: int):
        """Before a worker (re)joins: descriptors show the client's current position again (clears a stop)."""
        self._post_all_lanes(worker, self.seq[worker])

    # ------------------------------------------------------------------ dispatch / poll
    def dispatch(self, worker: int, lane: int, slot: int, n_valid: int, batch_id: int, kind: int = 0,
                 host_src: Optional[torch.Tensor] = None, fetch: bool = True):
        """Hand arena slot `slot` to (worker, lane).  Everything is queued on the (worker, lane) stream of the client:
        [upload of `host_src` into the slot] -> post descriptor -> wait for the done word -> read the result slot back
        into pinned memory -> event.  Returns at once; see `poll`."""
        lay = self.layout
        assert not self.busy[worker][lane]
        k = worker * lay.lanes + lane
        self.seq[worker][lane] += 1
        if self.seq[worker][lane] >= L.MBX_SEQ_STOP:
            self.seq[worker][lane] = 1
        seq = self.seq[worker][lane]
        c = self._cache[k]                 # per (worker, lane): addresses and handles resolved once
        desc = c["desc"]
        desc.seq, desc.epoch, desc.n_valid, desc.kind = seq, self.epoch[worker], n_valid, kind
        desc.src_off, desc.batch_id = lay.arena_off(slot), batch_id & 0xFFFFFFFF
        stream = self.streams[k]
        hs = c["hs"]
        self._hs_status[k] = 0xFFFF
        res = c["res"]
        tmo = int(self.done_timeout_s * 1e9)
        loc = self.local.get(worker)
        if loc is None:
            # hot path (7 of 8 dispatches on a full box): one native call, no Python context managers
            if kind == 1:
                host_src = None           # the worker uploads its own shard
            src_ptr, src_bytes = (host_src.data_ptr(), host_src.numel()) if host_src is not None else (None, 0)
            if host_src is not None:
                assert lay.raw_stride == lay.H * lay.W * 3, "host upload needs a dense frame pitch"
            rc = self.lib.dvc_mbx_dispatch(src_ptr, self.arena.ptr + desc.src_off, src_bytes, c["desc_addr"], c["desc_ref"],
                                           c["done_addr"], hs, tmo, res, c["h_res"] if fetch else None, lay.copy_bytes,
                                           c["event"], c["stream"])
            if rc:
                L.check(rc, "mbx_dispatch")
            self.busy[worker][lane] = True
            self.dispatches += 1
            return
        # the client's own GPU works too: same kernels, same words, queued inline on the same stream (a wait never spins
        # on something a later launch of this GPU has to produce)
        gexec = self._local_gexec.get((worker, lane))
        if gexec is None and loc.engine.cfg.use_cuda_graph:
            gexec = self._local_gexec[(worker, lane)] = loc.graph_exec(lane)
        if gexec is not None and host_src is None:
            rc = self.lib.dvc_mbx_local_dispatch(c["desc_addr"], c["desc_ref"], self.arena.ptr, loc.local_raw[lane].data_ptr(),
                                                 loc.ctl_ptr(lane), gexec, c["done_addr"], hs, tmo, res,
                                                 c["h_res"] if fetch else None, lay.copy_bytes, c["event"], c["stream"])
            if rc:
                L.check(rc, "mbx_local_dispatch")
            self.busy[worker][lane] = True
            self.dispatches += 1
            return
        with torch.cuda.device(self.dev), torch.cuda.stream(stream):
            L.set_device(self.dev)
            if host_src is not None and kind == 1:
                loc.local_raw[lane].view(lay.B, lay.raw_stride)[:, : lay.H * lay.W * 3].copy_(
                    host_src.reshape(lay.B, -1), non_blocking=True)
            elif host_src is not None:
                self.slot_tensor(slot).copy_(host_src, non_blocking=True)
            sp = c["stream"]
            L.check(self.lib.dvc_mbx_post(c["desc_addr_arr"], c["desc_ref"], 1, sp), "mbx_post")
            loc.enqueue_wait(lane, None, seq, 0.001)
            loc.enqueue_body(lane)
            c["seq_arr"][0] = seq
            L.check(self.lib.dvc_mbx_wait_done(c["done_addr_arr"], c["seq_arr"], 1, hs, tmo, sp), "mbx_wait_done")
            if fetch:
                L.check(self.lib.dvc_mbx_copy_out(res, c["h_res"], lay.copy_bytes, c["event"], sp), "mbx_copy_out")
            else:
                self.events[k].record(stream)
        self.busy[worker][lane] = True
        self.dispatches += 1

    def poll(self, worker: int, lane: int) -> Optional[str]:
        """None while the lane's dispatch is still running (or nothing is), else "ok" / "timeout" once."""
        if not self.busy[worker][lane]:
            return None
        k = worker * self.layout.lanes + lane
        if not self.events[k].query():
            return None
        self.busy[worker][lane] = False
        return "ok" if int(self.hs[k]["status"]) == L.MBX_GO else "timeout"

    def poll_all(self) -> List[Tuple[int, int, str]]:
        """Every busy lane whose dispatch has finished since the last call, as (worker, lane, "ok" | "timeout"); one
        native call queries all events."""
        lanes = self.layout.lanes
        n = len(self._busy_arr)
        busy = self.busy
        any_busy = False
        for k in range(n):
            b = busy[k // lanes][k % lanes]
            self._busy_arr[k] = b
            any_busy = any_busy or b
        if not any_busy:
            return []
        found = self.lib.dvc_mbx_poll(self._ev_arr, self._busy_arr, n, self._done_arr)
        if found < 0:
            L.check(found, "mbx_poll")
        out = []
        if found:
            for k in range(n):
                if self._done_arr[k]:
                    w, l = divmod(k, lanes)
                    busy[w][l] = False
                    out.append((w, l, "ok" if int(self._hs_status[k]) == L.MBX_GO else "timeout"))
        return out

    def wait_ns(self, worker: int, lane: int) -> int:
        return int(self.hs[worker * self.layout.lanes + lane]["wait_ns"])

    def result(self, worker: int, lane: int) -> Dict[str, np.ndarray]:
        """Views into the pinned copy of the lane's last result slot; valid until the lane's next dispatch."""
        k = worker * self.layout.lanes + lane
        v = self._views.get(k)
        if v is None:
            lay, B = self.layout, self.layout.B
            h = self.h_res[k].numpy()
            v = {"det": h[lay.det_off:lay.boxes_off].view(np.float32).reshape(B, 100, 7),
                 "boxes_i": h[lay.boxes_off:lay.det_count_off].view(np.int32).reshape(B, 100, 4),
                 "det_count": h[lay.det_count_off:lay.box_count_off].view(np.int32),
                 "box_count": h[lay.box_count_off:lay.copy_bytes].view(np.int32)}
            self._views[k] = v
        return v

    def device_result(self, worker: int, lane: int) -> Dict[str, torch.Tensor]:
        """The result slot itself (device tensors on the client GPU)."""
        lay, B = self.layout, self.layout.B
        off = lay.res_addr(0, worker, lane)
        return {"det": self.res.tensor(torch.float32, (B, 100, 7), offset_bytes=off + lay.det_off),
                "boxes_i": self.res.tensor(torch.int32, (B, 100, 4), offset_bytes=off + lay.boxes_off),
                "det_count": self.res.tensor(torch.int32, (B,), offset_bytes=off + lay.det_count_off),
                "box_count": self.res.tensor(torch.int32, (B,), offset_bytes=off + lay.box_count_off),
                "flags": self.res.tensor(torch.int32, (B,), offset_bytes=off + lay.flags_off)}

    def close(self):
        with torch.cuda.device(self.dev):
            torch.cuda.synchronize(self.dev)
            for b in (self.ctl, self.res, self.arena):
                b.free()


def open_box(export: Dict[str, Any], device) -> Tuple[BoxLayout, int, int, int]:
    """Worker process: map the client's three buffers.  Returns (layout, ctl_base, res_base, arena_base)."""
    from . import p2p
    d = dict(export["layout"])
    lay = BoxLayout(d["max_workers"], d["lanes"], d["B"], d["H"], d["W"], d["arena_slots"], d["weights_bytes"])
    return lay, p2p.open_handle(export["ctl"], device), p2p.open_handle(export["res"], device), p2p.open_handle(export["arena"], device)


# ====================================================================================== scheduler front-end
@dataclasses.dataclass
class _Flight:
    batch: Batch
    worker: int
    lane: int
    slot: int
    t0: float
    cancelled: bool = False


class ElasticBox:
    """The PyTorch-level scheduler on the mailbox data plane (replaces server.py:77-158 for the GPUs of one box).

    Workers are named ``gpu<index>``; `add_worker` / `remove_worker` may be called at any time, mid-stream.  A batch
    whose worker times out (stalled, killed) or leaves is re-queued at the front and goes to another worker; the
    departed worker's epoch word is bumped first, so rows it might still deliver later are rejected on its own GPU
    (K9 gate) and counted in its `stale` word.  Results come back through `on_result(batch, views, worker)`; `views`
    are numpy views into pinned memory that stay valid until the callback returns.

    `client` is a `ClientBox` or anything with the same dispatch / poll / result / set_epoch / stop_worker /
    reset_worker / alloc_slot / free_slot surface (tests use a CPU stand-in)."""

    def __init__(self, client, scheduler: Optional[Scheduler] = None, on_result: Optional[Callable] = None,
                 requester: str = "client", fetch: bool = True, log: Optional[Callable[[str], None]] = None, metrics=None):
        self.client = client
        self.lay = client.layout
        self.sched = scheduler or Scheduler(policy="least_loaded", requester_works=True, worker_timeout_s=3600.0)
        self.on_result = on_result
        self.requester = requester
        self.fetch = fetch
        self.log = log or (lambda m: None)
        self.metrics = metrics             # utils.metrics.Metrics (optional): per-batch latency, membership events
        self.flights: Dict[Tuple[int, int], _Flight] = {}
        self.alive: Dict[int, bool] = {}
        self.epoch_counter = 0
        self.stats = collections.Counter()
        self.failed_workers: List[int] = []
        self.procs: Dict[int, Any] = {}

    @staticmethod
    def name(w: int) -> str:
        return f"gpu{w}"

    # ------------------------------------------------------------------ membership
    def add_worker(self, w: int):
        """Worker `w` (re)joins: fresh epoch word, descriptors reset to the client's position, one credit per lane."""
        assert 0 <= w < self.lay.max_workers
        self.epoch_counter += 1
        self.client.reset_worker(w)
        self.client.set_epoch(w, self.epoch_counter)
        self.alive[w] = True
        self.sched.join(self.name(w), capacity=self.lay.lanes)
        self.stats["joins"] += 1
        if self.metrics is not None:
            self.metrics.emit("join", worker=self.name(w), epoch=self.epoch_counter)

    def remove_worker(self, w: int, reason: str = "leave", stop: bool = True) -> int:
        """Worker `w` leaves now (or was found dead).  Its in-flight batches are re-queued; returns how many."""
        if not self.alive.get(w):
            return 0
        self.alive[w] = False
        self.epoch_counter += 1
        self.client.set_epoch(w, self.epoch_counter)          # K14: whatever it still delivers is stale from here on
        requeued = self.sched.leave(self.name(w), reason)
        for (fw, fl), flight in self.flights.items():
            if fw == w and not flight.cancelled:
                flight.cancelled = True            # re-queued above; its done word will not come, do not wait for it
                self.client.cancel(fw, fl)
        if stop:
            self.client.stop_worker(w)
        self.stats["leaves" if reason == "leave" else "failures"] += 1
        self.stats["redispatched"] += len(requeued)
        if self.metrics is not None:
            self.metrics.emit("leave", worker=self.name(w), reason=reason, epoch=self.epoch_counter, redispatched=len(requeued))
        return len(requeued)

    def workers_alive(self) -> List[int]:
        return [w for w, a in self.alive.items() if a]

    def spawn_workers(self, workers: Sequence[Tuple[int, int]], graph, cfg, stall_after: Optional[int] = None,
                      ready_timeout_s: float = 300.0, local_feed_spec: Optional[Tuple] = None, delay_ms: float = 0.0) -> None:
        """Start one worker PROCESS per (worker index, device index) pair - all at once, they build their engines in
        parallel - and add each to the pool once its lanes are built."""
        import multiprocessing as mp
        ctx = mp.get_context("spawn")
        started = []
        for w, dev_index in workers:
            self.client.reset_worker(w)            # clears an earlier stop before the newcomer reads its position
            parent, child = ctx.Pipe()
            p = ctx.Process(target=worker_process_main,
                            args=(self.client.export(), w, dev_index, graph, cfg, child, stall_after, local_feed_spec, delay_ms), daemon=True)
            p.start()
            started.append((w, p, parent))
        deadline = time.monotonic() + ready_timeout_s
        for w, p, parent in started:
            if not parent.poll(max(0.0, deadline - time.monotonic())):
                p.kill()
                raise RuntimeError(f"worker gpu{w} did not come up within {ready_timeout_s} s")
            msg = parent.recv()
            assert msg[0] == "ready", msg
            self.procs[w] = (p, parent)
            self.add_worker(w)

    def spawn_worker(self, w: int, device_index: int, graph, cfg, stall_after: Optional[int] = None, ready_timeout_s: float = 300.0):
        """Start a worker PROCESS on GPU `device_index` and add it once its lanes are built.  Returns the process."""
        self.spawn_workers([(w, device_index)], graph, cfg, stall_after, ready_timeout_s)
        return self.procs[w][0]

    def rejoin_worker(self, w: int, ready_timeout_s: float = 60.0):
        """A worker process that left (its serving loop returned on stop) joins again: no new process, no new engine."""
        p, conn = self.procs[w]
        while conn.poll(0.0):
            conn.recv()                            # its "served" report of the previous membership
        self.client.reset_worker(w)
        conn.send("rejoin")
        if not conn.poll(ready_timeout_s):
            raise RuntimeError(f"worker gpu{w} did not answer the re-join")
        msg = conn.recv()
        assert msg[0] == "ready", msg
        self.add_worker(w)

    # ------------------------------------------------------------------ work intake
    def submit_slot(self, slot: int, frame_ids: Sequence[int], keep_slot: bool = True, requester: Optional[str] = None) -> Batch:
        """Frames already resident in arena slot `slot` (decoded or produced on the client GPU)."""
        return self.sched.submit(requester or self.requester, frame_ids, payload={"slot": slot, "host": None, "keep": keep_slot})

    def submit_local(self, frame_ids: Sequence[int], host: Optional[torch.Tensor] = None, requester: Optional[str] = None) -> Batch:
        """Sharded ingest: the frames are (or will be) in the WORKER's own memory - every worker process uploads its
        own shard over its own PCIe link (`WorkerLanes.serve(local_feed=...)`); the descriptor (kind 1) only carries
        the permission to deliver.  `host` is what the client's own GPU uploads when the batch lands on it."""
        return self.sched.submit(requester or self.requester, frame_ids, payload={"slot": 0, "host": host, "keep": True, "kind": 1})

    def submit_host(self, frames: torch.Tensor, frame_ids: Sequence[int], requester: Optional[str] = None) -> Batch:
        """A pinned host batch [B, H, W, 3]; it is uploaded into an arena slot on the stream of the lane it goes to."""
        return self.sched.submit(requester or self.requester, frame_ids, payload={"slot": None, "host": frames, "keep": False})

    # ------------------------------------------------------------------ the pump
    def pump(self) -> int:
        """Retire finished dispatches, then dispatch as much queued work as the workers' credits allow.
        Returns the number of batches completed by this call."""
        done = 0
        if hasattr(self.client, "poll_all"):
            finished = [((w, l), r) for w, l, r in self.client.poll_all() if (w, l) in self.flights]
        else:
            finished = [(key, r) for key, r in ((key, self.client.poll(*key)) for key in list(self.flights.keys())) if r is not None]
        for key, r in finished:
            fl = self.flights.pop(key)
            if fl.cancelled:
                continue
            if r == "ok":
                first = self.sched.complete(fl.batch.batch_id, self.name(fl.worker))
                if first is not None:
                    done += 1
                    self.stats["completed"] += 1
                    if self.metrics is not None:
                        self.metrics.observe("batch_latency_ms", (time.monotonic() - fl.t0) * 1e3)
                        self.metrics.inc("frames_out", len(fl.batch.frame_ids))
                    if self.on_result is not None:
                        self.on_result(fl.batch, self.client.result(fl.worker, fl.lane) if self.fetch else None, fl.worker)
                    pl = fl.batch.payload
                    if not pl["keep"] and pl["slot"] is not None:
                        self.client.free_slot(pl["slot"])
                        pl["slot"] = None
                else:
                    self.stats["duplicates"] += 1
            else:
                self.stats["timeouts"] += 1
                if self.alive.get(fl.worker):
                    self.log(f"worker gpu{fl.worker} did not answer batch {fl.batch.batch_id} in time: re-dispatching")
                    self.failed_workers.append(fl.worker)
                    self.remove_worker(fl.worker, reason="timeout", stop=True)
        while True:
            d = self.sched.next_dispatch()
            if d is None:
                break
            batch, winfo, _epoch = d
            w = int(winfo.addr[3:])
            lane = next((l for l in range(self.lay.lanes) if (w, l) not in self.flights), None)
            if lane is None or not self.alive.get(w):
                self.sched.undo_dispatch(batch.batch_id)      # an old flight of a revived worker still holds the lane
                break
            pl = batch.payload
            host = None
            if pl["host"] is not None:
                if pl["slot"] is None:
                    pl["slot"] = self.client.alloc_slot()
                    if pl["slot"] is None:
                        self.sched.undo_dispatch(batch.batch_id)
                        self.stats["arena_full"] += 1
                        break
                # a re-dispatched batch is uploaded again on ITS lane's stream: the first upload was ordered on another
                host = pl["host"]
            self.client.dispatch(w, lane, pl["slot"], len(batch.frame_ids), batch.batch_id, kind=pl.get("kind", 0),
                                 host_src=host, fetch=self.fetch)
            self.flights[(w, lane)] = _Flight(batch, w, lane, pl["slot"], time.monotonic())
            self.stats["dispatched"] += 1
        return done

    def outstanding(self) -> int:
        return self.sched.outstanding()

    def drain(self, timeout_s: float = 30.0, sleep_s: float = 0.00002) -> bool:
        """Pump until nothing is queued or in flight.  False on time-out (e.g. no worker left)."""
        deadline = time.monotonic() + timeout_s
        while self.sched.outstanding() or self.flights:
            if self.pump() == 0:
                if time.monotonic() > deadline:
                    return False
                time.sleep(sleep_s)
        return True

    def shutdown(self, join_s: float = 10.0):
        for w in list(self.workers_alive()):
            self.remove_worker(w, reason="leave", stop=True)
        for w, (p, conn) in list(self.procs.items()):
            try:
                conn.send("stop")
            except Exception:
                pass
            p.join(join_s)
            if p.is_alive():
                p.kill()
        self.procs.clear()


# ====================================================================================== process front-end
def worker_process_main(export: Dict[str, Any], worker: int, device_index: int, graph, cfg, conn=None,
                        stall_after: Optional[int] = None, local_feed_spec: Optional[Tuple] = None, delay_ms: float = 0.0):
    """Entry point of a spawned worker process (multiprocessing 'spawn'): map the client's box, adopt its weights,
    serve until told to stop; then wait on the pipe: "rejoin" serves again (workers leave and join at will,
    server.py:104-154), anything else ends the process.  "ready" goes out whenever the lanes are about to serve,
    ("served", batches, stale frames) whenever a membership ends.  `stall_after` (fault injection) makes the worker go
    silent after that many batches, `delay_ms` makes it pause that long after every batch (a slow volunteer).  `local_feed_spec` = ("synthetic", pool, seed): sharded ingest, the worker uploads
    its own synthetic frames (a pinned pool of `pool` batches) instead of pulling them from the client."""
    from ..engine import NativeEngine
    torch.cuda.set_device(device_index)
    dev = f"cuda:{device_index}"
    eng = NativeEngine(graph, dev, cfg, requester_id=getattr(cfg, "requester_id", "") or "client")
    lay, cb, rb, ab = open_box(export, dev)
    lanes = WorkerLanes(eng, lay, cb, rb, ab, worker, adopt_weights=True)
    feed = None
    if local_feed_spec is not None and local_feed_spec[0] == "synthetic":
        _, pool, seed = local_feed_spec
        rng = np.random.default_rng(seed + worker)
        small = rng.integers(0, 256, (pool, lay.B, lay.H // 8 + 1, lay.W // 8 + 1, 3), dtype=np.uint8)
        frames = np.repeat(np.repeat(small, 8, axis=2), 8, axis=3)[:, :, :lay.H, :lay.W]
        host = torch.from_numpy(np.ascontiguousarray(frames)).pin_memory()
        feed = lambda i: host[i % pool]
    stop_seen = [False]

    def stop() -> bool:
        return conn is not None and conn.poll()

    while True:
        ready = (lambda: conn.send(("ready", worker))) if conn is not None else None
        n = lanes.serve(should_stop=stop, max_batches=stall_after, local_feed=feed, on_ready=ready, throttle_s=delay_ms / 1e3)
        if stall_after is not None:
            while not stop():          # injected stall: alive, mapped, silent
                time.sleep(0.01)
        if conn is None:
            return
        try:
            conn.send(("served", n, lanes.stale_count()))
            msg = conn.recv()          # blocks: "rejoin" or anything else / EOF = exit
        except (EOFError, OSError):
            return
        if msg != "rejoin":
            return
