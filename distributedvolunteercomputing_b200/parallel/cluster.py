"""In-box multi-GPU pipeline: one process per GPU (torchrun), frames scattered and detections gathered
by the kernels themselves over NVLink.

Topology for one requester (`BoxCluster`):

    client GPU (rank c)                               worker GPU r (every rank, the client included)
    -------------------                               ----------------------------------------------
    raw frames  --K1a resize-->  frame ring  <===== K1b pulls its B frames with peer loads (scatter)
    result rings (det rows, counts, boxes, flags) <===== detection kernel pushes rows with peer stores
    K10 annotate all frames from the gathered boxes        + st.release.sys flag (gather)

torch.distributed (NCCL) is used for set-up (IPC handle exchange) and for two stream-ordered
one-element all-reduces per step that order "ring filled" -> "pull" and "rows pushed" -> "annotate";
no frame or detection bytes go through NCCL.  This replaces the reference's star relay, where every
frame crosses TCP four times through the server (server.py:56,89,60,73).

Two scatter modes:
  * ``pull_raw`` (default): the client keeps only the RAW frame ring.  Every GPU's K1a resize pulls its
    raw frames with coalesced peer loads (the scatter is fused into the resize; 270 KB per 300x300 frame
    on the wire), runs the network locally, pushes its detection rows to the client (fused gather) and
    overlays its own 400-wide frames, reading the integer boxes back from the client ring.  The client does
    no per-frame work for the other GPUs, so throughput scales with the number of GPUs.
  * ``pull_400``: the client resizes every frame to 400 wide; K1b on every GPU pulls from that ring
    (scatter fused into resize-to-300 + normalise + conv0) and the client overlays all frames.

`lanes` global batches are in flight at once: each lane has its own rings, plan, CUDA graph, stream
and NCCL communicator, so the small latency-bound kernels of one global batch overlap the large
kernels of the other on every GPU.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from ..ops import kernels as K
from ..ops import lib as L
from . import p2p


class _Lane:
    pass


def p2p_align(n: int) -> int:
    return (n + 15) & ~15


class BoxCluster:
    def __init__(self, engine, B: int, H: int, W: int, client_rank: int = 0, lanes: int = 2, mode: str = "pull_raw"):
        assert mode in ("pull_raw", "pull_400")
        self.mode = mode
        import torch.distributed as dist
        self.dist = dist
        self.engine = engine
        self.rank = dist.get_rank()
        self.world = dist.get_world_size()
        self.client_rank = client_rank
        self.is_client = self.rank == client_rank
        self.B, self.H, self.W = B, H, W
        self.dev = engine.device
        self.n = self.world * B
        self.step_id = 0
        self.epoch = 0
        self._setup_control()
        self.lanes: List[_Lane] = [self._make_lane(k) for k in range(max(1, lanes))]
        # convenience aliases to lane 0 (single-lane users and tests)
        l0 = self.lanes[0]
        self.st = l0.st
        self.frames400_local = l0.st.frames400          # pull_raw: this rank's annotated frames
        if self.is_client:
            self.raw_all, self.frames400, self.det, self.det_count = l0.raw_all, l0.frames400, l0.det, l0.det_count
            self.boxes_i, self.box_count, self.flags = l0.boxes_i, l0.box_count, l0.flags

    # ------------------------------------------------------------------ K13 / K14: weights and membership epoch
    def _setup_control(self):
        """Every rank owns a small exported control block: [0] membership epoch word, then the packed-weight
        blob.  The client pushes into all of them with its own stores (no NCCL, no host copies)."""
        layout, total = self.engine.weight_layout()
        self._wbytes = total
        self.ctrl = p2p.DevBuffer(256 + total, self.dev)
        ptrs = p2p.exchange({"ctrl": self.ctrl}, self.dev, None)["ctrl"]
        self.ctrl_ptrs = ptrs
        self.epoch_word = self.ctrl.tensor(torch.int32, (1,))
        self.epoch_word.zero_()
        # K14 gate: every lane's deliver kernel compares epoch_word (written by the client's broadcast) with
        # epoch_expect (the epoch this rank dispatches under) and counts rejected frames in stale_frames
        self.epoch_expect = torch.zeros(1, dtype=torch.int32, device=self.dev)
        self.stale_frames = torch.zeros(1, dtype=torch.int32, device=self.dev)
        torch.cuda.synchronize(self.dev)
        self.weight_blob = self.ctrl.tensor(torch.uint8, (total,), offset_bytes=256)
        if self.is_client:
            self._epoch_targets = p2p.device_ptr_array(ptrs, self.dev)
            self._weight_targets = p2p.device_ptr_array([p + 256 for r, p in enumerate(ptrs) if r != self.rank], self.dev)

    def broadcast_weights(self):
        """K13: the client's packed weights are pushed once into every peer (reference: each client loads the
        caffemodel from disk itself, worker.py:194); peers then adopt them.  Collective: call on every rank."""
        if self.is_client:
            self.engine.export_weights(self.weight_blob)
            if self.world > 1:
                L.check(self.engine.lib.dvc_push_multicast(L.ptr(self.weight_blob), L.ptr(self._weight_targets), self.world - 1,
                                                           (self._wbytes + 15) & ~15, L.stream_ptr()), "push_multicast")
        torch.cuda.current_stream().synchronize()
        self.dist.barrier()
        if not self.is_client:
            self.engine.import_weights(self.weight_blob)
            torch.cuda.current_stream().synchronize()
        self.dist.barrier()

    def publish_epoch(self, epoch: int):
        """K14: write the membership epoch word into every GPU's control block (release at system scope), then every
        rank adopts it as the epoch it dispatches under.  Collective (barrier inside): batches already in flight
        that reach their deliver kernel after the word changed are rejected and counted in ``stale_frames``."""
        self.epoch = epoch
        with torch.cuda.device(self.dev):
            if self.is_client:
                L.check(self.engine.lib.dvc_epoch_broadcast(L.ptr(self._epoch_targets), self.world, epoch, L.stream_ptr()), "epoch")
                torch.cuda.current_stream().synchronize()
            self.dist.barrier()
            self.dispatch_epoch(epoch)

    def dispatch_epoch(self, epoch: int):
        """The epoch this rank's following steps are dispatched under (stream-ordered device word)."""
        self.epoch_expect.fill_(int(epoch))

    def stale_count(self) -> int:
        """Frames this rank's deliver kernels rejected so far because the membership epoch had moved on."""
        return int(self.stale_frames.item())

    def _make_lane(self, k: int) -> _Lane:
        dist, eng, B, n = self.dist, self.engine, self.B, self.n
        ln = _Lane()
        ln.group = dist.new_group(list(range(self.world))) if k > 0 else None      # own communicator per lane
        ln.stream = torch.cuda.Stream(device=self.dev)
        st = eng.state(B, self.H, self.W, slot=200 + k)
        ln.st = st
        raw_bytes = p2p_align(self.H * self.W * 3)
        ln.raw_stride = raw_bytes
        bufs = {name: None for name in ("frames", "raw", "det", "det_count", "boxes", "box_count", "flags")}
        if self.is_client:
            if self.mode == "pull_400":
                bufs["frames"] = p2p.DevBuffer(n * st.stride400, self.dev)
            else:
                bufs["raw"] = p2p.DevBuffer(n * raw_bytes, self.dev)
            bufs["det"] = p2p.DevBuffer(n * 700 * 4, self.dev)
            bufs["det_count"] = p2p.DevBuffer(n * 4, self.dev)
            bufs["boxes"] = p2p.DevBuffer(n * 400 * 4, self.dev)
            bufs["box_count"] = p2p.DevBuffer(n * 4, self.dev)
            bufs["flags"] = p2p.DevBuffer(n * 4, self.dev)
        ln.bufs = bufs
        ptrs = p2p.exchange(bufs, self.dev, None)
        c, r = self.client_rank, self.rank
        base = {name: ptrs[name][c] for name in bufs}
        ln.base = base
        sinks = dict(det_ptr=base["det"] + r * B * 700 * 4, det_count_ptr=base["det_count"] + r * B * 4,
                     boxes_ptr=base["boxes"] + r * B * 400 * 4, box_count_ptr=base["box_count"] + r * B * 4,
                     flags_ptr=base["flags"] + r * B * 4, flag_value=1)
        if self.mode == "pull_400":
            eng.bind_remote(st, frames_ptr=base["frames"] + r * B * st.stride400, frame_stride=st.stride400, **sinks)
        else:
            eng.bind_remote(st, frames_ptr=L.ptr(st.frames400_flat), frame_stride=st.stride400, **sinks)
            # this rank's slice of the client's raw ring, as a tensor the resize kernel reads over NVLink
            ln.raw_mine = p2p.tensor_from_ptr(base["raw"] + r * B * raw_bytes, (B, raw_bytes), torch.uint8, self.dev)
            ln.raw_mine = torch.as_strided(ln.raw_mine, (B, self.H, self.W, 3), (raw_bytes, self.W * 3, 3, 1))
            ln.my_boxes = p2p.tensor_from_ptr(sinks["boxes_ptr"], (B, 100, 4), torch.int32, self.dev)
            ln.my_box_count = p2p.tensor_from_ptr(sinks["box_count_ptr"], (B,), torch.int32, self.dev)
        eng.set_epoch_gate(st, self.epoch_word, self.epoch_expect, self.stale_frames)
        ln.tok = torch.zeros(1, device=self.dev)
        if self.is_client:
            if self.mode == "pull_400":
                ln.raw_all = torch.zeros((n, self.H, self.W, 3), dtype=torch.uint8, device=self.dev)
                ln.frames_flat = bufs["frames"].tensor(torch.uint8, (n, st.stride400))
                ln.frames400 = torch.as_strided(ln.frames_flat, (n, st.h4, st.w4, 3), (st.stride400, st.w4 * 3, 3, 1))
            else:
                flat = bufs["raw"].tensor(torch.uint8, (n, raw_bytes))
                ln.raw_all = torch.as_strided(flat, (n, self.H, self.W, 3), (raw_bytes, self.W * 3, 3, 1))
                ln.frames_flat, ln.frames400 = None, None      # annotated frames stay on the GPU that produced them
            ln.det = bufs["det"].tensor(torch.float32, (n, 100, 7))
            ln.det_count = bufs["det_count"].tensor(torch.int32, (n,))
            ln.boxes_i = bufs["boxes"].tensor(torch.int32, (n, 100, 4))
            ln.box_count = bufs["box_count"].tensor(torch.int32, (n,))
            ln.flags = bufs["flags"].tensor(torch.int32, (n,))
        return ln

    def lane_of(self, step_index: int) -> _Lane:
        return self.lanes[step_index % len(self.lanes)]

    def _annotate(self, ln: _Lane, frames_flat, n_frames: int, boxes, counts):
        st = ln.st
        m = st.masks
        nl, lh, lw = m.label_masks.shape
        idh, idw = m.id_mask.shape
        L.check(self.engine.lib.dvc_annotate(L.ptr(frames_flat), st.stride400, None, n_frames, st.h4, st.w4,
                                             L.ptr(boxes), L.ptr(counts), None, L.ptr(m.id_mask), m.id_xy[0], m.id_xy[1],
                                             idw, idh, L.ptr(m.label_masks), nl, m.label_xy[0], m.label_xy[1], lw, lh,
                                             L.stream_ptr()), "annotate")

    def _run_lane(self, ln: _Lane, raw_all: Optional[torch.Tensor], annotate: bool, local_raw: Optional[torch.Tensor] = None):
        st = ln.st
        L.set_device(self.dev)
        if self.mode == "pull_400":
            if self.is_client:
                K.resize_u8(ln.raw_all if raw_all is None else raw_all, st.t_area, ln.frames_flat, st.stride400)
            self.dist.all_reduce(ln.tok, group=ln.group)      # ring filled -> everyone may pull (stream-ordered)
            self.engine.run(st, annotate=False, resize=False)
            self.dist.all_reduce(ln.tok, group=ln.group)      # rows pushed -> the client may draw
            if self.is_client and annotate:
                self._annotate(ln, ln.frames_flat, self.n, ln.boxes_i, ln.box_count)
        elif local_raw is not None:
            # "local ingest": this rank's frames arrived over its own PCIe link (sharded reader); nothing to
            # scatter, detections are still gathered on the client over NVLink
            K.resize_u8(local_raw, st.t_area, st.frames400_flat, st.stride400)
            self.engine.run(st, annotate=False, resize=False)
            if annotate:
                self._annotate(ln, st.frames400_flat, self.B, ln.my_boxes, ln.my_box_count)
            self.dist.all_reduce(ln.tok, group=ln.group)      # every rank's rows have landed -> the client may read
        else:
            if self.is_client and raw_all is not None and raw_all.data_ptr() != ln.raw_all.data_ptr():
                ln.raw_all.copy_(raw_all, non_blocking=True)  # frames enter the client's raw ring
            self.dist.all_reduce(ln.tok, group=ln.group)      # raw ring filled -> everyone may pull
            K.resize_u8(ln.raw_mine, st.t_area, st.frames400_flat, st.stride400,      # K1a: peer loads = the scatter,
                        epoch_gate=(self.epoch_word, self.epoch_expect))               # skipped for a stale dispatch
            self.engine.run(st, annotate=False, resize=False)                        # rows pushed to the client (gather)
            if annotate:
                self._annotate(ln, st.frames400_flat, self.B, ln.my_boxes, ln.my_box_count)
            self.dist.all_reduce(ln.tok, group=ln.group)      # every rank's rows have landed -> the client may read

    def step(self, raw_all: Optional[torch.Tensor] = None, annotate: bool = True, overlap: bool = False,
             local_raw: Optional[torch.Tensor] = None) -> _Lane:
        """One global batch: world*B frames (on the client GPU) -> detections + annotated frames on the
        client GPU.  Asynchronous.  With ``overlap`` the batch runs on its lane's stream so consecutive
        batches share the GPUs; call ``join()`` (or wait on the lane's stream) before reading results.
        Every rank must call step() the same number of times in the same order."""
        ln = self.lane_of(self.step_id if overlap else 0)
        self.step_id += 1
        with torch.cuda.device(self.dev):
            if not overlap:
                self._run_lane(ln, raw_all, annotate, local_raw)
                return ln
            ln.stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(ln.stream):
                self._run_lane(ln, raw_all, annotate, local_raw)
        return ln

    def join(self):
        for ln in self.lanes:
            torch.cuda.current_stream().wait_stream(ln.stream)

    def launches_per_step(self) -> int:
        n = self.st.n_ops
        if self.mode == "pull_raw" or self.is_client:
            n += 2
        return n
