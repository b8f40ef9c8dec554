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
"""
from __future__ import annotations

from typing import Optional

import torch

from ..ops import kernels as K
from ..ops import lib as L
from . import p2p


class BoxCluster:
    def __init__(self, engine, B: int, H: int, W: int, client_rank: int = 0, group=None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.engine = engine
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.client_rank = client_rank
        self.is_client = self.rank == client_rank
        self.B, self.H, self.W = B, H, W
        self.dev = engine.device
        st = engine.state(B, H, W)
        self.st = st
        n = self.world * B
        self.n = n
        bufs = {k: None for k in ("frames", "det", "det_count", "boxes", "box_count", "flags")}
        if self.is_client:
            bufs["frames"] = p2p.DevBuffer(n * st.stride400, self.dev)
            bufs["det"] = p2p.DevBuffer(n * 700 * 4, self.dev)
            bufs["det_count"] = p2p.DevBuffer(n * 4, self.dev)
            bufs["boxes"] = p2p.DevBuffer(n * 400 * 4, self.dev)
            bufs["box_count"] = p2p.DevBuffer(n * 4, self.dev)
            bufs["flags"] = p2p.DevBuffer(n * 4, self.dev)
        self.bufs = bufs
        self.ptrs = p2p.exchange(bufs, self.dev, group)
        c = client_rank
        r = self.rank
        base = {k: self.ptrs[k][c] for k in bufs}
        engine.bind_remote(st,
                           frames_ptr=base["frames"] + r * B * st.stride400, frame_stride=st.stride400,
                           det_ptr=base["det"] + r * B * 700 * 4, det_count_ptr=base["det_count"] + r * B * 4,
                           boxes_ptr=base["boxes"] + r * B * 400 * 4, box_count_ptr=base["box_count"] + r * B * 4,
                           flags_ptr=base["flags"] + r * B * 4, flag_value=1)
        self.step_id = 0
        self._tok = torch.zeros(1, device=self.dev)
        if self.is_client:
            self.raw_all = torch.zeros((n, H, W, 3), dtype=torch.uint8, device=self.dev)
            self.frames_flat = bufs["frames"].tensor(torch.uint8, (n, st.stride400))
            self.frames400 = torch.as_strided(self.frames_flat, (n, st.h4, st.w4, 3), (st.stride400, st.w4 * 3, 3, 1))
            self.det = bufs["det"].tensor(torch.float32, (n, 100, 7))
            self.det_count = bufs["det_count"].tensor(torch.int32, (n,))
            self.boxes_i = bufs["boxes"].tensor(torch.int32, (n, 100, 4))
            self.box_count = bufs["box_count"].tensor(torch.int32, (n,))
            self.flags = bufs["flags"].tensor(torch.int32, (n,))

    def _order(self):
        """Stream-ordered rendezvous of all ranks (no host synchronisation)."""
        self.dist.all_reduce(self._tok, group=self.group)

    def step(self, raw_all: Optional[torch.Tensor] = None, annotate: bool = True):
        """One global batch: world*B frames (on the client GPU) -> detections + annotated frames on the
        client GPU.  Asynchronous on the current stream of every rank."""
        st = self.st
        self.step_id += 1
        if self.is_client:
            K.resize_u8(self.raw_all if raw_all is None else raw_all, st.t_area, self.frames_flat, st.stride400)
        self._order()                                   # ring filled -> everyone may pull
        self.engine.run(st, annotate=False, resize=False)
        self._order()                                   # rows pushed -> the client may draw
        if self.is_client and annotate:
            m = st.masks
            nl, lh, lw = m.label_masks.shape
            idh, idw = m.id_mask.shape
            L.check(self.engine.lib.dvc_annotate(L.ptr(self.frames_flat), st.stride400, None, self.n, st.h4, st.w4,
                                                 L.ptr(self.boxes_i), L.ptr(self.box_count), None, L.ptr(m.id_mask), m.id_xy[0],
                                                 m.id_xy[1], idw, idh, L.ptr(m.label_masks), nl, m.label_xy[0], m.label_xy[1],
                                                 lw, lh, L.stream_ptr()), "annotate")

    def launches_per_step(self) -> int:
        n = self.st.n_ops
        if self.is_client:
            n += 2
        return n
