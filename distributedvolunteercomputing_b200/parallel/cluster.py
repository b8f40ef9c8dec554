"""In-box multi-GPU pipeline for processes started together (torchrun): one process per GPU, frames scattered and
detections gathered by the kernels themselves over NVLink, ordered by device-side mailboxes (parallel/mailbox.py).

    client GPU (rank c)                                worker GPU r (every rank, the client included)
    -------------------                                ----------------------------------------------
    frame arena (raw frames)         <===== K1a pulls its B frames with peer loads (scatter)
    descriptors (st.release.sys)     <----- wait kernel polls ld.acquire.sys over NVLink
    result slots + done words        <===== K9 pushes rows with peer stores, last CTA publishes the done word (gather)
    wait kernel polls the done word -> device->host copy of the slot

torch.distributed is used ONLY for set-up (one broadcast of the client's IPC handles) and for barriers between
phases; nothing inside a phase is a collective, so a slow or dead rank delays nobody but its own batches, which the
client's scheduler (`ElasticBox`) hands to another rank.  This replaces the reference's star relay, where every
frame crosses TCP four times through the server and every chunk waits for a `b'OK'` (server.py:56,62,89,60,73).

Annotated 400-wide frames stay on the GPU that produced them (`lanes.frames400(l)`); detections, integer boxes and
counts come back to the client.
"""
from __future__ import annotations

from typing import Callable, Optional

import torch

from .mailbox import ClientBox, ElasticBox, WorkerLanes, open_box


class BoxCluster:
    def __init__(self, engine, B: int, H: int, W: int, client_rank: int = 0, lanes: int = 2, arena_slots: Optional[int] = None,
                 done_timeout_s: float = 2.0, on_result: Optional[Callable] = None, fetch: bool = True, client_works: bool = True):
        import torch.distributed as dist
        self.dist = dist
        self.engine = engine
        self.rank = dist.get_rank()
        self.world = dist.get_world_size()
        self.client_rank = client_rank
        self.is_client = self.rank == client_rank
        self.B, self.H, self.W = B, H, W
        self.dev = engine.device
        self.client: Optional[ClientBox] = None
        self.box: Optional[ElasticBox] = None
        self.client_works = client_works
        exp = [None]
        if self.is_client:
            self.client = ClientBox(self.dev, B, H, W, max_workers=self.world, lanes=lanes, arena_slots=arena_slots,
                                    engine=engine, done_timeout_s=done_timeout_s)
            exp[0] = self.client.export()
        dist.broadcast_object_list(exp, src=client_rank)            # set-up only: CUDA IPC handles + layout
        if self.is_client:
            self.lanes = WorkerLanes(engine, self.client.layout, *self.client.bases(), self.rank) if client_works else None
            if self.lanes is not None:
                self.client.attach_local(self.rank, self.lanes)
            self.box = ElasticBox(self.client, on_result=on_result, fetch=fetch)
        else:
            lay, cb, rb, ab = open_box(exp[0], self.dev)
            self.lanes = WorkerLanes(engine, lay, cb, rb, ab, self.rank, adopt_weights=True)     # K13: weights from the client
        self.layout = self.client.layout if self.is_client else self.lanes.layout
        self.serve_ms = 0.0
        dist.barrier()

    # ------------------------------------------------------------------ phases (collective: call on every rank)
    def begin(self):
        """Start a phase: every rank becomes a member.  Collective."""
        if self.is_client:
            for w in range(self.world):
                if w == self.rank and not self.client_works:
                    continue
                self.box.add_worker(w)          # descriptors show the client's position, fresh epoch words
        self.dist.barrier()
        if not self.is_client:
            self._synced = self.lanes.sync_position()      # read BEFORE the client may post the phase's first descriptor
        self.dist.barrier()

    def serve(self, local_feed=None) -> int:
        """Non-client ranks: serve until the client ends the phase.  Returns the batches served; `serve_ms` is the
        device time between entering and leaving the loop (CUDA events on this rank)."""
        assert not self.is_client
        from ..utils.trace import DeviceTimer, span
        with torch.cuda.device(self.dev), span(f"mailbox.serve.rank{self.rank}"):
            with DeviceTimer() as t:
                n = self.lanes.serve(local_feed=local_feed, expect_work=True, synced=True) if self._synced else 0
            self.serve_ms = t.ms()
        return n

    def end(self):
        """Client: tell every other rank's serving loop to return.  Then a barrier on all ranks.  Collective."""
        if self.is_client:
            self.box.drain(30.0)
            for w in list(self.box.workers_alive()):
                self.box.remove_worker(w, reason="leave", stop=(w != self.rank))
        self.dist.barrier()

    def stale_count(self) -> int:
        return self.lanes.stale_count() if self.lanes is not None else 0

    def launches_per_batch(self) -> int:
        return self.lanes.launches_per_batch() if self.lanes is not None else 0

    def close(self):
        if self.is_client:
            self.client.close()
