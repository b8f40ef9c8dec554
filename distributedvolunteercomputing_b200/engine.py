"""Per-GPU inference engine: raw frames in, detections + annotated 400-wide frames out.

Replaces the body of the reference worker's `request` branch (worker.py:241-280:
imutils.resize -> cv2.resize -> blobFromImage -> net.forward -> filter -> draw).

Two implementations behind one interface:
  * ``NativeEngine`` – B200 path.  Static device buffers, weights packed once, the network
    lowered to a native plan (csrc/plan.cu: tcgen05/TMA GEMMs, depthwise, fused preproc+conv0,
    fused softmax/decode/NMS) replayed with ONE native call per batch and optionally captured
    in a CUDA graph.  Early (large-activation) layers run in L2-sized sub-batches.
  * ``CpuEngine``    – numpy/torch restatement for machines without a GPU (CI, remote CPU
    volunteers); also the oracle in tests.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Dict, Optional, Sequence, Tuple

import numpy as np
import torch

from .config import Config
from .models import detection as D
from .models import preprocess as P
from .models.mobilenet_ssd import SSDGraph, TorchSSD, packed_weights
from .utils import annotate as A


@dataclass
class BatchResult:
    """Per-batch outputs. Tensors (native) or numpy arrays (cpu)."""
    det: "np.ndarray | torch.Tensor"          # [B,100,7]
    det_count: "np.ndarray | torch.Tensor"    # [B]
    boxes_i: "np.ndarray | torch.Tensor"      # [B,100,4]
    box_count: "np.ndarray | torch.Tensor"    # [B]
    frames400: "np.ndarray | torch.Tensor"    # [B,h',w',3] annotated

    def numpy(self) -> "BatchResult":
        f = lambda t: t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else t
        return BatchResult(f(self.det), f(self.det_count), f(self.boxes_i), f(self.box_count), f(self.frames400))


def consider_ids(names: Sequence[str]) -> Tuple[int, ...]:
    return tuple(D.VOC_CLASSES.index(n) for n in names if n in D.VOC_CLASSES)


# ======================================================================================= CPU
class CpuEngine:
    def __init__(self, graph: SSDGraph, cfg: Optional[Config] = None, requester_id: str = ""):
        self.graph = graph
        self.cfg = cfg or Config()
        self.model = TorchSSD(graph).eval()
        self.requester_id = requester_id
        self.device = torch.device("cpu")

    def infer_blob(self, blob: np.ndarray):
        with torch.no_grad():
            loc, conf = self.model(torch.from_numpy(blob))
        return loc.numpy(), conf.numpy()

    def process(self, frames, requester_id: Optional[str] = None, annotate: bool = True) -> BatchResult:
        frames = np.asarray(frames)
        B = frames.shape[0]
        rid = self.requester_id if requester_id is None else requester_id
        f400s, blobs = [], []
        for f in frames:
            f400 = P.resize_width_400(f, self.cfg.out_width)
            img = P.resize_linear_u8(f400, P.NET_SIZE, P.NET_SIZE)
            f400s.append(f400)
            blobs.append(P.blob_from_image(img, self.cfg.mean, self.cfg.scale)[0])
        loc, conf = self.infer_blob(np.stack(blobs))
        g = self.graph
        det = np.zeros((B, 100, 7), np.float32)
        det_count = np.zeros((B,), np.int32)
        boxes_i = np.zeros((B, 100, 4), np.int32)
        box_count = np.zeros((B,), np.int32)
        ids = consider_ids(self.cfg.consider)
        for b in range(B):
            d = D.detection_output(loc[b], D.softmax(conf[b]), g.priors, g.variances, g.det)
            det[b, : len(d)] = d
            det_count[b] = len(d)
            h, w = f400s[b].shape[:2]
            bx, n = D.app_filter(d, w, h, self.cfg.confidence, ids)
            boxes_i[b, :n] = bx
            box_count[b] = n
            if annotate:
                label = A.label_text({self.cfg.consider[0]: n}) if len(self.cfg.consider) == 1 else \
                    A.label_text({c: sum(1 for r in d if r[2] > self.cfg.confidence and int(r[1]) == D.VOC_CLASSES.index(c))
                                  for c in self.cfg.consider})
                A.annotate_cpu(f400s[b], bx, rid, label)
        return BatchResult(det, det_count, boxes_i, box_count, np.stack(f400s))


# ==================================================================================== native
import threading

# CUDA-graph capture is process-global state: engines on different GPUs driven from different threads
# must not capture at the same time, and other threads' CUDA calls must not invalidate a capture.
_CAPTURE_LOCK = threading.Lock()
# which MobileNet blocks run as ONE fused depthwise + pointwise launch when DVC_FUSE_DWPW is not set ("0": none)
FUSE_DWPW_DEFAULT = "0"


class _ShapeState:
    """Everything that depends on (batch, source frame size): buffers, tables, plan, graph."""
    pass


class NativeEngine:
    def __init__(self, graph: SSDGraph, device="cuda:0", cfg: Optional[Config] = None, requester_id: str = "",
                 max_batch: Optional[int] = None):
        from .ops import kernels as K
        from .ops import lib as L
        if not torch.cuda.is_available():
            raise RuntimeError("NativeEngine needs a CUDA device")
        self.K, self.L = K, L
        self.lib = L.load()
        self.graph = graph
        self.cfg = cfg or Config()
        self.device = torch.device(device)
        self.requester_id = requester_id
        self.max_batch = max_batch or self.cfg.batch
        # LRU: a long-lived worker sees many (frame size, requester) pairs; states held by streams / lanes (slot > 0)
        # belong to their caller and are never evicted
        from .utils.lru import LruCache
        self._states: "LruCache[_ShapeState]" = LruCache(max(1, self.cfg.max_cached_states), on_evict=self._release_state,
                                                         can_evict=lambda st: getattr(st, "evictable", False))
        with torch.cuda.device(self.device):
            L.set_device(self.device)
            self._pack_weights()
            self.priors, self.variances = K.priorbox(graph.priors_spec, graph.input_hw, self.device)

    # ------------------------------------------------------------------ weights
    def _pack_weights(self):
        g, dev = self.graph, self.device
        pk = packed_weights(g)
        self.w: Dict[str, torch.Tensor] = {}
        self.b: Dict[str, torch.Tensor] = {}
        for c in g.convs:
            w, b = pk[c.name]
            if c.name == "conv0" or (c.k == 3 and not c.depthwise and c.cin * 9 < 64):
                # first layer: fp32 [27][Cout]
                self.w[c.name] = torch.from_numpy(np.ascontiguousarray(w.T)).to(dev, torch.float32)
            elif c.depthwise:
                self.w[c.name] = torch.from_numpy(w).to(dev, torch.float32)
            else:
                self.w[c.name] = torch.from_numpy(w).to(dev, torch.float16).contiguous()
            self.b[c.name] = torch.from_numpy(np.ascontiguousarray(b)).to(dev, torch.float32)
        for h in g.heads:
            w, b = pk[f"{h.tap}_mbox"]
            self.w[f"{h.tap}_mbox"] = torch.from_numpy(w).to(dev, torch.float16).contiguous()
            self.b[f"{h.tap}_mbox"] = torch.from_numpy(b).to(dev, torch.float32)

    def weight_layout(self):
        """Deterministic flat layout of every packed weight/bias tensor: [(key, kind, offset, nbytes)], total bytes.
        Each entry is 16-byte aligned so the blob can be pushed with 128-bit stores (K13)."""
        out, off = [], 0
        for kind, d in (("w", self.w), ("b", self.b)):
            for key in sorted(d):
                t = d[key]
                n = t.numel() * t.element_size()
                out.append((key, kind, off, n))
                off += (n + 15) & ~15
        return out, off

    def export_weights(self, flat: torch.Tensor):
        """Copy all packed weights into `flat` (uint8, device) following weight_layout()."""
        layout, total = self.weight_layout()
        assert flat.numel() >= total
        for key, kind, off, n in layout:
            t = (self.w if kind == "w" else self.b)[key]
            flat[off:off + n].copy_(t.contiguous().view(torch.uint8).reshape(-1))

    def import_weights(self, flat: torch.Tensor):
        """Overwrite this engine's packed weights from a blob written by export_weights (same model)."""
        layout, total = self.weight_layout()
        for key, kind, off, n in layout:
            t = (self.w if kind == "w" else self.b)[key]
            t.view(torch.uint8).reshape(-1).copy_(flat[off:off + n])

    # ------------------------------------------------------------------ per-shape state
    def state(self, B: int, H: int, W: int, requester_id: Optional[str] = None, slot: int = 0) -> _ShapeState:
        rid = self.requester_id if requester_id is None else requester_id
        key = (B, H, W, rid, slot)
        st = self._states.get(key)
        if st is None:
            with torch.cuda.device(self.device):
                self.L.set_device(self.device)
                st = self._make_state(B, H, W, rid)
            st.evictable = slot == 0
            self._states.put(key, st)
        return st

    def _release_state(self, key, st: _ShapeState):
        """LRU eviction: the state is idle (process() is synchronous), but make sure before its buffers are recycled."""
        torch.cuda.synchronize(self.device)
        for name in ("graph_ann", "graph_noann"):
            if hasattr(st, name):
                delattr(st, name)
        if getattr(st, "plan", None):
            self.lib.dvc_plan_destroy(st.plan)
            st.plan = None

    def _make_state(self, B: int, H: int, W: int, rid: str) -> _ShapeState:
        K, L, g, dev, cfg = self.K, self.L, self.graph, self.device, self.cfg
        st = _ShapeState()
        st.B, st.H, st.W = B, H, W
        st.h4, st.w4 = P.out_size_400(H, W, cfg.out_width)
        assert st.h4 >= 2, "frame too flat"
        st.raw = torch.zeros((B, H, W, 3), dtype=torch.uint8, device=dev)
        st.stride400 = K.align16(st.h4 * st.w4 * 3)
        st.frames400_flat = torch.zeros((B, st.stride400), dtype=torch.uint8, device=dev)
        st.frames400 = torch.as_strided(st.frames400_flat, (B, st.h4, st.w4, 3), (st.stride400, st.w4 * 3, 3, 1))
        st.t_area = K.make_area_tables(H, W, st.h4, st.w4, dev)
        st.t_lin = K.make_linear_tables(st.h4, st.w4, P.NET_SIZE, P.NET_SIZE, dev)
        st.loc = torch.zeros((B, g.num_priors * 4), dtype=torch.float32, device=dev)
        st.conf = torch.zeros((B, g.num_priors * g.num_classes), dtype=torch.float32, device=dev)
        # one contiguous result block (det | boxes | det_count | box_count): the pipelined API reads it back with ONE copy
        o1, o2, o3, o4 = B * 2800, B * 2800 + B * 1600, B * 2800 + B * 1600 + B * 4, B * 2800 + B * 1600 + B * 8
        st.res_flat = torch.zeros((o4,), dtype=torch.uint8, device=dev)
        st.res_offsets = (o1, o2, o3, o4)
        st.det = st.res_flat[:o1].view(torch.float32).view(B, 100, 7)
        st.boxes_i = st.res_flat[o1:o2].view(torch.int32).view(B, 100, 4)
        st.det_count = st.res_flat[o2:o3].view(torch.int32)
        st.box_count = st.res_flat[o3:o4].view(torch.int32)
        st.flags = torch.zeros((B,), dtype=torch.int32, device=dev)
        st.det_scratch = torch.empty((K.detection_scratch_bytes(B, g.num_priors, g.num_classes),), dtype=torch.uint8, device=dev)
        st.det_params = K.det_params(g.det, g.num_priors, st.w4, st.h4, cfg.confidence, consider_ids(cfg.consider))
        idm, ixy, lm, lxy = A.text_masks(rid, st.h4, st.w4, cfg.consider)
        st.masks = K.TextMasks(torch.from_numpy(idm).to(dev), ixy, torch.from_numpy(lm).to(dev), lxy)
        st.mean = (C.c_float * 3)(*[float(v) for v in cfg.mean])
        self._build_plan(st)
        st.graph = None
        return st

    # ------------------------------------------------------------------ plan lowering
    def _build_plan(self, st: _ShapeState):
        L, g, dev, lib = self.L, self.graph, self.device, self.lib
        B = st.B
        # sub_batch = 0 (default): no split.  Measured on B200 (batch 64, two batches in flight): 57.5k / 65.5k /
        # 70.1k / 71.7k frames/s for sub-batches of 8 / 16 / 32 / 64 - fewer, larger launches beat L2 residency.
        sub = self.cfg.sub_batch if self.cfg.sub_batch > 0 else B
        sub = min(sub, B)
        body = g.body
        # split point: everything up to and including `split_name` runs per sub-batch
        names = [c.name for c in body]
        split_name = "conv6" if "conv6" in names else names[min(len(names) - 1, 12)]
        split_idx = names.index(split_name)
        per_frame = [c.out_hw[0] * c.out_hw[1] * c.cout for c in body]
        early_max = max(per_frame[: split_idx + 1])
        late_max = max(per_frame[split_idx:])
        st.sA = torch.empty((sub * early_max,), dtype=torch.float16, device=dev)
        st.sB = torch.empty((sub * early_max,), dtype=torch.float16, device=dev)
        st.fA = torch.empty((B * late_max,), dtype=torch.float16, device=dev)
        st.fB = torch.empty((B * late_max,), dtype=torch.float16, device=dev)
        implicit_conv = os.environ.get("DVC_IMPLICIT_CONV", "1") != "0"     # 0: explicit im2col buffer + GEMM
        # patch buffer only for the dense 3x3 layers that do not take the implicit-GEMM path
        col_max = max([c.out_hw[0] * c.out_hw[1] * 9 * c.cin for c in body if c.k == 3 and not c.depthwise and c.name != "conv0"
                       and not (implicit_conv and c.cin % 64 == 0)] + [8])
        st.col = torch.empty((B * col_max,), dtype=torch.float16, device=dev)
        plan = lib.dvc_plan_create()
        st.plan = plan
        st._keep = []
        esz = 2
        heads = {h.tap: h for h in g.heads}

        def add_conv(c, src_ptr, dst_ptr, nb):
            if c.depthwise:
                L.check(lib.dvc_plan_add_dwconv(plan, src_ptr, nb, c.in_hw[0], c.in_hw[1], c.cin, c.stride,
                                                L.ptr(self.w[c.name]), L.ptr(self.b[c.name]), int(c.relu), dst_ptr), c.name)
            elif c.k == 1:
                M = nb * c.out_hw[0] * c.out_hw[1]
                L.check(lib.dvc_plan_add_gemm(plan, src_ptr, L.ptr(self.w[c.name]), L.ptr(self.b[c.name]), M, c.cout, c.cin,
                                              c.cin, 0, dst_ptr, c.cout, int(c.relu), None, 0), c.name)
            else:
                assert c.k == 3 and c.stride == 2 and c.pad == 1, c
                if implicit_conv and c.cin % 64 == 0:
                    # implicit GEMM: the patch rows are gathered by im2col-mode TMA loads inside the GEMM kernel
                    L.check(lib.dvc_plan_add_conv3x3s2(plan, src_ptr, nb, c.in_hw[0], c.in_hw[1], c.cin, L.ptr(self.w[c.name]),
                                                       L.ptr(self.b[c.name]), c.cout, dst_ptr, c.cout, int(c.relu), 0), c.name)
                    return
                L.check(lib.dvc_plan_add_im2col(plan, src_ptr, nb, c.in_hw[0], c.in_hw[1], c.cin, L.ptr(st.col)), c.name + "/im2col")
                M = nb * c.out_hw[0] * c.out_hw[1]
                L.check(lib.dvc_plan_add_gemm(plan, L.ptr(st.col), L.ptr(self.w[c.name]), L.ptr(self.b[c.name]), M, c.cout,
                                              9 * c.cin, 9 * c.cin, 0, dst_ptr, c.cout, int(c.relu), None, 0), c.name)

        def add_head(h, src_ptr, nb, b0):
            hw = h.grid_hw[0] * h.grid_hw[1]
            w = self.w[f"{h.tap}_mbox"]
            head = L.HeadOut(L.ptr(st.loc) + b0 * st.loc.stride(0) * 4, L.ptr(st.conf) + b0 * st.conf.stride(0) * 4,
                             st.loc.stride(0), st.conf.stride(0), hw, h.loc.cout, h.conf.cout,
                             h.prior_offset * 4, h.prior_offset * g.num_classes)
            st._keep.append(head)
            L.check(lib.dvc_plan_add_gemm(plan, src_ptr, L.ptr(w), L.ptr(self.b[f"{h.tap}_mbox"]), nb * hw, w.shape[0],
                                          w.shape[1], w.shape[1], 1, None, 0, 0, C.byref(head), 0), h.tap + "_mbox")

        assert body[0].name == "conv0" and body[0].k == 3 and body[0].cout == 32, "first layer must be the 3x3/s2 stem"
        # fused dw -> pw launches (csrc/dwpw_fused.cu): "0" none, "1" every eligible block, or a comma list of the
        # pointwise layers' names ("conv3": only that block)
        fuse_env = os.environ.get("DVC_FUSE_DWPW", FUSE_DWPW_DEFAULT)
        fuse = fuse_env not in ("", "0")
        fuse_only = None if fuse_env in ("0", "1", "") else {n.strip() for n in fuse_env.split(",")}

        def fusable(dw, pw):
            return (dw.depthwise and dw.stride == 1 and dw.relu and pw.k == 1 and not pw.depthwise and pw.relu
                    and pw.cin == dw.cout and pw.cout % 64 == 0 and 64 <= pw.cout <= 256 and (dw.cin == 32 or dw.cin % 64 == 0))
        # ---- early layers, per sub-batch, in the small (L2-resident) ping-pong buffers
        for b0 in range(0, B, sub):
            nb = min(sub, B - b0)
            cur, other = st.sA, st.sB
            L.check(lib.dvc_plan_add_preproc_conv0(
                plan, L.ptr(st.frames400_flat), st.stride400, None, nb, st.h4, st.w4,
                L.ptr(st.t_lin.xs), L.ptr(st.t_lin.xq), L.ptr(st.t_lin.ys), L.ptr(st.t_lin.yq), L.ptr(self.w["conv0"]),
                L.ptr(self.b["conv0"]), C.cast(st.mean, C.c_void_p), float(self.cfg.scale), L.ptr(cur), b0), "conv0")
            li = 1
            while li <= split_idx:
                c = body[li]
                nxt = body[li + 1] if li + 1 <= split_idx else None
                if fuse and nxt is not None and fusable(c, nxt) and (fuse_only is None or nxt.name in fuse_only):
                    # EXPERIMENTAL (DVC_FUSE_DWPW=1): depthwise + pointwise of one block as a single launch
                    dst = L.ptr(st.fA) + b0 * per_frame[li + 1] * esz if li + 1 == split_idx else L.ptr(other)
                    L.check(lib.dvc_plan_add_dwpw_fused(plan, L.ptr(cur), nb, c.in_hw[0], c.in_hw[1], c.cin, L.ptr(self.w[c.name]),
                                                        L.ptr(self.b[c.name]), L.ptr(self.w[nxt.name]), L.ptr(self.b[nxt.name]),
                                                        nxt.cout, dst), c.name + "+" + nxt.name)
                    cur, other = other, cur
                    if nxt.top in heads and li + 1 != split_idx:
                        add_head(heads[nxt.top], dst, nb, b0)
                    li += 2
                    continue
                if li == split_idx:
                    dst = L.ptr(st.fA) + b0 * per_frame[li] * esz
                else:
                    dst = L.ptr(other)
                add_conv(c, L.ptr(cur), dst, nb)
                cur, other = other, cur
                if c.top in heads and li != split_idx:
                    add_head(heads[c.top], dst, nb, b0)
                li += 1
        # ---- late layers on the full batch.  Every tap layer (an input of an SSD head) writes into a buffer of its own
        # instead of the ping-pong pair, so nothing overwrites a head's input and the heads can run on the plan's side
        # branch, off the critical path (csrc/plan.cu, dvc_plan_set_head_branch).
        cur = st.fA
        st.taps = {}
        if body[split_idx].top in heads:
            add_head(heads[body[split_idx].top], L.ptr(cur), B, 0)
        for li in range(split_idx + 1, len(body)):
            c = body[li]
            if c.top in heads:
                dst = st.taps[c.name] = torch.empty((B * per_frame[li],), dtype=torch.float16, device=dev)
            else:
                dst = st.fA if cur is not st.fA else st.fB
            add_conv(c, L.ptr(cur), L.ptr(dst), B)
            cur = dst
            if c.top in heads:
                add_head(heads[c.top], L.ptr(cur), B, 0)
        L.check(lib.dvc_plan_add_detection(plan, L.ptr(st.loc), L.ptr(st.conf), L.ptr(self.priors), L.ptr(self.variances), B,
                                           C.byref(st.det_params), None, L.ptr(st.det), L.ptr(st.det_count), L.ptr(st.boxes_i),
                                           L.ptr(st.box_count), None, 0, L.ptr(st.det_scratch),
                                           st.det_scratch.numel()), "detection")
        st.n_ops = lib.dvc_plan_num_ops(plan)
        head_branch = os.environ.get("DVC_HEAD_BRANCH", "1") != "0" and body[split_idx].top not in heads
        L.check(lib.dvc_plan_set_head_branch(plan, 1 if head_branch else 0), "plan_set_head_branch")

    # ------------------------------------------------------------------ execution
    def launches_per_batch(self, st: _ShapeState, annotate: bool = True) -> int:
        return st.n_ops + 2 + 1 + (1 if annotate else 0)   # plan ops (detection = 3 kernels) + K1a resize + K10 annotate

    def _enqueue_resize(self, st: _ShapeState, src: Optional[torch.Tensor] = None):
        """K1a: raw frames (st.raw, or any other device batch of the same shape) -> 400-wide ring."""
        self.K.resize_u8(st.raw if src is None else src, st.t_area, st.frames400_flat, st.stride400)

    def _enqueue_body(self, st: _ShapeState, annotate: bool = True):
        self.L.check(self.lib.dvc_plan_run(st.plan, self.L.stream_ptr()), "plan_run")
        if annotate:
            self._annotate_strided(st)

    def _annotate_strided(self, st):
        L = self.L
        m = st.masks
        nl, lh, lw = m.label_masks.shape
        idh, idw = m.id_mask.shape
        L.check(self.lib.dvc_annotate(L.ptr(st.frames400_flat), st.stride400, None, st.B, st.h4, st.w4, L.ptr(st.boxes_i),
                                      L.ptr(st.box_count), None, L.ptr(m.id_mask), m.id_xy[0], m.id_xy[1], idw, idh,
                                      L.ptr(m.label_masks), nl, m.label_xy[0], m.label_xy[1], lw, lh, L.stream_ptr()), "annotate")

    def bind_remote(self, st: _ShapeState, frames_ptr: int, frame_stride: int, det_ptr: int, det_count_ptr: int,
                    boxes_ptr: int, box_count_ptr: int, flags_ptr: int = 0, flag_value: int = 0):
        """Point this state's plan at frames that live elsewhere (e.g. the ingest GPU's ring, mapped over
        NVLink) and at result rings that live elsewhere (the client GPU).  K1b then pulls its frames with
        peer loads and the detection kernel pushes its rows with peer stores.  Must precede graph capture."""
        L = self.L
        L.check(self.lib.dvc_plan_set_source(st.plan, frames_ptr, frame_stride, None), "plan_set_source")
        L.check(self.lib.dvc_plan_set_sinks(st.plan, det_ptr, det_count_ptr, boxes_ptr, box_count_ptr, flags_ptr or None, None),
                "plan_set_sinks")
        L.check(self.lib.dvc_plan_set_flag_value(st.plan, flag_value), "plan_set_flag_value")
        for key in ("graph_ann", "graph_noann"):
            if hasattr(st, key):
                delattr(st, key)
        st.remote = True

    def bind_mailbox(self, st: _ShapeState, ctl_ptr: int, live_epoch_ptr: int, done_word_ptr: int, done_counter: torch.Tensor,
                     stale_count: Optional[torch.Tensor] = None):
        """Bind this state's plan to one mailbox lane (parallel/mailbox.py, csrc/mailbox.cu): K1b runs only when the
        lane's wait kernel saw a dispatch, K9 publishes the lane's sequence number in the client's done word once the
        whole batch is delivered and rejects the batch when the client's epoch word for this worker has moved on.
        Call after `bind_remote`; must precede graph capture."""
        L = self.L
        assert done_counter.is_cuda and done_counter.dtype == torch.int32
        st.mailbox = (done_counter, stale_count)            # keep the tensors alive
        L.check(self.lib.dvc_plan_set_mailbox(st.plan, ctl_ptr, live_epoch_ptr, done_word_ptr, L.ptr(done_counter),
                                              L.ptr(stale_count) if stale_count is not None else None), "plan_set_mailbox")
        for key in ("graph_ann", "graph_noann"):
            if hasattr(st, key):
                delattr(st, key)

    def enable_timestamps(self, st: _ShapeState, on: bool = True):
        """Device-side trace points: per frame, the %globaltimer nanosecond when K1b starts pulling it and when K9
        has delivered its rows (``st.t_ingest`` / ``st.t_done``, int64 [B] on the device).  The reference only
        prints one wall-clock line per job (worker.py:105,219).  Must precede graph capture of this state."""
        L = self.L
        if on:
            B = st.raw.shape[0]
            st.t_ingest = torch.zeros(B, dtype=torch.int64, device=self.device)
            st.t_done = torch.zeros(B, dtype=torch.int64, device=self.device)
            L.check(self.lib.dvc_plan_set_timestamps(st.plan, L.ptr(st.t_ingest), L.ptr(st.t_done)), "plan_set_timestamps")
        else:
            L.check(self.lib.dvc_plan_set_timestamps(st.plan, None, None), "plan_set_timestamps")
            st.t_ingest = st.t_done = None
        for key in ("graph_ann", "graph_noann"):
            if hasattr(st, key):
                delattr(st, key)

    def frame_latency_us(self, st: _ShapeState) -> np.ndarray:
        """Per-frame device latency (ingest -> delivered) of the last batch run on ``st``, in microseconds."""
        assert getattr(st, "t_done", None) is not None, "call enable_timestamps(st) first"
        return ((st.t_done - st.t_ingest).cpu().numpy().astype(np.float64)) / 1e3

    def set_epoch_gate(self, st: _ShapeState, epoch_now: Optional[torch.Tensor], epoch_expect: Optional[torch.Tensor],
                       stale_count: Optional[torch.Tensor] = None):
        """K14: gate this state's delivery on the membership epoch.  ``epoch_now`` is the int32 device word
        `publish_epoch` / `dvc_epoch_broadcast` writes on this GPU, ``epoch_expect`` the int32 device word the
        dispatcher sets to the epoch a batch is sent under; when they differ the deliver kernel writes nothing
        (no rows, counts or flags) and adds one per frame to ``stale_count``.  Pointers are baked into the plan
        (and any CUDA graph captured later), the values are read at run time.  ``None`` switches the gate off."""
        L = self.L
        for t in (epoch_now, epoch_expect, stale_count):
            assert t is None or (t.is_cuda and t.dtype == torch.int32 and t.numel() >= 1)
        st.epoch_gate = (epoch_now, epoch_expect, stale_count)        # keep the tensors alive
        L.check(self.lib.dvc_plan_set_epoch_gate(st.plan, L.ptr(epoch_now) if epoch_now is not None else None,
                                                 L.ptr(epoch_expect) if epoch_expect is not None else None,
                                                 L.ptr(stale_count) if stale_count is not None else None), "plan_set_epoch_gate")
        for key in ("graph_ann", "graph_noann"):
            if hasattr(st, key):
                delattr(st, key)

    def run(self, st: _ShapeState, annotate: bool = True, use_graph: Optional[bool] = None,
            src: Optional[torch.Tensor] = None, resize: bool = True):
        """Process the frames in ``st.raw`` (or in ``src``, a device batch of the same shape).
        Asynchronous on the current stream.  With CUDA graphs the resize launch stays eager
        (its source pointer may change per call) and everything after it is one graph replay."""
        use_graph = self.cfg.use_cuda_graph if use_graph is None else use_graph
        with torch.cuda.device(self.device):
            self.L.set_device(self.device)
            if resize:
                self._enqueue_resize(st, src)
            if not use_graph:
                self._enqueue_body(st, annotate)
                return
            key = "graph_ann" if annotate else "graph_noann"
            gr = getattr(st, key, None)
            if gr is None:
                # warm up outside capture (first-launch attribute setup), then capture
                with _CAPTURE_LOCK:
                    self._enqueue_body(st, annotate)
                    torch.cuda.synchronize(self.device)
                    gr = torch.cuda.CUDAGraph()
                    # an explicit capture stream on THIS device: torch's default capture stream is created
                    # once per process on whichever device was current first
                    cap_stream = torch.cuda.Stream(device=self.device)
                    with torch.cuda.graph(gr, stream=cap_stream, capture_error_mode="thread_local"):
                        self._enqueue_body(st, annotate)
                    setattr(st, key, gr)
                # the warm-up pass annotated the 400-wide frames in place: regenerate them before the replay
                if resize:
                    self._enqueue_resize(st, src)
            gr.replay()

    def debug_layerwise(self, frames400: torch.Tensor) -> Dict[str, torch.Tensor]:
        """Run the network one native kernel at a time (per-op wrappers, separate output tensors).
        Returns {layer name: NHWC fp16 activation} plus 'loc'/'conf'.  Test/debug aid only."""
        K, g = self.K, self.graph
        B = frames400.shape[0]
        acts: Dict[str, torch.Tensor] = {}
        with torch.cuda.device(self.device):
            body = g.body
            x = K.preproc_conv0(frames400, self.w["conv0"], self.b["conv0"], self.cfg.mean, self.cfg.scale)
            acts["conv0"] = x
            tops = {"conv0": x}
            for c in body[1:]:
                src = tops[c.bottom]
                if c.depthwise:
                    y = K.dwconv3x3(src, self.w[c.name], self.b[c.name], c.stride, c.relu)
                elif c.k == 1:
                    y = K.gemm_f16(src.view(-1, c.cin), self.w[c.name], self.b[c.name], c.relu).view(B, c.out_hw[0], c.out_hw[1], c.cout)
                elif c.cin % 64 == 0 and os.environ.get("DVC_IMPLICIT_CONV", "1") != "0":
                    y = K.conv3x3s2_f16(src, self.w[c.name], self.b[c.name], c.relu)
                else:
                    col = K.im2col3x3s2(src)
                    y = K.gemm_f16(col, self.w[c.name], self.b[c.name], c.relu).view(B, c.out_hw[0], c.out_hw[1], c.cout)
                tops[c.top] = y
                acts[c.name] = y
            loc = torch.zeros((B, g.num_priors * 4), dtype=torch.float32, device=self.device)
            conf = torch.zeros((B, g.num_priors * g.num_classes), dtype=torch.float32, device=self.device)
            for h in g.heads:
                t = tops[h.tap]
                K.gemm_head(t.view(-1, t.shape[-1]), self.w[f"{h.tap}_mbox"], self.b[f"{h.tap}_mbox"], loc, conf,
                            h.grid_hw[0] * h.grid_hw[1], h.loc.cout, h.conf.cout, h.prior_offset, g.num_classes)
            acts["loc"], acts["conf"] = loc, conf
            torch.cuda.synchronize(self.device)
        return acts

    def result(self, st: _ShapeState) -> BatchResult:
        return BatchResult(st.det, st.det_count, st.boxes_i, st.box_count, st.frames400)

    def process(self, frames, requester_id: Optional[str] = None, annotate: bool = True) -> BatchResult:
        """frames: uint8 [B,H,W,3] (numpy / CPU tensor / CUDA tensor). Synchronous convenience API."""
        from .utils.trace import span
        t = torch.as_tensor(frames) if not isinstance(frames, torch.Tensor) else frames
        B, H, W, _ = t.shape
        st = self.state(B, H, W, requester_id)
        with torch.cuda.device(self.device), span("engine.process", getattr(self, "metrics", None), frames=B):
            st.raw.copy_(t, non_blocking=True)         # NVTX range "engine.process" brackets upload, kernels and the wait
            self.run(st, annotate)
            torch.cuda.current_stream().synchronize()
        return self.result(st)

    def process_stream(self, batches, depth: int = 2, annotate: bool = True, fetch_frames: bool = False,
                       requester_id: Optional[str] = None):
        """Pipelined public API: iterate over host batches (uint8 [B,H,W,3], ideally pinned) and yield one
        host-side ``BatchResult`` per batch, in order.  ``depth`` independent slots (own buffers, plan, CUDA
        graph and compute stream each) plus a copy-in and a copy-out stream overlap the host->device copy of
        batch i+1, the kernels of batches i and i-1 (which share the SMs) and the device->host read of batch i-2.
        ``fetch_frames`` also brings the annotated 400-wide frames back (large); detections always come back.

        Lifetime: a yielded ``BatchResult`` is a view of its slot's pinned host buffers.  It is valid until the consumer
        asks for the NEXT item - the slot is then re-used and an asynchronous copy overwrites it - so copy what must
        outlive the iteration (``list(engine.process_stream(...))`` keeps aliases, not results).

        Live sources: an item may be ``None`` ("nothing more for now"): everything in flight is then finished and
        yielded before the generator is asked again, so the tail of a stream is not held back waiting for a batch that
        may never come.  An item may also be ``(frames, requester_id)`` to overlay a per-batch requester id."""
        dev = self.device
        with torch.cuda.device(dev):
            s_in, s_out = (torch.cuda.Stream(device=dev) for _ in range(2))
            run_lanes = [torch.cuda.Stream(device=dev) for _ in range(depth)]    # batches in flight overlap on the SMs
            slots = {}
            pending = []      # (slot record, ready event)

            def finish(rec, ev):
                ev.synchronize()
                return BatchResult(rec["h_det"], rec["h_cnt"], rec["h_boxes"], rec["h_bcnt"],
                                   rec["h_frames"] if fetch_frames else None)

            i = -1
            for batch in batches:
                if batch is None:                       # live source has nothing right now: flush what is in flight
                    while pending:
                        yield finish(*pending.pop(0))
                    continue
                i += 1
                rid = requester_id
                if isinstance(batch, tuple):
                    batch, rid = batch
                t = batch if isinstance(batch, torch.Tensor) else torch.as_tensor(batch)
                B, H, W, _ = t.shape
                k = i % depth
                key = (k, B, H, W, rid)
                rec = slots.get(key)
                if rec is None:
                    st = self.state(B, H, W, rid, slot=1 + k)
                    o1, o2, o3, o4 = st.res_offsets
                    h_res = torch.empty((o4,), dtype=torch.uint8).pin_memory()
                    rec = {"st": st, "free": None, "h_res": h_res,
                           "h_det": h_res[:o1].view(torch.float32).view(B, 100, 7),
                           "h_boxes": h_res[o1:o2].view(torch.int32).view(B, 100, 4),
                           "h_cnt": h_res[o2:o3].view(torch.int32),
                           "h_bcnt": h_res[o3:o4].view(torch.int32),
                           "h_frames": torch.empty((B, st.h4, st.w4, 3), dtype=torch.uint8).pin_memory() if fetch_frames else None}
                    slots[key] = rec
                # a slot is reused only after its previous result has been handed out
                while pending and len(pending) >= depth:
                    yield finish(*pending.pop(0))
                st = rec["st"]
                with torch.cuda.stream(s_in):
                    if rec["free"] is not None:
                        s_in.wait_event(rec["free"])
                    st.raw.copy_(t, non_blocking=True)
                    ev_in = torch.cuda.Event()
                    ev_in.record(s_in)
                s_run = run_lanes[k]
                with torch.cuda.stream(s_run):
                    s_run.wait_event(ev_in)
                    self.run(st, annotate)
                    ev_run = torch.cuda.Event()
                    ev_run.record(s_run)
                with torch.cuda.stream(s_out):
                    s_out.wait_event(ev_run)
                    rec["h_res"].copy_(st.res_flat, non_blocking=True)      # det, boxes and both counts in one copy
                    if fetch_frames:
                        rec["h_frames"].copy_(st.frames400, non_blocking=True)
                    ev_out = torch.cuda.Event()
                    ev_out.record(s_out)
                rec["free"] = ev_out
                pending.append((rec, ev_out))
            while pending:
                yield finish(*pending.pop(0))

    def close(self):
        for st in self._states.values():
            if getattr(st, "plan", None):
                self.lib.dvc_plan_destroy(st.plan)
                st.plan = None
        self._states.clear()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def create_engine(graph: SSDGraph, cfg: Optional[Config] = None, device: Optional[str] = None, requester_id: str = "",
                  max_batch: Optional[int] = None):
    cfg = cfg or Config()
    dev = device or cfg.device
    if dev == "auto":
        dev = "cuda:0" if torch.cuda.is_available() else "cpu"
    if str(dev).startswith("cuda"):
        return NativeEngine(graph, dev, cfg, requester_id, max_batch)
    return CpuEngine(graph, cfg, requester_id)
