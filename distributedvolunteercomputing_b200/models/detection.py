"""PriorBox and DetectionOutput: exact numpy restatements (CPU oracle + CPU fallback).

Reference semantics: MobileNetSSD_deploy.prototxt.txt:1239-1256 (PriorBox, x6),
:1872-1880 (Softmax) and :1890-1912 (DetectionOutput), executed by cv2.dnn at
worker.py:249.  Formulas are the ones validated against cv2.dnn in SURVEY.md §2.4.

The CUDA kernels csrc/priorbox.cu and csrc/detect.cu implement the same recipe;
ordering is made deterministic (score desc, then prior index asc, then class asc) so
CPU and GPU agree exactly and set-comparison against cv2 is stable.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np

from ..caffe_io import NetDef, conv_table

VOC_CLASSES = ["background", "aeroplane", "bicycle", "bird", "boat", "bottle", "bus", "car", "cat", "chair",
               "cow", "diningtable", "dog", "horse", "motorbike", "person", "pottedplant", "sheep", "sofa",
               "train", "tvmonitor"]  # worker.py:188-191


@dataclass
class PriorSpec:
    tap: str
    grid_h: int
    grid_w: int
    min_size: float
    max_size: Optional[float]
    aspect_ratios: List[float]          # as listed in the prototxt (before flip)
    flip: bool = True
    clip: bool = False
    variance: Tuple[float, float, float, float] = (0.1, 0.1, 0.2, 0.2)
    offset: float = 0.5
    step: Optional[float] = None

    @property
    def full_ratios(self) -> List[float]:
        out: List[float] = []
        for ar in self.aspect_ratios:
            if any(abs(ar - o) < 1e-6 for o in out) or abs(ar - 1.0) < 1e-6:
                continue
            out.append(ar)
            if self.flip:
                out.append(1.0 / ar)
        return out

    @property
    def anchors(self) -> int:
        return 1 + (1 if self.max_size is not None else 0) + len(self.full_ratios)

    @property
    def count(self) -> int:
        return self.grid_h * self.grid_w * self.anchors

    def anchor_wh(self) -> List[Tuple[float, float]]:
        """Box (w, h) in pixels, in the order cv2.dnn / caffe-ssd emits them."""
        wh = [(self.min_size, self.min_size)]
        if self.max_size is not None:
            s = math.sqrt(self.min_size * self.max_size)
            wh.append((s, s))
        for ar in self.full_ratios:
            r = math.sqrt(ar)
            wh.append((self.min_size * r, self.min_size / r))
        return wh


@dataclass
class DetectionParams:
    num_classes: int = 21
    background_label_id: int = 0
    nms_threshold: float = 0.45
    top_k: int = 100
    keep_top_k: int = 100
    confidence_threshold: float = 0.25
    share_location: bool = True
    code_type: str = "CENTER_SIZE"
    variance_encoded_in_target: bool = False


def prior_specs(net: NetDef) -> List[PriorSpec]:
    shapes = {c.top: c.out_hw for c in conv_table(net)}
    specs = []
    for l in net.of_type("PriorBox"):
        p = l.p("prior_box_param")
        gh, gw = shapes[l.bottoms[0]]
        var = tuple(float(v) for v in p.all("variance")) or (0.1,)
        if len(var) == 1:
            var = var * 4
        specs.append(PriorSpec(
            tap=l.bottoms[0], grid_h=gh, grid_w=gw,
            min_size=float(p.one("min_size")),
            max_size=(float(p.one("max_size")) if p.one("max_size") is not None else None),
            aspect_ratios=[float(a) for a in p.all("aspect_ratio")],
            flip=bool(p.one("flip", True)), clip=bool(p.one("clip", False)),
            variance=var, offset=float(p.one("offset", 0.5)),
            step=(float(p.one("step")) if p.one("step") is not None else None)))
    # honour the Concat order of mbox_priorbox if present
    for l in net.of_type("Concat"):
        if l.name == "mbox_priorbox":
            order = [b.replace("_mbox_priorbox", "") for b in l.bottoms]
            specs.sort(key=lambda s: order.index(s.tap))
    return specs


def detection_params(net: NetDef) -> DetectionParams:
    ls = net.of_type("DetectionOutput")
    if not ls:
        return DetectionParams()
    p = ls[0].p("detection_output_param")
    nms = p.one("nms_param") or {}
    nms_get = nms.one if hasattr(nms, "one") else (lambda k, d=None: d)
    return DetectionParams(
        num_classes=int(p.one("num_classes", 21)),
        background_label_id=int(p.one("background_label_id", 0)),
        nms_threshold=float(nms_get("nms_threshold", 0.45)),
        top_k=int(nms_get("top_k", 100)),
        keep_top_k=int(p.one("keep_top_k", 100)),
        confidence_threshold=float(p.one("confidence_threshold", 0.25)),
        share_location=bool(p.one("share_location", True)),
        code_type=str(p.one("code_type", "CENTER_SIZE")),
        variance_encoded_in_target=bool(p.one("variance_encoded_in_target", False)))


def priorbox(specs: Sequence[PriorSpec], img_h: int = 300, img_w: int = 300) -> Tuple[np.ndarray, np.ndarray]:
    """Return (boxes [P,4] normalised xmin,ymin,xmax,ymax ; variances [P,4]) float32."""
    boxes: List[np.ndarray] = []
    vars_: List[np.ndarray] = []
    for s in specs:
        step_w = s.step if s.step is not None else img_w / float(s.grid_w)
        step_h = s.step if s.step is not None else img_h / float(s.grid_h)
        wh = np.asarray(s.anchor_wh(), np.float32)                       # [A,2]
        ys, xs = np.meshgrid(np.arange(s.grid_h, dtype=np.float32), np.arange(s.grid_w, dtype=np.float32),
                             indexing="ij")
        cx = ((xs + np.float32(s.offset)) * np.float32(step_w))[..., None]  # [H,W,1]
        cy = ((ys + np.float32(s.offset)) * np.float32(step_h))[..., None]
        bw = wh[:, 0][None, None, :] * np.float32(0.5)
        bh = wh[:, 1][None, None, :] * np.float32(0.5)
        b = np.stack([(cx - bw) / np.float32(img_w), (cy - bh) / np.float32(img_h),
                      (cx + bw) / np.float32(img_w), (cy + bh) / np.float32(img_h)], axis=-1)  # [H,W,A,4]
        if s.clip:
            b = np.clip(b, 0.0, 1.0)
        boxes.append(b.reshape(-1, 4).astype(np.float32))
        vars_.append(np.tile(np.asarray(s.variance, np.float32), (s.count, 1)))
    return np.concatenate(boxes, 0), np.concatenate(vars_, 0)


def softmax(x: np.ndarray, axis: int = -1) -> np.ndarray:
    m = x.max(axis=axis, keepdims=True)
    e = np.exp(x - m)
    return e / e.sum(axis=axis, keepdims=True)


def decode_boxes(loc: np.ndarray, priors: np.ndarray, variances: np.ndarray) -> np.ndarray:
    """CENTER_SIZE decode, variance not encoded in target, no clipping. loc [...,P,4]."""
    pw = priors[:, 2] - priors[:, 0]
    ph = priors[:, 3] - priors[:, 1]
    pcx = (priors[:, 0] + priors[:, 2]) * np.float32(0.5)
    pcy = (priors[:, 1] + priors[:, 3]) * np.float32(0.5)
    cx = variances[:, 0] * loc[..., 0] * pw + pcx
    cy = variances[:, 1] * loc[..., 1] * ph + pcy
    w = np.exp(variances[:, 2] * loc[..., 2]) * pw
    h = np.exp(variances[:, 3] * loc[..., 3]) * ph
    return np.stack([cx - w * np.float32(0.5), cy - h * np.float32(0.5),
                     cx + w * np.float32(0.5), cy + h * np.float32(0.5)], axis=-1).astype(np.float32)


def iou(a: np.ndarray, b: np.ndarray) -> float:
    if b[0] > a[2] or b[2] < a[0] or b[1] > a[3] or b[3] < a[1]:
        return 0.0
    ix0, iy0 = max(a[0], b[0]), max(a[1], b[1])
    ix1, iy1 = min(a[2], b[2]), min(a[3], b[3])
    inter = float(np.float32(ix1 - ix0) * np.float32(iy1 - iy0))
    aa = float(np.float32(a[2] - a[0]) * np.float32(a[3] - a[1]))
    bb = float(np.float32(b[2] - b[0]) * np.float32(b[3] - b[1]))
    return inter / (aa + bb - inter) if inter > 0 else 0.0


def detection_output(loc: np.ndarray, conf_prob: np.ndarray, priors: np.ndarray, variances: np.ndarray,
                     params: DetectionParams = DetectionParams()) -> np.ndarray:
    """One image. loc [P,4], conf_prob [P,C] (already softmaxed). Returns [n,7] rows
    (image_id=0, label, conf, xmin, ymin, xmax, ymax) sorted by conf desc."""
    boxes = decode_boxes(loc.astype(np.float32), priors, variances)
    kept: List[Tuple[float, int, int]] = []   # (score, prior idx, label)
    for c in range(params.num_classes):
        if c == params.background_label_id:
            continue
        sc = conf_prob[:, c]
        idx = np.nonzero(sc > params.confidence_threshold)[0]
        if idx.size == 0:
            continue
        order = idx[np.lexsort((idx, -sc[idx]))]     # score desc, index asc
        if params.top_k > 0:
            order = order[: params.top_k]
        keep: List[int] = []
        for i in order:
            ok = True
            for j in keep:
                if iou(boxes[i], boxes[j]) > params.nms_threshold:
                    ok = False
                    break
            if ok:
                keep.append(int(i))
        kept.extend((float(sc[i]), i, c) for i in keep)
    kept.sort(key=lambda t: (-t[0], t[1], t[2]))
    if params.keep_top_k > 0:
        kept = kept[: params.keep_top_k]
    out = np.zeros((len(kept), 7), np.float32)
    for r, (s, i, c) in enumerate(kept):
        out[r] = (0, c, s, *boxes[i])
    return out


def match_detections(a: np.ndarray, b: np.ndarray, tol_conf: float = 2e-3, tol_box: float = 2e-3):
    """Order-insensitive comparison of two [n,7] detection sets (zero-conf rows ignored).
    Returns (matched, only_a, only_b)."""
    a = a[a[:, 2] > 0]
    b = b[b[:, 2] > 0]
    used = np.zeros(len(b), bool)
    matched = 0
    only_a = []
    for r in a:
        hit = -1
        for j, q in enumerate(b):
            if used[j] or int(q[1]) != int(r[1]):
                continue
            if abs(q[2] - r[2]) <= tol_conf and np.all(np.abs(q[3:7] - r[3:7]) <= tol_box):
                hit = j
                break
        if hit >= 0:
            used[hit] = True
            matched += 1
        else:
            only_a.append(r)
    only_b = [q for j, q in enumerate(b) if not used[j]]
    return matched, only_a, only_b


def app_filter(dets: np.ndarray, w: int, h: int, min_conf: float = 0.2, consider=(15,)):
    """Reference post-filter (worker.py:253-269): conf > 0.2, class in CONSIDER, box *
    [w,h,w,h] truncated toward zero. Returns (int boxes [m,4], count)."""
    boxes = []
    for r in dets:
        if r[2] > min_conf and int(r[1]) in consider:
            b = (r[3:7] * np.array([w, h, w, h], np.float32)).astype("int")
            boxes.append(b)
    arr = np.asarray(boxes, np.int32).reshape(-1, 4)
    return arr, len(boxes)
