"""Annotation helpers: CPU oracle (cv2) and pre-rendered text masks for the CUDA raster (K10).

Reference overlay (worker.py:268-276): blue rectangles (thickness 2) around every kept
detection, the requester id in red at (10, 25) and "person: N" in green at (10, h - 20),
FONT_HERSHEY_SIMPLEX scale 0.5 thickness 2.
"""
from __future__ import annotations

from typing import Sequence, Tuple

import numpy as np

try:  # cv2 is only needed for text rendering / the CPU path
    import cv2
except Exception:  # pragma: no cover
    cv2 = None

BOX_COLOR = (255, 0, 0)
ID_COLOR = (0, 0, 255)
LABEL_COLOR = (0, 255, 0)


def label_text(counts: dict) -> str:
    return ", ".join("{}: {}".format(k, v) for k, v in counts.items())


def annotate_cpu(frame400: np.ndarray, boxes_i: np.ndarray, requester_id: str, label: str) -> np.ndarray:
    """In-place overlay with the reference's own drawing calls. boxes_i: [n,4] ints."""
    h = frame400.shape[0]
    for (x0, y0, x1, y1) in np.asarray(boxes_i).reshape(-1, 4):
        cv2.rectangle(frame400, (int(x0), int(y0)), (int(x1), int(y1)), BOX_COLOR, 2)
    cv2.putText(frame400, requester_id, (10, 25), cv2.FONT_HERSHEY_SIMPLEX, 0.5, ID_COLOR, 2)
    cv2.putText(frame400, label, (10, h - 20), cv2.FONT_HERSHEY_SIMPLEX, 0.5, LABEL_COLOR, 2)
    return frame400


def _render_mask(text: str, org: Tuple[int, int], h: int, w: int) -> np.ndarray:
    img = np.zeros((h, w), np.uint8)
    cv2.putText(img, text, org, cv2.FONT_HERSHEY_SIMPLEX, 0.5, 255, 2)
    return img


def text_masks(requester_id: str, h: int, w: int, names: Sequence[str] = ("person",), max_count: int = 100):
    """Pre-render the two text overlays once (cv2 on the host) as tight binary masks.

    Returns (id_mask [ih,iw] u8, (ix,iy), label_masks [max_count+1, lh, lw] u8, (lx,ly)).
    Only a single-class CONSIDER set maps to a count-indexed label atlas; with several
    classes the label atlas is rendered for the first name (the GPU path then shows its count).
    """
    idm = _render_mask(requester_id, (10, 25), h, w)
    ys, xs = np.nonzero(idm)
    if len(ys) == 0:
        id_mask, id_xy = np.zeros((1, 1), np.uint8), (0, 0)
    else:
        id_xy = (int(xs.min()), int(ys.min()))
        id_mask = (idm[ys.min(): ys.max() + 1, xs.min(): xs.max() + 1] > 0).astype(np.uint8)
    full = [_render_mask(label_text({names[0]: n}), (10, h - 20), h, w) for n in range(max_count + 1)]
    stack = np.stack(full)
    nz = np.nonzero(stack.max(axis=0))
    if len(nz[0]) == 0:
        return id_mask, id_xy, np.zeros((max_count + 1, 1, 1), np.uint8), (0, 0)
    y0, y1, x0, x1 = nz[0].min(), nz[0].max() + 1, nz[1].min(), nz[1].max() + 1
    label_masks = (stack[:, y0:y1, x0:x1] > 0).astype(np.uint8)
    return id_mask, id_xy, np.ascontiguousarray(label_masks), (int(x0), int(y0))
