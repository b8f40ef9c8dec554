"""Bit-level restatement of the reference's frame preprocessing (numpy, CPU).

Reference pipeline (worker.py:243-246):
    frame400 = imutils.resize(frame, width=400)          # cv2.resize(..., INTER_AREA)
    img300   = cv2.resize(frame400, (300, 300))          # INTER_LINEAR, u8 fixed point
    blob     = cv2.dnn.blobFromImage(img300, 0.007843, (300, 300), 127.5)

Everything here is written from the documented arithmetic of cv2.resize so the CUDA
kernels (csrc/preproc.cu) have an exact integer/float recipe to follow; tests compare
both against cv2 itself with a +-1 LSB tolerance on the u8 stages.

Quirk kept on purpose: a scalar mean of 127.5 is applied to channel 0 only in
OpenCV-Python, so MEAN_COMPAT = (127.5, 0, 0); MEAN_CANONICAL = (127.5,)*3 is the
north-star formula.
"""
from __future__ import annotations

import math
from typing import List, Tuple

import numpy as np

SCALE = 0.007843
MEAN_COMPAT = (127.5, 0.0, 0.0)
MEAN_CANONICAL = (127.5, 127.5, 127.5)
OUT_WIDTH = 400
NET_SIZE = 300

COEF_BITS = 11
COEF_SCALE = 1 << COEF_BITS


def out_size_400(h: int, w: int, width: int = OUT_WIDTH) -> Tuple[int, int]:
    """(h', w') of imutils.resize(frame, width=400): dim = (400, int(h * 400/float(w)))."""
    r = width / float(w)
    return int(h * r), width


# ---------------------------------------------------------------- linear (fixed point)

def linear_taps(ssize: int, dsize: int, area_mode: bool) -> Tuple[np.ndarray, np.ndarray]:
    """Per destination index: source index sx (int32) and the second-tap weight as a
    Q11 short (alpha1); alpha0 = 2048 - alpha1 is NOT assumed - both are rounded
    independently exactly as cv2 does (saturate_cast<short>(c * 2048))."""
    inv_scale = dsize / float(ssize)
    scale = 1.0 / inv_scale
    sx = np.zeros(dsize, np.int32)
    a = np.zeros((dsize, 2), np.int32)
    for dx in range(dsize):
        if not area_mode:
            fx = np.float32((dx + 0.5) * scale - 0.5)
            s = int(math.floor(fx))
            fx = np.float32(fx - s)
        else:
            s = int(math.floor(dx * scale))
            fx = np.float32((dx + 1) - (s + 1) * inv_scale)
            fx = np.float32(0.0) if fx <= 0 else np.float32(fx - math.floor(fx))
        if s < 0:
            fx, s = np.float32(0.0), 0
        if s >= ssize - 1:
            fx, s = np.float32(0.0), ssize - 1
        sx[dx] = s
        c0 = np.float32(1.0) - fx
        a[dx, 0] = int(np.rint(np.float32(c0 * COEF_SCALE)))
        a[dx, 1] = int(np.rint(np.float32(fx * COEF_SCALE)))
    return sx, a


def resize_linear_u8(img: np.ndarray, dh: int, dw: int, area_mode: bool = False) -> np.ndarray:
    """cv2.resize(img, (dw, dh), interpolation=INTER_LINEAR) for uint8 HWC (SIMD variant
    of the vertical pass: ((b0*(r0>>4))>>16 + (b1*(r1>>4))>>16 + 2) >> 2)."""
    sh, sw = img.shape[:2]
    xs, xa = linear_taps(sw, dw, area_mode)
    ys, ya = linear_taps(sh, dh, area_mode)
    src = img.astype(np.int32)
    x1 = np.minimum(xs + 1, sw - 1)
    # horizontal pass -> int rows (Q11)
    rows = src[:, xs, :] * xa[:, 0][None, :, None] + src[:, x1, :] * xa[:, 1][None, :, None]
    y1 = np.minimum(ys + 1, sh - 1)
    r0 = rows[ys] >> 4
    r1 = rows[y1] >> 4
    b0 = ya[:, 0][:, None, None]
    b1 = ya[:, 1][:, None, None]
    out = (((b0 * r0) >> 16) + ((b1 * r1) >> 16) + 2) >> 2
    return np.clip(out, 0, 255).astype(np.uint8)


# ---------------------------------------------------------------- area (decimation)

def area_tab(ssize: int, dsize: int) -> List[List[Tuple[int, float]]]:
    """cv2's computeResizeAreaTab: per destination index a list of (src index, alpha)."""
    scale = ssize / float(dsize)
    tab: List[List[Tuple[int, float]]] = []
    for dx in range(dsize):
        fsx1 = dx * scale
        fsx2 = fsx1 + scale
        cell = min(scale, ssize - fsx1)
        sx1 = int(math.ceil(fsx1))
        sx2 = int(math.floor(fsx2))
        sx2 = min(sx2, ssize - 1)
        sx1 = min(sx1, sx2)
        ent: List[Tuple[int, float]] = []
        if sx1 - fsx1 > 1e-3:
            ent.append((sx1 - 1, float(np.float32((sx1 - fsx1) / cell))))
        for s in range(sx1, sx2):
            ent.append((s, float(np.float32(1.0 / cell))))
        if fsx2 - sx2 > 1e-3:
            ent.append((sx2, float(np.float32(min(min(fsx2 - sx2, 1.0), cell) / cell))))
        tab.append(ent)
    return tab


def area_tab_arrays(ssize: int, dsize: int, max_taps: int = 0):
    """Dense form of area_tab for the GPU: start index, tap count, alpha[dsize, max_taps]."""
    tab = area_tab(ssize, dsize)
    mt = max(max_taps, max(len(e) for e in tab))
    start = np.zeros(dsize, np.int32)
    cnt = np.zeros(dsize, np.int32)
    alpha = np.zeros((dsize, mt), np.float32)
    for d, ent in enumerate(tab):
        if not ent:
            continue
        start[d] = ent[0][0]
        cnt[d] = len(ent)
        for k, (s, a) in enumerate(ent):
            assert s == ent[0][0] + k
            alpha[d, k] = a
    return start, cnt, alpha


def resize_area_u8(img: np.ndarray, dh: int, dw: int) -> np.ndarray:
    """cv2.resize(img, (dw, dh), interpolation=INTER_AREA) for uint8 HWC."""
    sh, sw = img.shape[:2]
    scale_x = sw / float(dw)
    scale_y = sh / float(dh)
    if scale_x < 1.0 or scale_y < 1.0:
        # enlarging: cv2 falls back to the bilinear kernel with "area" coordinates
        return resize_linear_u8(img, dh, dw, area_mode=True)
    isx, isy = int(round(scale_x)), int(round(scale_y))
    eps = np.finfo(np.float64).eps
    if abs(scale_x - isx) < eps and abs(scale_y - isy) < eps:
        s = img.astype(np.int32).reshape(dh, isy, dw, isx, -1).sum(axis=(1, 3))
        if isx == 2 and isy == 2:
            return ((s + 2) >> 2).astype(np.uint8)
        f = (s.astype(np.float32) * np.float32(1.0 / (isx * isy)))
        return np.clip(np.rint(f), 0, 255).astype(np.uint8)
    xstart, xcnt, xalpha = area_tab_arrays(sw, dw)
    ystart, ycnt, yalpha = area_tab_arrays(sh, dh)
    src = img.astype(np.float32)
    c = img.shape[2] if img.ndim == 3 else 1
    src = src.reshape(sh, sw, c)
    # horizontal reduce
    hbuf = np.zeros((sh, dw, c), np.float32)
    for k in range(xalpha.shape[1]):
        idx = np.minimum(xstart + k, sw - 1)
        hbuf += src[:, idx, :] * xalpha[:, k][None, :, None]
    out = np.zeros((dh, dw, c), np.float32)
    for k in range(yalpha.shape[1]):
        idx = np.minimum(ystart + k, sh - 1)
        out += hbuf[idx] * yalpha[:, k][:, None, None]
    out = np.clip(np.rint(out), 0, 255).astype(np.uint8)
    return out.reshape(dh, dw, c) if img.ndim == 3 else out.reshape(dh, dw)


def resize_width_400(frame: np.ndarray, width: int = OUT_WIDTH) -> np.ndarray:
    """imutils.resize(frame, width=400) (reference worker.py:243)."""
    h, w = frame.shape[:2]
    dh, dw = out_size_400(h, w, width)
    return resize_area_u8(frame, dh, dw)


def blob_from_image(img300: np.ndarray, mean=MEAN_COMPAT, scale: float = SCALE) -> np.ndarray:
    """cv2.dnn.blobFromImage(img, scale, size, mean) -> NCHW float32 [1,3,H,W]."""
    x = img300.astype(np.float32)
    m = np.asarray(mean, np.float32).reshape(1, 1, 3)
    x = (x - m) * np.float32(scale)
    return np.ascontiguousarray(x.transpose(2, 0, 1)[None])


def preprocess_frame(frame: np.ndarray, mean=MEAN_COMPAT):
    """Full reference preprocessing. Returns (frame400 u8 HWC, blob f32 NCHW)."""
    f400 = resize_width_400(frame)
    img300 = resize_linear_u8(f400, NET_SIZE, NET_SIZE)
    return f400, blob_from_image(img300, mean)
