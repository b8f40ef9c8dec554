"""Python wrappers over the native kernels (one function per C entry point).

These are thin: they validate shapes/dtypes, pass raw device pointers and the current CUDA
stream, and raise ``NativeError`` on failure.  ``engine.Engine`` uses the plan API instead
(one native call per forward pass); these per-op wrappers exist for unit tests and for ad-hoc
use.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from ..models import preprocess as P
from . import lib as L


def _dev_i32(a, device) -> torch.Tensor:
    return torch.as_tensor(np.ascontiguousarray(a, dtype=np.int32), device=device)


def _dev_f32(a, device) -> torch.Tensor:
    return torch.as_tensor(np.ascontiguousarray(a, dtype=np.float32), device=device)


def align16(n: int) -> int:
    return (n + 15) & ~15


# ------------------------------------------------------------------------------------ K1a
@dataclass
class ResizeTables:
    """Device-resident coefficient tables for one (src h,w) -> (dst h,w) u8 resize."""
    mode: int
    sh: int
    sw: int
    dh: int
    dw: int
    xs: Optional[torch.Tensor] = None
    xc: Optional[torch.Tensor] = None
    xa: Optional[torch.Tensor] = None
    ys: Optional[torch.Tensor] = None
    yc: Optional[torch.Tensor] = None
    ya: Optional[torch.Tensor] = None
    xq: Optional[torch.Tensor] = None
    yq: Optional[torch.Tensor] = None
    ix: int = 0
    iy: int = 0
    x_taps: int = 0
    y_taps: int = 0


def make_area_tables(sh: int, sw: int, dh: int, dw: int, device) -> ResizeTables:
    """Tables for cv2.resize(..., INTER_AREA) (all three regimes, see models/preprocess.py)."""
    scale_x, scale_y = sw / float(dw), sh / float(dh)
    if scale_x < 1.0 or scale_y < 1.0:
        return make_linear_tables(sh, sw, dh, dw, device, area_mode=True)
    isx, isy = int(round(scale_x)), int(round(scale_y))
    eps = np.finfo(np.float64).eps
    if abs(scale_x - isx) < eps and abs(scale_y - isy) < eps:
        return ResizeTables(1, sh, sw, dh, dw, ix=isx, iy=isy)
    xs, xc, xa = P.area_tab_arrays(sw, dw)
    ys, yc, ya = P.area_tab_arrays(sh, dh)
    return ResizeTables(0, sh, sw, dh, dw, xs=_dev_i32(xs, device), xc=_dev_i32(xc, device), xa=_dev_f32(xa, device),
                        ys=_dev_i32(ys, device), yc=_dev_i32(yc, device), ya=_dev_f32(ya, device),
                        x_taps=xa.shape[1], y_taps=ya.shape[1])


def make_linear_tables(sh: int, sw: int, dh: int, dw: int, device, area_mode: bool = False) -> ResizeTables:
    xs, xq = P.linear_taps(sw, dw, area_mode)
    ys, yq = P.linear_taps(sh, dh, area_mode)
    return ResizeTables(2, sh, sw, dh, dw, xs=_dev_i32(xs, device), ys=_dev_i32(ys, device),
                        xq=_dev_i32(xq, device), yq=_dev_i32(yq, device))


def resize_u8_indirect(src_word: int, src_frame_stride: int, B: int, t: ResizeTables, dst: torch.Tensor,
                       dst_frame_stride: int, gate_now: int, gate_expect: int, stream=None) -> None:
    """K1a for a mailbox lane: the source base pointer is read by the kernel from the device word at ``src_word``
    (``&lane_ctl.src``) and the pull is skipped unless the two gate words (device addresses) are equal."""
    lib = L.load()
    L.check(lib.dvc_resize_u8_indirect(src_word, src_frame_stride, B, t.sh, t.sw, L.ptr(dst), dst_frame_stride, t.dh, t.dw,
                                       t.mode, L.ptr(t.xs), L.ptr(t.xc), L.ptr(t.xa), t.x_taps, L.ptr(t.ys), L.ptr(t.yc),
                                       L.ptr(t.ya), t.y_taps, L.ptr(t.xq), L.ptr(t.yq), t.ix, t.iy, gate_now, gate_expect,
                                       L.stream_ptr(stream)), "resize_u8_indirect")


def resize_u8(src: torch.Tensor, t: ResizeTables, dst: Optional[torch.Tensor] = None, dst_frame_stride: int = 0,
              stream=None, epoch_gate=None) -> torch.Tensor:
    """cv2 INTER_AREA on the device.  ``epoch_gate=(epoch_now, epoch_expect)`` (two int32 device words) makes the
    kernel skip the read of ``src`` - the NVLink pull in the multi-GPU path - when the membership epoch moved on."""
    assert src.dtype == torch.uint8 and src.is_cuda and src.dim() == 4 and src.shape[3] == 3
    B = src.shape[0]
    assert (src.shape[1], src.shape[2]) == (t.sh, t.sw)
    assert src.stride()[1:] == (t.sw * 3, 3, 1), "frames must be dense HWC (a frame pitch is allowed)"
    if dst is None:
        dst = torch.empty((B, t.dh, t.dw, 3), dtype=torch.uint8, device=src.device)
        dst_frame_stride = t.dh * t.dw * 3
    lib = L.load()
    args = (L.ptr(src), src.stride(0), B, t.sh, t.sw, L.ptr(dst), dst_frame_stride, t.dh, t.dw, t.mode,
            L.ptr(t.xs), L.ptr(t.xc), L.ptr(t.xa), t.x_taps, L.ptr(t.ys), L.ptr(t.yc), L.ptr(t.ya),
            t.y_taps, L.ptr(t.xq), L.ptr(t.yq), t.ix, t.iy)
    if epoch_gate is None:
        L.check(lib.dvc_resize_u8(*args, L.stream_ptr(stream)), "resize_u8")
    else:
        L.check(lib.dvc_resize_u8_gated(*args, L.ptr(epoch_gate[0]), L.ptr(epoch_gate[1]), L.stream_ptr(stream)), "resize_u8_gated")
    return dst


# ------------------------------------------------------------------------------------ K1b
def preproc_conv0(frames400: torch.Tensor, w0_27x32: torch.Tensor, b0: torch.Tensor, mean=P.MEAN_COMPAT,
                  scale: float = P.SCALE, tables: Optional[ResizeTables] = None, frame_index=None, B: Optional[int] = None,
                  stream=None) -> torch.Tensor:
    assert frames400.dtype == torch.uint8 and frames400.is_cuda and frames400.dim() == 4
    sh, sw = frames400.shape[1], frames400.shape[2]
    if tables is None:
        tables = make_linear_tables(sh, sw, P.NET_SIZE, P.NET_SIZE, frames400.device)
    B = frames400.shape[0] if B is None else B
    out = torch.empty((B, 150, 150, 32), dtype=torch.float16, device=frames400.device)
    m = (C.c_float * 3)(*[float(v) for v in mean])
    lib = L.load()
    L.check(lib.dvc_preproc_conv0(L.ptr(frames400), frames400.stride(0), L.ptr(frame_index), B, sh, sw, L.ptr(tables.xs),
                                  L.ptr(tables.xq), L.ptr(tables.ys), L.ptr(tables.yq), L.ptr(w0_27x32), L.ptr(b0),
                                  C.cast(m, C.c_void_p), float(scale), L.ptr(out), L.stream_ptr(stream)), "preproc_conv0")
    return out


# ------------------------------------------------------------------------------------ K3 / K4 / K2
def dwconv3x3(x: torch.Tensor, w9c: torch.Tensor, b: torch.Tensor, stride: int, relu: bool = True, stream=None):
    assert x.dtype == torch.float16 and x.is_cuda and x.is_contiguous() and x.dim() == 4
    B, H, W, Cc = x.shape
    Ho, Wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    out = torch.empty((B, Ho, Wo, Cc), dtype=torch.float16, device=x.device)
    L.check(L.load().dvc_dwconv3x3(L.ptr(x), B, H, W, Cc, stride, L.ptr(w9c), L.ptr(b), int(relu), L.ptr(out),
                                   L.stream_ptr(stream)), "dwconv3x3")
    return out


def im2col3x3s2(x: torch.Tensor, stream=None) -> torch.Tensor:
    assert x.dtype == torch.float16 and x.is_cuda and x.is_contiguous() and x.dim() == 4
    B, H, W, Cc = x.shape
    Ho, Wo = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    out = torch.empty((B * Ho * Wo, 9 * Cc), dtype=torch.float16, device=x.device)
    L.check(L.load().dvc_im2col3x3s2(L.ptr(x), B, H, W, Cc, L.ptr(out), L.stream_ptr(stream)), "im2col3x3s2")
    return out


def conv3x3s2_f16(x: torch.Tensor, wt: torch.Tensor, bias: torch.Tensor, relu: bool = True, block_n: int = 0, stream=None) -> torch.Tensor:
    """Dense 3x3 / stride-2 / pad-1 convolution (SSD extras, prototxt:922-1146) as an implicit GEMM: NHWC fp16 in,
    wt [Cout, 9*Cin] in (kh, kw, cin) order, no patch buffer (the GEMM's activation tiles are im2col-mode TMA loads)."""
    assert x.dtype == torch.float16 and x.is_cuda and x.is_contiguous() and x.dim() == 4
    B, H, W, Cc = x.shape
    assert wt.dtype == torch.float16 and wt.shape[1] == 9 * Cc and wt.is_contiguous() and bias.dtype == torch.float32
    Ho, Wo = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    N = wt.shape[0]
    out = torch.empty((B, Ho, Wo, N), dtype=torch.float16, device=x.device)
    L.check(L.load().dvc_conv3x3s2_f16(L.ptr(x), B, H, W, Cc, L.ptr(wt), L.ptr(bias), N, L.ptr(out), N, int(relu), block_n,
                                       L.stream_ptr(stream)), "conv3x3s2_f16")
    return out


def debug_im2col_tile(x: torch.Tensor, m0: int, kb: int):
    """Operand tile [128, 64] the implicit convolution's producer fetches for (first output pixel m0, K block kb),
    and the load status (0 = ok).  Test hook for the im2col tensor map."""
    assert x.dtype == torch.float16 and x.is_cuda and x.is_contiguous() and x.dim() == 4
    B, H, W, Cc = x.shape
    tile = torch.zeros((128, 64), dtype=torch.float16, device=x.device)
    status = torch.full((1,), -1, dtype=torch.int32, device=x.device)
    L.check(L.load().dvc_debug_im2col_tile(L.ptr(x), B, H, W, Cc, int(m0), int(kb), L.ptr(tile), L.ptr(status),
                                           L.stream_ptr(None)), "debug_im2col_tile")
    torch.cuda.synchronize(x.device)
    return tile, int(status.item())


def gemm_f16(a: torch.Tensor, wt: torch.Tensor, bias: torch.Tensor, relu: bool = True, block_n: int = 0,
             out: Optional[torch.Tensor] = None, stream=None) -> torch.Tensor:
    """out[M,N] = act(a[M,K] @ wt[N,K]^T + bias) on tcgen05 tensor cores (fp16 in, fp32 accumulate)."""
    assert a.dtype == torch.float16 and wt.dtype == torch.float16 and bias.dtype == torch.float32
    assert a.is_cuda and a.dim() == 2 and wt.dim() == 2 and a.shape[1] == wt.shape[1] and a.stride(1) == 1
    M, K = a.shape
    N = wt.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=torch.float16, device=a.device)
    L.check(L.load().dvc_gemm_f16(L.ptr(a), L.ptr(wt), L.ptr(bias), M, N, K, a.stride(0), 0, L.ptr(out), out.stride(0),
                                  int(relu), None, block_n, L.stream_ptr(stream)), "gemm_f16")
    return out


def gemm_head(a: torch.Tensor, wt: torch.Tensor, bias: torch.Tensor, loc: torch.Tensor, conf: torch.Tensor, hw: int,
              n_loc: int, n_conf: int, prior_off: int, num_classes: int, block_n: int = 0, stream=None) -> None:
    """SSD head GEMM with the fused [loc ; conf] weights, split-stored at the concat offsets."""
    M, K = a.shape
    N = wt.shape[0]
    head = L.HeadOut(L.ptr(loc), L.ptr(conf), loc.stride(0), conf.stride(0), hw, n_loc, n_conf, prior_off * 4,
                     prior_off * num_classes)
    L.check(L.load().dvc_gemm_f16(L.ptr(a), L.ptr(wt), L.ptr(bias), M, N, K, a.stride(0), 1, None, 0, 0, C.byref(head),
                                  block_n, L.stream_ptr(stream)), "gemm_head")


def dwpw_fused(x: torch.Tensor, wdw_9c: torch.Tensor, bdw: torch.Tensor, wpw: torch.Tensor, bpw: torch.Tensor, stride: int = 1,
               out: Optional[torch.Tensor] = None, stream=None) -> torch.Tensor:
    """Depthwise 3x3 (stride 1 or 2) + ReLU fused into the pointwise 1x1 + ReLU of the same block (csrc/dwpw_fused.cu).
    x [B,H,W,C_in] fp16 -> [B,Ho,Wo,C_out] fp16."""
    assert x.dtype == torch.float16 and x.is_cuda and x.dim() == 4 and x.is_contiguous()
    assert wpw.dtype == torch.float16 and wpw.is_contiguous() and wpw.shape[1] == x.shape[3]
    B, H, W, C_in = x.shape
    C_out = wpw.shape[0]
    Ho, Wo = (H + 2 - 3) // stride + 1, (W + 2 - 3) // stride + 1
    if out is None:
        out = torch.empty((B, Ho, Wo, C_out), dtype=torch.float16, device=x.device)
    L.check(L.load().dvc_dwpw_fused(L.ptr(x), B, H, W, C_in, stride, L.ptr(wdw_9c), L.ptr(bdw), L.ptr(wpw), L.ptr(bpw), C_out,
                                    L.ptr(out), L.stream_ptr(stream)), "dwpw_fused")
    return out


# ------------------------------------------------------------------------------------ K6..K10
def priorbox(specs, img_hw: Tuple[int, int], device) -> Tuple[torch.Tensor, torch.Tensor]:
    total = sum(s.count for s in specs)
    boxes = torch.empty((total, 4), dtype=torch.float32, device=device)
    var = torch.empty((total, 4), dtype=torch.float32, device=device)
    off = 0
    lib = L.load()
    keep = []
    for s in specs:
        wh = _dev_f32(np.asarray(s.anchor_wh(), np.float32).reshape(-1), device)
        keep.append(wh)
        v4 = (C.c_float * 4)(*[float(v) for v in s.variance])
        L.check(lib.dvc_priorbox(L.ptr(boxes), L.ptr(var), off, s.grid_h, s.grid_w, img_hw[0], img_hw[1], L.ptr(wh),
                                 s.anchors, float(s.offset), C.cast(v4, C.c_void_p), int(s.clip), L.stream_ptr()), "priorbox")
        off += s.count
    torch.cuda.current_stream().synchronize()
    return boxes, var


def det_params(det, num_priors: int, out_w: int, out_h: int, app_conf: float = 0.2, consider: Sequence[int] = (15,)):
    mask = 0
    for c in consider:
        mask |= 1 << int(c)
    return L.DetParams(num_priors, det.num_classes, det.background_label_id, det.confidence_threshold, det.nms_threshold,
                       det.top_k, det.keep_top_k, app_conf, mask, out_w, out_h)


def detection_scratch_bytes(B: int, num_priors: int, num_classes: int) -> int:
    return int(L.load().dvc_detection_scratch_bytes(B, num_priors, num_classes))


def detection_output(loc: torch.Tensor, conf: torch.Tensor, priors: torch.Tensor, variances: torch.Tensor, params,
                     scratch: Optional[torch.Tensor] = None, slot_index=None, flags=None, flag_value: int = 0, outs=None,
                     stream=None):
    B = loc.shape[0]
    dev = loc.device
    if outs is None:
        det = torch.empty((B, 100, 7), dtype=torch.float32, device=dev)
        det_count = torch.empty((B,), dtype=torch.int32, device=dev)
        boxes_i = torch.zeros((B, 100, 4), dtype=torch.int32, device=dev)
        box_count = torch.empty((B,), dtype=torch.int32, device=dev)
    else:
        det, det_count, boxes_i, box_count = outs
    need = detection_scratch_bytes(B, params.num_priors, params.num_classes)
    if scratch is None:
        scratch = torch.empty((need,), dtype=torch.uint8, device=dev)
    L.check(L.load().dvc_detection_output(L.ptr(loc), L.ptr(conf), L.ptr(priors), L.ptr(variances), B, C.byref(params),
                                          L.ptr(slot_index), L.ptr(det), L.ptr(det_count), L.ptr(boxes_i), L.ptr(box_count),
                                          L.ptr(flags), flag_value, L.ptr(scratch), scratch.numel() * scratch.element_size(),
                                          L.stream_ptr(stream)), "detection_output")
    return det, det_count, boxes_i, box_count


@dataclass
class TextMasks:
    id_mask: torch.Tensor
    id_xy: Tuple[int, int]
    label_masks: torch.Tensor
    label_xy: Tuple[int, int]


def annotate(frames400: torch.Tensor, boxes_i: torch.Tensor, box_count: torch.Tensor, masks: Optional[TextMasks],
             frame_index=None, slot_index=None, B: Optional[int] = None, stream=None) -> torch.Tensor:
    assert frames400.dtype == torch.uint8 and frames400.is_cuda and frames400.dim() == 4
    h, w = frames400.shape[1], frames400.shape[2]
    B = frames400.shape[0] if B is None else B
    if masks is None:
        idm, ixy, lm, lxy, nl, idw, idh, lw, lh = None, (0, 0), None, (0, 0), 0, 0, 0, 0, 0
    else:
        idm, ixy, lm, lxy = masks.id_mask, masks.id_xy, masks.label_masks, masks.label_xy
        nl, lh, lw = lm.shape
        idh, idw = idm.shape
    L.check(L.load().dvc_annotate(L.ptr(frames400), frames400.stride(0), L.ptr(frame_index), B, h, w, L.ptr(boxes_i),
                                  L.ptr(box_count), L.ptr(slot_index), L.ptr(idm), ixy[0], ixy[1], idw, idh, L.ptr(lm), nl,
                                  lxy[0], lxy[1], lw, lh, L.stream_ptr(stream)), "annotate")
    return frames400
