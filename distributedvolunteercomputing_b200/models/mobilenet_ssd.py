"""MobileNet-SSD 300x300: graph description built from the prototxt + fp32 torch model.

The reference runs this network through cv2.dnn (worker.py:194,248-249); topology is
MobileNetSSD_deploy.prototxt.txt:1-1912 (47 convolutions, 6 taps, 1917 priors).

``SSDGraph`` is the framework's model IR: a topologically ordered list of fused
conv(+ReLU) nodes plus head/prior/detection metadata.  It is consumed by
  * ``TorchSSD``   – plain PyTorch fp32/fp16 module (oracle for kernel tests, CPU
                     fallback, and the cuDNN arm of baseline (b)),
  * ``engine.Engine`` – which lowers it to the native sm_100a plan (csrc/plan.cu).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

from ..caffe_io import (ConvSpec, NetDef, conv_table, fold_batchnorm, load_netdef, mobilenet_ssd_prototxt,
                        random_init_weights, read_caffemodel)
from .detection import DetectionParams, PriorSpec, detection_params, prior_specs, priorbox


@dataclass
class HeadSpec:
    tap: str
    loc: ConvSpec
    conf: ConvSpec
    anchors: int
    prior_offset: int      # first prior index of this tap in the concatenated outputs
    grid_hw: Tuple[int, int]


class SSDGraph:
    def __init__(self, net: NetDef, weights: Dict[str, List[np.ndarray]]):
        self.net = net
        self.weights = fold_batchnorm(net, weights)
        self.convs: List[ConvSpec] = conv_table(net)
        self.by_name: Dict[str, ConvSpec] = {c.name: c for c in self.convs}
        missing = [c.name for c in self.convs if c.name not in self.weights]
        if missing:
            raise KeyError(f"caffemodel is missing blobs for layers: {missing[:4]}{'...' if len(missing) > 4 else ''}")
        for c in self.convs:
            w = self.weights[c.name][0]
            exp = (c.cout, c.cin // c.group, c.k, c.k)
            if tuple(w.shape) != exp:
                self.weights[c.name][0] = w.reshape(exp)
        self.priors_spec: List[PriorSpec] = prior_specs(net)
        self.det: DetectionParams = detection_params(net)
        self.num_classes = self.det.num_classes
        self.input_hw = (net.input_shape[2], net.input_shape[3]) if len(net.input_shape) == 4 else (300, 300)
        # heads in mbox concat order
        self.heads: List[HeadSpec] = []
        off = 0
        for ps in self.priors_spec:
            loc = self.by_name[f"{ps.tap}_mbox_loc"]
            conf = self.by_name[f"{ps.tap}_mbox_conf"]
            assert loc.cout == ps.anchors * 4 and conf.cout == ps.anchors * self.num_classes, (ps.tap, loc.cout, conf.cout)
            self.heads.append(HeadSpec(ps.tap, loc, conf, ps.anchors, off, (ps.grid_h, ps.grid_w)))
            off += ps.count
        self.num_priors = off
        head_names = {h.loc.name for h in self.heads} | {h.conf.name for h in self.heads}
        self.body: List[ConvSpec] = [c for c in self.convs if c.name not in head_names]
        self.priors, self.variances = priorbox(self.priors_spec, *self.input_hw)

    # ------------------------------------------------------------------ constructors
    @staticmethod
    def from_files(prototxt: str, caffemodel: str) -> "SSDGraph":
        return SSDGraph(load_netdef(prototxt), read_caffemodel(caffemodel))

    @staticmethod
    def random(seed: int = 0, conf_gain: float = 1.0, prototxt: Optional[str] = None) -> "SSDGraph":
        net = load_netdef(prototxt if prototxt is not None else mobilenet_ssd_prototxt())
        return SSDGraph(net, random_init_weights(net, seed=seed, conf_gain=conf_gain))

    # ------------------------------------------------------------------ stats
    @property
    def macs(self) -> int:
        return sum(c.macs for c in self.convs)

    @property
    def n_params(self) -> int:
        return sum(c.n_params for c in self.convs)


class TorchSSD(torch.nn.Module):
    """Plain PyTorch restatement (NCHW in, [B,P,4] loc + [B,P,C] conf logits out)."""

    def __init__(self, graph: SSDGraph, dtype=torch.float32, channels_last: bool = False):
        super().__init__()
        self.graph = graph
        self.channels_last = channels_last
        self.params_w = torch.nn.ParameterDict()
        self.params_b = torch.nn.ParameterDict()
        for c in graph.convs:
            w, b = graph.weights[c.name]
            wt = torch.from_numpy(np.ascontiguousarray(w)).to(dtype)
            if channels_last:
                wt = wt.contiguous(memory_format=torch.channels_last)
            self.params_w[c.name.replace("/", "__")] = torch.nn.Parameter(wt, requires_grad=False)
            self.params_b[c.name.replace("/", "__")] = torch.nn.Parameter(
                torch.from_numpy(np.ascontiguousarray(b)).to(dtype), requires_grad=False)

    def conv(self, c: ConvSpec, x: torch.Tensor) -> torch.Tensor:
        key = c.name.replace("/", "__")
        y = F.conv2d(x, self.params_w[key], self.params_b[key], stride=c.stride, padding=c.pad, groups=c.group)
        return F.relu(y) if c.relu else y

    def forward(self, x: torch.Tensor, capture: Optional[Dict[str, torch.Tensor]] = None):
        g = self.graph
        if self.channels_last:
            x = x.contiguous(memory_format=torch.channels_last)
        acts: Dict[str, torch.Tensor] = {g.net.input_name: x}
        for c in g.body:
            acts[c.top] = self.conv(c, acts[c.bottom])
            if capture is not None:
                capture[c.name] = acts[c.top]
        B = x.shape[0]
        locs, confs = [], []
        for h in g.heads:
            t = acts[h.tap]
            locs.append(self.conv(h.loc, t).permute(0, 2, 3, 1).reshape(B, -1, 4))
            confs.append(self.conv(h.conf, t).permute(0, 2, 3, 1).reshape(B, -1, g.num_classes))
        return torch.cat(locs, 1), torch.cat(confs, 1)


def packed_weights(graph: SSDGraph):
    """Layout transforms shared by the native engine and its tests.

    Returns dict name -> (w, b) with
      * pointwise 1x1      : w [Cout, Cin]                       (K-major B operand)
      * depthwise 3x3      : w [9, C]   (tap-major, channel innermost – NHWC friendly)
      * dense 3x3 (conv0)  : w [Cout, 27]  in (kh, kw, cin) order
      * dense 3x3 (extras) : w [Cout, 9*Cin] in (kh, kw, cin) order (im2col K order)
    and fused heads name f"{tap}_mbox": w [pad16(A*4 + A*C), Cin] = [loc ; conf], b likewise.
    """
    out = {}
    for c in graph.convs:
        w, b = graph.weights[c.name]
        if c.depthwise:
            out[c.name] = (np.ascontiguousarray(w.reshape(c.cout, 9).T), b)
        elif c.k == 1:
            out[c.name] = (np.ascontiguousarray(w.reshape(c.cout, c.cin)), b)
        else:
            out[c.name] = (np.ascontiguousarray(w.transpose(0, 2, 3, 1).reshape(c.cout, -1)), b)
    for h in graph.heads:
        wl, bl = out[h.loc.name]
        wc, bc = out[h.conf.name]
        n = wl.shape[0] + wc.shape[0]
        npad = (n + 15) // 16 * 16
        w = np.zeros((npad, wl.shape[1]), np.float32)
        b = np.zeros((npad,), np.float32)
        w[: wl.shape[0]] = wl
        w[wl.shape[0]: n] = wc
        b[: wl.shape[0]] = bl
        b[wl.shape[0]: n] = bc
        out[f"{h.tap}_mbox"] = (w, b)
    return out
