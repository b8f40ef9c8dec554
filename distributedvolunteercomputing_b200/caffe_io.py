"""Caffe checkpoint I/O: prototxt text parser, caffemodel binary reader/writer.

The public surface that must be kept is "load MobileNetSSD_deploy.prototxt.txt +
MobileNetSSD_deploy.caffemodel" (reference: worker.py:186,194 reads both through
cv2.dnn.readNetFromCaffe).  Nothing here depends on caffe or on protobuf codegen:
the text format is parsed by a small recursive-descent parser and the binary
format is decoded with a hand-written protobuf wire reader that only understands
the handful of fields we need:

    NetParameter   : name=1, layers(V1)=2, layer=100
    LayerParameter : name=1, type=2, bottom=3, top=4, blobs=7
    V1Layer        : name=4, blobs=6
    BlobProto      : num=1 channels=2 height=3 width=4 data=5(packed f32)
                     shape=7{dim=1 packed i64} double_data=8
                     raw_data_type=10 raw_data=12 (OpenCV fp16 extension)

Also provided:
  * ``write_caffemodel``           – emit a file cv2.dnn accepts,
  * ``random_init_weights``        – MSRA random init of the SSD (the real blob is
                                     absent from the reference mount,
                                     .MISSING_LARGE_BLOBS:1),
  * ``fold_batchnorm``             – BatchNorm(+Scale) folding for train-style
                                     prototxts (no-op for the deploy file),
  * ``mobilenet_ssd_prototxt``     – generator for an equivalent deploy prototxt so
                                     the repo does not have to vendor the file.
"""
from __future__ import annotations

import math
import struct
from dataclasses import dataclass
from typing import Any, Dict, Iterable, List, Optional, Tuple

import numpy as np

# --------------------------------------------------------------------------------------
# prototxt (protobuf text format)
# --------------------------------------------------------------------------------------


class Message(dict):
    """A parsed text-format message: key -> list of values (repeated-safe)."""

    def one(self, key: str, default: Any = None) -> Any:
        v = self.get(key)
        if not v:
            return default
        return v[-1]

    def all(self, key: str) -> List[Any]:
        return list(self.get(key, []))


def _tokenize(text: str) -> List[str]:
    toks: List[str] = []
    i, n = 0, len(text)
    while i < n:
        c = text[i]
        if c in " \t\r\n,;":
            i += 1
        elif c == "#":
            while i < n and text[i] != "\n":
                i += 1
        elif c in "{}:<>":
            toks.append(c)
            i += 1
        elif c in "\"'":
            q = c
            j = i + 1
            buf = []
            while j < n and text[j] != q:
                if text[j] == "\\" and j + 1 < n:
                    buf.append(text[j + 1])
                    j += 2
                else:
                    buf.append(text[j])
                    j += 1
            toks.append('"' + "".join(buf))
            i = j + 1
        else:
            j = i
            while j < n and text[j] not in " \t\r\n{}:<>,;#\"'":
                j += 1
            toks.append(text[i:j])
            i = j
    return toks


def _scalar(tok: str) -> Any:
    if tok.startswith('"'):
        return tok[1:]
    low = tok.lower()
    if low == "true":
        return True
    if low == "false":
        return False
    try:
        return int(tok)
    except ValueError:
        pass
    try:
        return float(tok)
    except ValueError:
        return tok  # bare enum (CAFFE, TEST, CENTER_SIZE, ...)


def parse_prototxt(text: str) -> Message:
    toks = _tokenize(text)
    pos = 0

    def parse_msg(closer: Optional[str]) -> Message:
        nonlocal pos
        msg = Message()
        while pos < len(toks):
            t = toks[pos]
            if closer is not None and t == closer:
                pos += 1
                return msg
            key = t
            pos += 1
            if pos >= len(toks):
                raise ValueError(f"prototxt: dangling key {key!r}")
            if toks[pos] == ":":
                pos += 1
                if toks[pos] in "{<":
                    close = "}" if toks[pos] == "{" else ">"
                    pos += 1
                    msg.setdefault(key, []).append(parse_msg(close))
                else:
                    msg.setdefault(key, []).append(_scalar(toks[pos]))
                    pos += 1
            elif toks[pos] in "{<":
                close = "}" if toks[pos] == "{" else ">"
                pos += 1
                msg.setdefault(key, []).append(parse_msg(close))
            else:
                raise ValueError(f"prototxt: expected ':' or '{{' after {key!r}")
        if closer is not None:
            raise ValueError("prototxt: unbalanced braces")
        return msg

    return parse_msg(None)


@dataclass
class LayerDef:
    name: str
    type: str
    bottoms: List[str]
    tops: List[str]
    params: Message

    def p(self, block: str) -> Message:
        return self.params.one(block, Message())


@dataclass
class NetDef:
    name: str
    input_name: str
    input_shape: Tuple[int, ...]
    layers: List[LayerDef]

    def by_name(self, name: str) -> LayerDef:
        for l in self.layers:
            if l.name == name:
                return l
        raise KeyError(name)

    def of_type(self, t: str) -> List[LayerDef]:
        return [l for l in self.layers if l.type == t]


def load_netdef(text_or_path: str) -> NetDef:
    text = text_or_path
    if "\n" not in text_or_path and "{" not in text_or_path:
        with open(text_or_path, "r") as f:
            text = f.read()
    root = parse_prototxt(text)
    layers = []
    for lm in root.all("layer") + root.all("layers"):
        layers.append(
            LayerDef(
                name=str(lm.one("name", "")),
                type=str(lm.one("type", "")),
                bottoms=[str(b) for b in lm.all("bottom")],
                tops=[str(t) for t in lm.all("top")],
                params=lm,
            )
        )
    shape: Tuple[int, ...] = ()
    if root.one("input_shape") is not None:
        shape = tuple(int(d) for d in root.one("input_shape").all("dim"))
    elif root.all("input_dim"):
        shape = tuple(int(d) for d in root.all("input_dim"))
    input_name = str(root.one("input", "data"))
    # "Input" layer style
    for l in layers:
        if l.type == "Input" and not shape:
            ip = l.p("input_param")
            if ip.one("shape") is not None:
                shape = tuple(int(d) for d in ip.one("shape").all("dim"))
                input_name = l.tops[0]
    return NetDef(str(root.one("name", "")), input_name, shape, layers)


# --------------------------------------------------------------------------------------
# protobuf wire format (binary caffemodel)
# --------------------------------------------------------------------------------------


def _read_varint(buf: memoryview, pos: int) -> Tuple[int, int]:
    result = 0
    shift = 0
    while True:
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not (b & 0x80):
            return result, pos
        shift += 7
        if shift > 70:
            raise ValueError("caffemodel: varint too long")


def _iter_fields(buf: memoryview) -> Iterable[Tuple[int, int, Any]]:
    """Yield (field_number, wire_type, value) – value is int or memoryview."""
    pos, n = 0, len(buf)
    while pos < n:
        key, pos = _read_varint(buf, pos)
        fno, wt = key >> 3, key & 7
        if wt == 0:
            v, pos = _read_varint(buf, pos)
            yield fno, wt, v
        elif wt == 1:
            yield fno, wt, buf[pos : pos + 8]
            pos += 8
        elif wt == 2:
            ln, pos = _read_varint(buf, pos)
            if pos + ln > n:
                raise ValueError("caffemodel: truncated length-delimited field")
            yield fno, wt, buf[pos : pos + ln]
            pos += ln
        elif wt == 5:
            yield fno, wt, buf[pos : pos + 4]
            pos += 4
        else:
            raise ValueError(f"caffemodel: unsupported wire type {wt}")


def _packed_varints(buf: memoryview) -> List[int]:
    out, pos = [], 0
    while pos < len(buf):
        v, pos = _read_varint(buf, pos)
        out.append(v)
    return out


def _decode_blob(buf: memoryview) -> np.ndarray:
    legacy = {}
    shape: Optional[List[int]] = None
    chunks: List[np.ndarray] = []
    singles: List[float] = []
    dchunks: List[np.ndarray] = []
    raw_type = None
    raw = None
    for fno, wt, v in _iter_fields(buf):
        if fno in (1, 2, 3, 4) and wt == 0:
            legacy[fno] = v
        elif fno == 5:
            if wt == 2:
                chunks.append(np.frombuffer(v, dtype="<f4"))
            elif wt == 5:
                singles.append(struct.unpack("<f", bytes(v))[0])
        elif fno == 7 and wt == 2:
            for f2, w2, v2 in _iter_fields(v):
                if f2 == 1:
                    if w2 == 2:
                        shape = _packed_varints(v2)
                    else:
                        shape = (shape or []) + [v2]
        elif fno == 8:
            if wt == 2:
                dchunks.append(np.frombuffer(v, dtype="<f8"))
            elif wt == 1:
                dchunks.append(np.frombuffer(v, dtype="<f8"))
        elif fno == 10 and wt == 0:
            raw_type = v
        elif fno == 12 and wt == 2:
            raw = v
    if chunks or singles:
        parts = chunks + ([np.asarray(singles, dtype=np.float32)] if singles else [])
        data = np.concatenate(parts).astype(np.float32, copy=False)
    elif dchunks:
        data = np.concatenate(dchunks).astype(np.float32)
    elif raw is not None:
        # OpenCV extension: raw_data_type 2 == FLOAT16 (caffe.proto Type enum in cv2)
        dt = {2: "<f2"}.get(raw_type, "<f4")
        data = np.frombuffer(raw, dtype=dt).astype(np.float32)
    else:
        data = np.zeros((0,), np.float32)
    if shape is None:
        if legacy:
            shape = [legacy.get(k, 1) for k in (1, 2, 3, 4)]
        else:
            shape = [data.size]
    if int(np.prod(shape)) != data.size:
        raise ValueError(f"caffemodel: blob shape {shape} != {data.size} values")
    return np.array(data, dtype=np.float32).reshape(shape)


def read_caffemodel(path_or_bytes) -> Dict[str, List[np.ndarray]]:
    """Return {layer name: [blob, ...]} for every layer that carries blobs."""
    if isinstance(path_or_bytes, (bytes, bytearray, memoryview)):
        data = bytes(path_or_bytes)
    else:
        with open(path_or_bytes, "rb") as f:
            data = f.read()
    out: Dict[str, List[np.ndarray]] = {}
    for fno, wt, v in _iter_fields(memoryview(data)):
        if wt != 2 or fno not in (100, 2):
            continue
        name_field, blob_field = (1, 7) if fno == 100 else (4, 6)
        name = None
        blobs: List[np.ndarray] = []
        for f2, w2, v2 in _iter_fields(v):
            if f2 == name_field and w2 == 2:
                name = bytes(v2).decode("utf-8")
            elif f2 == blob_field and w2 == 2:
                blobs.append(_decode_blob(v2))
        if name is not None and blobs:
            out[name] = blobs
    return out


def _varint(v: int) -> bytes:
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _ld(fno: int, payload: bytes) -> bytes:
    return _varint((fno << 3) | 2) + _varint(len(payload)) + payload


def _encode_blob(arr: np.ndarray) -> bytes:
    arr = np.ascontiguousarray(arr, dtype="<f4")
    shape = b"".join(_varint(int(d)) for d in arr.shape)
    return _ld(7, _ld(1, shape)) + _ld(5, arr.tobytes())


def write_caffemodel(path: str, weights: Dict[str, List[np.ndarray]], net_name: str = "MobileNet-SSD",
                     layer_types: Optional[Dict[str, str]] = None) -> int:
    """Write a binary NetParameter cv2.dnn.readNetFromCaffe accepts. Returns bytes written."""
    layer_types = layer_types or {}
    body = bytearray(_ld(1, net_name.encode()))
    for name, blobs in weights.items():
        lp = _ld(1, name.encode()) + _ld(2, layer_types.get(name, "Convolution").encode())
        for b in blobs:
            lp += _ld(7, _encode_blob(b))
        body += _ld(100, lp)
    with open(path, "wb") as f:
        f.write(body)
    return len(body)


# --------------------------------------------------------------------------------------
# graph helpers: conv table, random init, BN folding
# --------------------------------------------------------------------------------------


@dataclass
class ConvSpec:
    name: str
    bottom: str
    top: str
    cin: int
    cout: int
    k: int
    stride: int
    pad: int
    group: int
    relu: bool = False
    in_hw: Tuple[int, int] = (0, 0)
    out_hw: Tuple[int, int] = (0, 0)

    @property
    def depthwise(self) -> bool:
        return self.group > 1 and self.group == self.cin == self.cout

    @property
    def macs(self) -> int:
        return self.out_hw[0] * self.out_hw[1] * self.cout * (self.cin // self.group) * self.k * self.k

    @property
    def n_params(self) -> int:
        return self.cout * (self.cin // self.group) * self.k * self.k + self.cout


def conv_table(net: NetDef) -> List[ConvSpec]:
    """Shape-propagate the graph and return one ConvSpec per Convolution (ReLU fused in)."""
    shapes: Dict[str, Tuple[int, int, int]] = {}
    if len(net.input_shape) == 4:
        shapes[net.input_name] = (net.input_shape[1], net.input_shape[2], net.input_shape[3])
    specs: List[ConvSpec] = []
    last_by_top: Dict[str, ConvSpec] = {}
    for l in net.layers:
        if l.type == "Convolution":
            cp = l.p("convolution_param")
            k = int(cp.one("kernel_size", 1))
            s = int(cp.one("stride", 1))
            pad = int(cp.one("pad", 0))
            g = int(cp.one("group", 1))
            cout = int(cp.one("num_output"))
            cin, h, w = shapes[l.bottoms[0]]
            oh = (h + 2 * pad - k) // s + 1
            ow = (w + 2 * pad - k) // s + 1
            spec = ConvSpec(l.name, l.bottoms[0], l.tops[0], cin, cout, k, s, pad, g, False, (h, w), (oh, ow))
            shapes[l.tops[0]] = (cout, oh, ow)
            specs.append(spec)
            last_by_top[l.tops[0]] = spec
        elif l.type == "ReLU":
            if l.bottoms[0] in last_by_top and l.tops[0] == l.bottoms[0]:
                last_by_top[l.bottoms[0]].relu = True
            shapes[l.tops[0]] = shapes[l.bottoms[0]]
        elif l.type in ("BatchNorm", "Scale", "Dropout"):
            if l.bottoms[0] in shapes:
                shapes[l.tops[0]] = shapes[l.bottoms[0]]
    return specs


def random_init_weights(net: NetDef, seed: int = 0, conf_gain: float = 1.0) -> Dict[str, List[np.ndarray]]:
    """MSRA-style random init (std = sqrt(2/fan_in)); ``conf_gain`` scales *_mbox_conf
    weights so the random model actually produces detections."""
    rng = np.random.default_rng(seed)
    out: Dict[str, List[np.ndarray]] = {}
    for c in conv_table(net):
        fan_in = (c.cin // c.group) * c.k * c.k
        w = rng.standard_normal((c.cout, c.cin // c.group, c.k, c.k), dtype=np.float32) * math.sqrt(2.0 / fan_in)
        b = (rng.standard_normal((c.cout,), dtype=np.float32) * 0.05).astype(np.float32)
        if c.name.endswith("_mbox_conf"):
            w *= conf_gain
        out[c.name] = [w.astype(np.float32), b]
    return out


def fold_batchnorm(net: NetDef, weights: Dict[str, List[np.ndarray]], eps_default: float = 1e-5
                   ) -> Dict[str, List[np.ndarray]]:
    """Fold BatchNorm (mean,var,scale_factor) and Scale (gamma,beta) layers that follow a
    Convolution into that convolution's weight/bias.  Identity for the deploy prototxt."""
    out = {k: [np.array(b, dtype=np.float32) for b in v] for k, v in weights.items()}
    producer: Dict[str, str] = {}
    for l in net.layers:
        if l.type == "Convolution":
            producer[l.tops[0]] = l.name
            if l.name in out and len(out[l.name]) == 1:
                out[l.name].append(np.zeros(out[l.name][0].shape[0], np.float32))
        elif l.type == "BatchNorm" and l.bottoms[0] in producer and l.name in weights:
            conv = producer[l.bottoms[0]]
            mean, var, sf = weights[l.name][:3]
            sf = float(np.asarray(sf).reshape(-1)[0])
            sf = 1.0 / sf if sf != 0 else 0.0
            eps = float(l.p("batch_norm_param").one("eps", eps_default))
            inv = 1.0 / np.sqrt(var.reshape(-1) * sf + eps)
            w, b = out[conv]
            out[conv] = [w * inv.reshape(-1, 1, 1, 1), (b - mean.reshape(-1) * sf) * inv]
            out.pop(l.name, None)
            producer[l.tops[0]] = conv
        elif l.type == "Scale" and l.bottoms[0] in producer and l.name in weights:
            conv = producer[l.bottoms[0]]
            blobs = weights[l.name]
            gamma = blobs[0].reshape(-1)
            beta = blobs[1].reshape(-1) if len(blobs) > 1 else np.zeros_like(gamma)
            w, b = out[conv]
            out[conv] = [w * gamma.reshape(-1, 1, 1, 1), b * gamma + beta]
            out.pop(l.name, None)
            producer[l.tops[0]] = conv
        elif l.type == "ReLU" and l.bottoms[0] in producer and l.tops[0] != l.bottoms[0]:
            pass
    return out


# --------------------------------------------------------------------------------------
# deploy prototxt generator (equivalent topology; training-only fields omitted)
# --------------------------------------------------------------------------------------

_BACKBONE = [  # (index, cin, cout, dw stride)
    (1, 32, 64, 1), (2, 64, 128, 2), (3, 128, 128, 1), (4, 128, 256, 2), (5, 256, 256, 1),
    (6, 256, 512, 2), (7, 512, 512, 1), (8, 512, 512, 1), (9, 512, 512, 1), (10, 512, 512, 1),
    (11, 512, 512, 1), (12, 512, 1024, 2), (13, 1024, 1024, 1),
]
_EXTRAS = [("conv14", 1024, 256, 512), ("conv15", 512, 128, 256), ("conv16", 256, 128, 256), ("conv17", 256, 64, 128)]
_TAPS = [  # (tap, anchors, min, max, aspect ratios)
    ("conv11", 3, 60.0, None, [2.0]),
    ("conv13", 6, 105.0, 150.0, [2.0, 3.0]),
    ("conv14_2", 6, 150.0, 195.0, [2.0, 3.0]),
    ("conv15_2", 6, 195.0, 240.0, [2.0, 3.0]),
    ("conv16_2", 6, 240.0, 285.0, [2.0, 3.0]),
    ("conv17_2", 6, 285.0, 300.0, [2.0, 3.0]),
]


def mobilenet_ssd_prototxt(num_classes: int = 21, input_size: int = 300) -> str:
    """Emit a deploy prototxt with the same layer names/shapes as MobileNetSSD_deploy
    (reference: MobileNetSSD_deploy.prototxt.txt:1-1912) minus lr/filler noise."""
    L: List[str] = ['name: "MobileNet-SSD"', 'input: "data"',
                    f"input_shape {{ dim: 1 dim: 3 dim: {input_size} dim: {input_size} }}"]

    def conv(name, bottom, cout, k=1, stride=1, pad=0, group=1, relu=True):
        extra = ""
        if pad:
            extra += f" pad: {pad}"
        if stride != 1:
            extra += f" stride: {stride}"
        if group != 1:
            extra += f" group: {group} engine: CAFFE"
        L.append(f'layer {{ name: "{name}" type: "Convolution" bottom: "{bottom}" top: "{name}" '
                 f"convolution_param {{ num_output: {cout} kernel_size: {k}{extra} }} }}")
        if relu:
            L.append(f'layer {{ name: "{name}/relu" type: "ReLU" bottom: "{name}" top: "{name}" }}')

    conv("conv0", "data", 32, k=3, stride=2, pad=1)
    prev = "conv0"
    for i, cin, cout, s in _BACKBONE:
        conv(f"conv{i}/dw", prev, cin, k=3, stride=s, pad=1, group=cin)
        conv(f"conv{i}", f"conv{i}/dw", cout)
        prev = f"conv{i}"
    for base, cin, mid, cout in _EXTRAS:
        conv(f"{base}_1", prev, mid)
        conv(f"{base}_2", f"{base}_1", cout, k=3, stride=2, pad=1)
        prev = f"{base}_2"
    for tap, a, mn, mx, ars in _TAPS:
        for kind, per in (("loc", 4), ("conf", num_classes)):
            n = f"{tap}_mbox_{kind}"
            conv(n, tap, a * per, relu=False)
            L.append(f'layer {{ name: "{n}_perm" type: "Permute" bottom: "{n}" top: "{n}_perm" '
                     f"permute_param {{ order: 0 order: 2 order: 3 order: 1 }} }}")
            L.append(f'layer {{ name: "{n}_flat" type: "Flatten" bottom: "{n}_perm" top: "{n}_flat" '
                     f"flatten_param {{ axis: 1 }} }}")
        mxs = f" max_size: {mx}" if mx is not None else ""
        arss = "".join(f" aspect_ratio: {a_}" for a_ in ars)
        L.append(f'layer {{ name: "{tap}_mbox_priorbox" type: "PriorBox" bottom: "{tap}" bottom: "data" '
                 f'top: "{tap}_mbox_priorbox" prior_box_param {{ min_size: {mn}{mxs}{arss} flip: true clip: false '
                 f"variance: 0.1 variance: 0.1 variance: 0.2 variance: 0.2 offset: 0.5 }} }}")
    for kind, axis, suffix in (("loc", 1, "_flat"), ("conf", 1, "_flat"), ("priorbox", 2, "")):
        bots = " ".join(f'bottom: "{t[0]}_mbox_{kind}{suffix}"' for t in _TAPS)
        L.append(f'layer {{ name: "mbox_{kind}" type: "Concat" {bots} top: "mbox_{kind}" concat_param {{ axis: {axis} }} }}')
    L.append('layer { name: "mbox_conf_reshape" type: "Reshape" bottom: "mbox_conf" top: "mbox_conf_reshape" '
             f"reshape_param {{ shape {{ dim: 0 dim: -1 dim: {num_classes} }} }} }}")
    L.append('layer { name: "mbox_conf_softmax" type: "Softmax" bottom: "mbox_conf_reshape" top: "mbox_conf_softmax" '
             "softmax_param { axis: 2 } }")
    L.append('layer { name: "mbox_conf_flatten" type: "Flatten" bottom: "mbox_conf_softmax" top: "mbox_conf_flatten" '
             "flatten_param { axis: 1 } }")
    L.append('layer { name: "detection_out" type: "DetectionOutput" bottom: "mbox_loc" bottom: "mbox_conf_flatten" '
             'bottom: "mbox_priorbox" top: "detection_out" include { phase: TEST } detection_output_param { '
             f"num_classes: {num_classes} share_location: true background_label_id: 0 "
             "nms_param { nms_threshold: 0.45 top_k: 100 } code_type: CENTER_SIZE keep_top_k: 100 "
             "confidence_threshold: 0.25 } }")
    return "\n".join(L) + "\n"


def count_params(weights: Dict[str, List[np.ndarray]]) -> int:
    return int(sum(int(b.size) for blobs in weights.values() for b in blobs))
