import os

import numpy as np
import pytest

from distributedvolunteercomputing_b200 import caffe_io

REF_PROTO = "/root/reference/MobileNetSSD_deploy.prototxt.txt"


def test_generated_prototxt_topology():
    net = caffe_io.load_netdef(caffe_io.mobilenet_ssd_prototxt())
    types = {}
    for l in net.layers:
        types[l.type] = types.get(l.type, 0) + 1
    assert types == {"Convolution": 47, "ReLU": 35, "Permute": 12, "Flatten": 13, "PriorBox": 6,
                     "Concat": 3, "Reshape": 1, "Softmax": 1, "DetectionOutput": 1}
    convs = caffe_io.conv_table(net)
    assert sum(c.n_params for c in convs) == 5_783_417
    assert abs(sum(c.macs for c in convs) / 1e6 - 1148.5) < 0.5
    assert net.input_shape == (1, 3, 300, 300)


@pytest.mark.skipif(not os.path.exists(REF_PROTO), reason="reference mount absent")
def test_reference_prototxt_equivalent():
    ref = caffe_io.load_netdef(REF_PROTO)
    gen = caffe_io.load_netdef(caffe_io.mobilenet_ssd_prototxt())
    assert [(l.name, l.type, l.bottoms, l.tops) for l in ref.layers] == \
           [(l.name, l.type, l.bottoms, l.tops) for l in gen.layers]
    a, b = caffe_io.conv_table(ref), caffe_io.conv_table(gen)
    assert a == b
    from distributedvolunteercomputing_b200.models.detection import prior_specs, detection_params
    assert prior_specs(ref) == prior_specs(gen)
    assert detection_params(ref) == detection_params(gen)


def test_caffemodel_roundtrip(tmp_path):
    net = caffe_io.load_netdef(caffe_io.mobilenet_ssd_prototxt())
    w = caffe_io.random_init_weights(net, seed=3)
    p = tmp_path / "m.caffemodel"
    n = caffe_io.write_caffemodel(str(p), w)
    assert 23_000_000 < n < 23_300_000
    r = caffe_io.read_caffemodel(str(p))
    assert set(r) == set(w) and caffe_io.count_params(r) == 5_783_417
    for k in w:
        for x, y in zip(w[k], r[k]):
            assert x.shape == y.shape and np.array_equal(x, y)


def test_caffemodel_legacy_and_unpacked_fields():
    # V1 'layers' (field 2: name=4, blobs=6), legacy dims, unpacked float data, unknown fields skipped
    from distributedvolunteercomputing_b200.caffe_io import _ld, _varint
    import struct
    vals = [1.5, -2.0, 3.25, 4.0, 5.0, 6.0]
    blob = b"".join(_varint((f << 3) | 0) + _varint(v) for f, v in ((1, 1), (2, 1), (3, 2), (4, 3)))
    blob += b"".join(_varint((5 << 3) | 5) + struct.pack("<f", v) for v in vals)
    layer = _ld(4, b"fc") + _varint((9 << 3) | 0) + _varint(77) + _ld(6, blob)
    data = _ld(1, b"net") + _ld(2, layer) + _varint((55 << 3) | 1) + b"\0" * 8
    r = caffe_io.read_caffemodel(data)
    assert r["fc"][0].shape == (1, 1, 2, 3) and np.allclose(r["fc"][0].ravel(), vals)


def test_cv2_loads_generated_model(ssd_files):
    cv2 = pytest.importorskip("cv2")
    net = cv2.dnn.readNetFromCaffe(*ssd_files)
    out = net.forward() if False else None
    blob = np.zeros((1, 3, 300, 300), np.float32)
    net.setInput(blob)
    assert net.forward().shape[-1] == 7


def test_bn_scale_folding():
    txt = '''name: "t" input: "data" input_shape { dim: 1 dim: 2 dim: 4 dim: 4 }
    layer { name: "c" type: "Convolution" bottom: "data" top: "c" convolution_param { num_output: 3 kernel_size: 1 } }
    layer { name: "c/bn" type: "BatchNorm" bottom: "c" top: "c" batch_norm_param { eps: 0.001 } }
    layer { name: "c/scale" type: "Scale" bottom: "c" top: "c" scale_param { bias_term: true } }
    layer { name: "c/relu" type: "ReLU" bottom: "c" top: "c" }'''
    net = caffe_io.load_netdef(txt)
    rng = np.random.default_rng(0)
    w = rng.standard_normal((3, 2, 1, 1)).astype(np.float32)
    b = rng.standard_normal(3).astype(np.float32)
    mean, var = rng.standard_normal(3).astype(np.float32), rng.random(3).astype(np.float32) + 0.5
    sf = np.array([2.0], np.float32)
    gamma, beta = rng.standard_normal(3).astype(np.float32), rng.standard_normal(3).astype(np.float32)
    folded = caffe_io.fold_batchnorm(net, {"c": [w, b], "c/bn": [mean * 2, var * 2, sf], "c/scale": [gamma, beta]})
    assert set(folded) == {"c"}
    x = rng.standard_normal((2, 5)).astype(np.float32)
    y = (w.reshape(3, 2) @ x) + b[:, None]
    y = (y - mean[:, None]) / np.sqrt(var[:, None] + 0.001) * gamma[:, None] + beta[:, None]
    y2 = folded["c"][0].reshape(3, 2) @ x + folded["c"][1][:, None]
    assert np.allclose(y, y2, atol=1e-5)
    assert caffe_io.conv_table(net)[0].relu
