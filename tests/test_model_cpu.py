import numpy as np
import pytest
import torch

from distributedvolunteercomputing_b200.models import detection as D
from distributedvolunteercomputing_b200.models import preprocess as P
from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph, TorchSSD, packed_weights

cv2 = pytest.importorskip("cv2")


@pytest.fixture(scope="module")
def cvnet(ssd_files):
    return cv2.dnn.readNetFromCaffe(*ssd_files)


def _frame(seed=0, h=300, w=400):
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, (h // 8, w // 8, 3), dtype=np.uint8)
    return cv2.resize(img, (w, h), interpolation=cv2.INTER_CUBIC)


def test_graph_stats(graph):
    assert graph.num_priors == 1917 and graph.n_params == 5_783_417
    assert [h.anchors for h in graph.heads] == [3, 6, 6, 6, 6, 6]
    assert [h.prior_offset for h in graph.heads] == [0, 1083, 1683, 1833, 1887, 1911]
    pk = packed_weights(graph)
    assert pk["conv11_mbox"][0].shape == (80, 512) and pk["conv13_mbox"][0].shape == (160, 1024)
    assert pk["conv1/dw"][0].shape == (9, 32) and pk["conv0"][0].shape == (32, 27)


def test_priorbox_matches_cv2(graph, cvnet):
    cvnet.setInput(np.zeros((1, 3, 300, 300), np.float32))
    ref = cvnet.forward("mbox_priorbox")          # [1,2,7668]
    assert np.abs(ref[0, 0].reshape(-1, 4) - graph.priors).max() < 1e-6
    assert np.abs(ref[0, 1].reshape(-1, 4) - graph.variances).max() < 1e-6
    px = graph.priors[:3] * 300
    assert np.allclose(px[0], [-22.105, -22.105, 37.895, 37.895], atol=1e-3)
    assert np.allclose(px[1], [-34.532, -13.318, 50.321, 29.108], atol=1e-3)


@pytest.mark.parametrize("mean", [P.MEAN_COMPAT, P.MEAN_CANONICAL])
def test_torch_model_matches_cv2_layers(graph, cvnet, mean):
    frame = _frame(1)
    f400 = cv2.resize(frame, (400, 300), interpolation=cv2.INTER_AREA)
    img300 = cv2.resize(f400, (300, 300))
    blob = P.blob_from_image(img300, mean)
    if mean == P.MEAN_COMPAT:
        assert np.array_equal(blob, cv2.dnn.blobFromImage(img300, 0.007843, (300, 300), 127.5))
    cvnet.setInput(blob)
    names = ["conv0", "conv1/dw", "conv5", "conv11", "conv13", "conv17_2", "mbox_loc", "mbox_conf"]
    refs = dict(zip(names, cvnet.forward(names)))
    model = TorchSSD(graph)
    cap = {}
    with torch.no_grad():
        loc, conf = model(torch.from_numpy(blob), capture=cap)
    for n in names[:6]:
        r, m = np.maximum(refs[n], 0), cap[n].numpy()   # cv2 returns the pre-ReLU blob for the conv layer id
        assert r.shape == m.shape
        assert np.abs(r - m).max() <= 2e-4 * max(1.0, np.abs(r).max()), n
    assert np.abs(refs["mbox_loc"].reshape(1917, 4) - loc[0].numpy()).max() < 1e-3
    assert np.abs(refs["mbox_conf"].reshape(1917, 21) - conf[0].numpy()).max() < 1e-3


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_detection_output_matches_cv2(graph, cvnet, seed):
    frame = _frame(seed, 360, 640)
    f400, blob = P.preprocess_frame(frame)
    cvnet.setInput(blob)
    ref = cvnet.forward("detection_out").reshape(-1, 7)
    loc, conf = [o.reshape(1917, -1) for o in cvnet.forward(["mbox_loc", "mbox_conf"])]
    mine = D.detection_output(loc, D.softmax(conf), graph.priors, graph.variances, graph.det)
    nref = int((ref[:, 2] > 0).sum())
    assert nref > 3, "random model should produce detections (conf_gain)"
    matched, oa, ob = D.match_detections(mine, ref, 1e-5, 1e-5)
    assert matched == nref == len(mine), (matched, nref, len(mine))


def test_full_pipeline_vs_reference_recipe(graph, cvnet):
    """The reference worker's exact per-frame recipe (worker.py:243-269) vs ours end to end."""
    frame = _frame(5, 480, 640)
    f400_ref = cv2.resize(frame, (400, 300), interpolation=cv2.INTER_AREA)
    blob_ref = cv2.dnn.blobFromImage(cv2.resize(f400_ref, (300, 300)), 0.007843, (300, 300), 127.5)
    cvnet.setInput(blob_ref)
    ref = cvnet.forward().reshape(-1, 7)
    f400, blob = P.preprocess_frame(frame)
    assert np.abs(f400.astype(int) - f400_ref.astype(int)).max() <= 1
    model = TorchSSD(graph)
    with torch.no_grad():
        loc, conf = model(torch.from_numpy(blob))
    mine = D.detection_output(loc[0].numpy(), D.softmax(conf[0].numpy()), graph.priors, graph.variances, graph.det)
    matched, oa, ob = D.match_detections(mine, ref, 5e-3, 5e-3)
    nref = int((ref[:, 2] > 0).sum())
    assert matched >= 0.9 * nref and len(oa) + len(ob) <= max(2, 0.1 * nref)
    boxes, count = D.app_filter(mine, 400, 300)
    assert boxes.shape[1] == 4 and count == len(boxes)
