"""The allocation-free detection entry point, called through the C ABI.

Some embedders cannot give the library scratch memory (their own pool owns the whole GPU, or
they cannot tolerate cudaMalloc's device-wide stall, or they are plain C with no allocator to
lend).  `dvc_detection_output_static` keeps its scratch in a device array compiled into
the library.  This test drives it exactly as such a caller would - raw pointers through ctypes -
and checks it against the ordinary entry point, twice in a row with different inputs, so that a
pool still holding the first call's data cannot leak into the second.

Collected last on purpose (file name): if it fails under `-x`, every other test has already run.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _call_pair(lib, L, graph, B, gain, seed):
    dev = "cuda:0"
    rng = np.random.default_rng(seed)
    P_, C_ = graph.num_priors, graph.num_classes
    loc = torch.from_numpy(rng.standard_normal((B, P_ * 4)).astype(np.float32)).to(dev)
    conf = torch.from_numpy((rng.standard_normal((B, P_ * C_)) * gain).astype(np.float32)).to(dev)
    pri = torch.from_numpy(graph.priors).to(dev)
    var = torch.from_numpy(graph.variances).to(dev)
    d = graph.det
    prm = L.DetParams(P_, C_, d.background_label_id, d.confidence_threshold, d.nms_threshold, d.top_k, d.keep_top_k,
                      0.2, 1 << 15, 400, 300)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def outputs():
        return (torch.full((B, 100, 7), -1.0, device=dev), torch.full((B,), -1, dtype=torch.int32, device=dev),
                torch.zeros((B, 100, 4), dtype=torch.int32, device=dev), torch.full((B,), -1, dtype=torch.int32, device=dev))

    need = lib.dvc_detection_scratch_bytes(B, P_, C_)
    scratch = torch.empty((need,), dtype=torch.uint8, device=dev)
    a = outputs()
    rc = lib.dvc_detection_output(loc.data_ptr(), conf.data_ptr(), pri.data_ptr(), var.data_ptr(), B, C.byref(prm), None,
                                  a[0].data_ptr(), a[1].data_ptr(), a[2].data_ptr(), a[3].data_ptr(), None, 0,
                                  scratch.data_ptr(), need, stream)
    assert rc == 0, lib.dvc_last_error()
    b = outputs()
    rc = lib.dvc_detection_output_static(loc.data_ptr(), conf.data_ptr(), pri.data_ptr(), var.data_ptr(), B, C.byref(prm), None,
                                         b[0].data_ptr(), b[1].data_ptr(), b[2].data_ptr(), b[3].data_ptr(), None, 0, stream)
    assert rc == 0, lib.dvc_last_error()
    torch.cuda.synchronize()
    return [t.cpu().numpy() for t in a], [t.cpu().numpy() for t in b]


def test_static_scratch_entry_point_matches_and_rejects_oversize(graph):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from distributedvolunteercomputing_b200.ops import lib as L
    lib = L.load()
    # two calls in a row: a crowded batch first (pool left full of candidates), then a sparse one
    for B, gain, seed in ((6, 6.0, 0), (3, 1.5, 1)):
        ref, got = _call_pair(lib, L, graph, B, gain, seed)
        assert ref[1].sum() > 0, "inputs should produce detections"
        assert np.array_equal(ref[1], got[1]) and np.array_equal(ref[3], got[3])          # counts
        assert np.abs(ref[0] - got[0]).max() <= 1e-5                                       # detection rows
        assert np.array_equal(ref[2], got[2])                                              # integer boxes
    # beyond the built-in pool: a clear refusal, nothing launched
    P_, C_ = graph.num_priors, graph.num_classes
    B_big = 64
    assert lib.dvc_detection_scratch_bytes(B_big, P_, C_) > 4 * 1024 * 1024
    d = graph.det
    prm = L.DetParams(P_, C_, d.background_label_id, d.confidence_threshold, d.nms_threshold, d.top_k, d.keep_top_k,
                      0.2, 1 << 15, 400, 300)
    x = torch.zeros((16,), device="cuda:0")
    rc = lib.dvc_detection_output_static(x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), B_big, C.byref(prm), None,
                                         x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), None, 0,
                                         C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = lib.dvc_last_error().decode()
    assert "built-in" in msg and str(4 * 1024 * 1024) in msg, msg
    torch.cuda.synchronize()
