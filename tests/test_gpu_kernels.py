"""GPU numerics: every native kernel against a plain PyTorch / numpy fp32 restatement."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from distributedvolunteercomputing_b200.models import detection as D
from distributedvolunteercomputing_b200.models import preprocess as P


@pytest.fixture(scope="module")
def K():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from distributedvolunteercomputing_b200.ops import kernels, lib
    lib.load()   # hard failure if the in-tree library is missing/unloadable
    return kernels


DEV = "cuda:0"


def _rand(shape, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale)


@pytest.mark.parametrize("M,N,K_,bn", [
    (128, 64, 64, 0), (300, 128, 64, 0), (1000, 64, 32, 0), (5625, 128, 128, 128), (1444, 256, 256, 256),
    (361 * 3, 512, 512, 128), (100, 1024, 1024, 256), (100, 256, 1024, 64), (25, 512, 2304, 0), (9, 128, 256, 0),
    (4, 64, 256, 0), (1, 128, 576, 0), (22500, 64, 32, 64), (700, 80, 512, 80), (300, 160, 1024, 160), (513, 48, 96, 48),
    (2166, 512, 512, 16), (5776, 256, 256, 32), (2166, 512, 512, 0), (5776, 256, 256, 0), (4 * 22500, 64, 32, 0), (16 * 5625, 128, 128, 0),
])
def test_gemm_matches_fp32(K, M, N, K_, bn):
    a = _rand((M, K_), 1.0, 1).half().to(DEV)
    w = _rand((N, K_), 1.0 / np.sqrt(K_), 2).half().to(DEV)
    bias = _rand((N,), 0.5, 3).to(DEV)
    out = K.gemm_f16(a, w, bias, relu=True, block_n=bn)
    torch.cuda.synchronize()
    ref = torch.relu(a.float() @ w.float().t() + bias)
    err = (out.float() - ref).abs().max().item()
    assert err <= 2e-3 * max(1.0, ref.abs().max().item()) + 2e-3, (err, ref.abs().max().item())
    out2 = K.gemm_f16(a, w, bias, relu=False, block_n=bn)
    ref2 = a.float() @ w.float().t() + bias
    assert (out2.float() - ref2).abs().max().item() <= 2e-3 * max(1.0, ref2.abs().max().item()) + 2e-3


@pytest.mark.parametrize("M,N,K_,bn", [(1000, 256, 256, 128), (23104 // 4 + 77, 512, 512, 256), (900, 1024, 1024, 256)])
def test_gemm_cta_pair_matches_single(K, M, N, K_, bn, monkeypatch):
    """tcgen05 cta_group::2: a CTA pair multiplies one 256 x block_n tile (odd M-tile counts and an M tail included);
    must be bit-identical to the single-CTA path (same K order) and match fp32."""
    a = _rand((M, K_), 1.0, 11).half().to(DEV)
    w = _rand((N, K_), 1.0 / np.sqrt(K_), 12).half().to(DEV)
    bias = _rand((N,), 0.5, 13).to(DEV)
    monkeypatch.setenv("DVC_GEMM_2CTA", "0")
    solo = K.gemm_f16(a, w, bias, relu=True, block_n=bn)
    monkeypatch.setenv("DVC_GEMM_2CTA", "1")
    pair = K.gemm_f16(a, w, bias, relu=True, block_n=bn)
    torch.cuda.synchronize()
    assert torch.equal(solo, pair)
    ref = torch.relu(a.float() @ w.float().t() + bias)
    assert (pair.float() - ref).abs().max().item() <= 2e-3 * max(1.0, ref.abs().max().item()) + 2e-3


def test_gemm_head_split_store(K):
    B, hw, Cin, A_, ncls, P_ = 3, 19 * 19, 512, 3, 21, 1917
    n_loc, n_conf = A_ * 4, A_ * ncls
    N = (n_loc + n_conf + 15) // 16 * 16
    a = _rand((B * hw, Cin), 1.0, 4).half().to(DEV)
    w = torch.zeros((N, Cin)); w[: n_loc + n_conf] = _rand((n_loc + n_conf, Cin), 0.05, 5)
    bias = torch.zeros((N,)); bias[: n_loc + n_conf] = _rand((n_loc + n_conf,), 0.5, 6)
    w, bias = w.half().to(DEV), bias.to(DEV)
    loc = torch.full((B, P_ * 4), -7.0, device=DEV)
    conf = torch.full((B, P_ * ncls), -7.0, device=DEV)
    off = 100
    K.gemm_head(a, w, bias, loc, conf, hw, n_loc, n_conf, off, ncls)
    torch.cuda.synchronize()
    ref = (a.float() @ w.float().t() + bias).view(B, hw, N)
    rl = ref[:, :, :n_loc].reshape(B, -1)
    rc = ref[:, :, n_loc:n_loc + n_conf].reshape(B, -1)
    assert (loc[:, off * 4: off * 4 + hw * n_loc] - rl).abs().max() < 5e-3
    assert (conf[:, off * ncls: off * ncls + hw * n_conf] - rc).abs().max() < 5e-3
    assert (loc[:, : off * 4] == -7).all() and (loc[:, off * 4 + hw * n_loc:] == -7).all()
    assert (conf[:, : off * ncls] == -7).all() and (conf[:, off * ncls + hw * n_conf:] == -7).all()


@pytest.mark.parametrize("B,H,W,C,stride", [(2, 150, 150, 32, 1), (2, 150, 150, 64, 2), (3, 75, 75, 128, 1), (2, 38, 38, 256, 2),
                                            (2, 19, 19, 512, 1), (3, 19, 19, 512, 2), (2, 10, 10, 1024, 1), (1, 7, 5, 8, 1)])
def test_dwconv_matches_fp32(K, B, H, W, C, stride):
    x = _rand((B, H, W, C), 1.0, 7).half().to(DEV)
    w = _rand((C, 1, 3, 3), 0.3, 8).to(DEV)
    b = _rand((C,), 0.2, 9).to(DEV)
    w9c = w.view(C, 9).t().contiguous()
    out = K.dwconv3x3(x, w9c, b, stride, relu=True)
    torch.cuda.synchronize()
    ref = F.relu(F.conv2d(x.float().permute(0, 3, 1, 2), w, b, stride=stride, padding=1, groups=C)).permute(0, 2, 3, 1)
    assert out.shape == ref.shape
    assert (out.float() - ref).abs().max().item() <= 3e-3 * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("B,H,W,Cin,Cout,stride", [
    (4, 150, 150, 32, 64, 1), (3, 75, 75, 128, 128, 1), (4, 38, 38, 256, 256, 1),          # conv1 / conv3 / conv5 (overhanging tiles)
    (3, 150, 150, 64, 128, 2), (3, 75, 75, 128, 256, 2),                                   # conv2 / conv4 (stride 2)
    (1, 8, 16, 64, 64, 1), (2, 23, 37, 256, 192, 1), (2, 19, 19, 128, 256, 2), (1, 23, 37, 64, 64, 2)])
def test_dwpw_fused_matches_fp32_and_two_kernel_path(K, B, H, W, Cin, Cout, stride):
    """Fused depthwise 3x3 -> pointwise 1x1 (csrc/dwpw_fused.cu) against an fp32 restatement (the intermediate rounded
    to fp16 like both GPU paths do) and bit for bit against dwconv3x3 -> gemm_f16."""
    x = _rand((B, H, W, Cin), 1.0, 40).half().to(DEV)
    wdw = _rand((9, Cin), 0.3, 41).to(DEV)
    bdw = _rand((Cin,), 0.2, 42).to(DEV)
    wpw = (_rand((Cout, Cin), 1.0, 43) / Cin ** 0.5).half().to(DEV)
    bpw = _rand((Cout,), 0.2, 44).to(DEV)
    mid = K.dwconv3x3(x, wdw, bdw, stride, True)
    Ho, Wo = mid.shape[1], mid.shape[2]
    two = K.gemm_f16(mid.view(-1, Cin), wpw, bpw, True).view(B, Ho, Wo, Cout)
    fused = K.dwpw_fused(x, wdw, bdw, wpw, bpw, stride)
    torch.cuda.synchronize()
    w4 = wdw.t().reshape(Cin, 1, 3, 3)
    m32 = torch.relu(F.conv2d(x.float().permute(0, 3, 1, 2), w4, bdw, padding=1, stride=stride, groups=Cin)).half().float()
    ref = torch.relu(m32.permute(0, 2, 3, 1).reshape(-1, Cin) @ wpw.float().t() + bpw).view(B, Ho, Wo, Cout)
    assert fused.shape == ref.shape
    assert (fused.float() - ref).abs().max().item() <= 3e-3 * max(1.0, ref.abs().max().item()) + 2e-3
    assert torch.equal(fused, two)


@pytest.mark.parametrize("B,H,W,C", [(2, 10, 10, 256), (3, 5, 5, 128), (2, 3, 3, 128), (4, 2, 2, 64)])
def test_im2col_matches_unfold(K, B, H, W, C):
    x = _rand((B, H, W, C), 1.0, 10).half().to(DEV)
    col = K.im2col3x3s2(x)
    torch.cuda.synchronize()
    u = F.unfold(x.float().permute(0, 3, 1, 2), kernel_size=3, padding=1, stride=2)      # [B, C*9, L]  (c, kh, kw)
    L_ = u.shape[2]
    u = u.view(B, C, 9, L_).permute(0, 3, 2, 1).reshape(B * L_, 9 * C)                      # (kh,kw,c)
    assert torch.equal(col.float(), u)


@pytest.mark.parametrize("B,H,W,C", [(3, 19, 19, 256), (5, 10, 10, 128), (2, 5, 5, 128), (70, 3, 3, 64), (64, 2, 2, 64)])
def test_im2col_tma_tile_matches_patch_rows(K, B, H, W, C):
    """The im2col-mode tensor map of the implicit 3x3/s2 convolution: every operand tile the GEMM producer would fetch
    (128 consecutive output pixels x 64 channels of one tap) equals the matching block of the explicit patch matrix,
    including tiles that start mid-row, span several images, touch the padding halo and run past the last image (those rows are never stored)."""
    x = _rand((B, H, W, C), 1.0, 20).half().to(DEV)
    col = K.im2col3x3s2(x)
    M = col.shape[0]
    m_tiles = (M + 127) // 128
    kbs = sorted({0, 1, 4 * C // 64, 9 * C // 64 - 1, 5 * C // 64 + (C // 64 - 1)})
    for mt in sorted({0, m_tiles // 2, m_tiles - 1}):
        for kb in kbs:
            tile, status = K.debug_im2col_tile(x, mt * 128, kb)
            assert status == 0, (mt, kb, status)
            rows = min(128, M - mt * 128)
            assert torch.equal(tile[:rows], col[mt * 128:mt * 128 + rows, kb * 64:(kb + 1) * 64]), (mt, kb)


@pytest.mark.parametrize("B,H,W,Cin,Cout", [(64, 10, 10, 256, 512), (64, 5, 5, 128, 256), (64, 3, 3, 128, 256), (64, 2, 2, 64, 128),
                                            (64, 19, 19, 256, 512), (3, 19, 19, 256, 512), (1, 3, 3, 64, 128)])
def test_conv3x3s2_implicit_gemm(K, B, H, W, Cin, Cout):
    """SSD extras conv14_2 .. conv17_2 (prototxt:922-1146) without the patch buffer: bit-identical to im2col + GEMM (same
    operand tiles, same accumulation order) and within fp16 tolerance of an fp32 convolution."""
    x = _rand((B, H, W, Cin), 1.0, 21).half().to(DEV)
    w = _rand((Cout, Cin, 3, 3), 0.05, 22)
    b = _rand((Cout,), 0.2, 23).to(DEV)
    wt = w.permute(0, 2, 3, 1).reshape(Cout, 9 * Cin).contiguous().half().to(DEV)        # (kh, kw, cin)
    y = K.conv3x3s2_f16(x, wt, b, relu=True)
    two = K.gemm_f16(K.im2col3x3s2(x), wt, b, relu=True).view_as(y)
    torch.cuda.synchronize()
    assert torch.equal(y, two)
    ref = F.relu(F.conv2d(x.float().permute(0, 3, 1, 2), wt.float().view(Cout, 3, 3, Cin).permute(0, 3, 1, 2), b, stride=2, padding=1))
    ref = ref.permute(0, 2, 3, 1)
    assert (y.float() - ref).abs().max().item() <= 3e-3 * max(1.0, ref.abs().max().item()) + 2e-3


@pytest.mark.parametrize("H,W", [(300, 300), (480, 640), (720, 1280), (600, 800), (1080, 1920), (333, 517), (240, 320), (200, 301), (150, 202), (97, 399)])
def test_resize_area_matches_cv2(K, H, W):
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(H)
    frames = rng.integers(0, 256, (2, H, W, 3), dtype=np.uint8)
    dh, dw = P.out_size_400(H, W)
    t = K.make_area_tables(H, W, dh, dw, DEV)
    out = K.resize_u8(torch.from_numpy(frames).to(DEV), t).cpu().numpy()
    for i in range(2):
        ref = cv2.resize(frames[i], (dw, dh), interpolation=cv2.INTER_AREA)
        d = np.abs(out[i].astype(int) - ref.astype(int))
        assert d.max() <= 1 and (d > 0).mean() < 0.01, (d.max(), (d > 0).mean())
        assert np.array_equal(out[i], P.resize_area_u8(frames[i], dh, dw))


@pytest.mark.parametrize("h4", [300, 225, 400, 257])
@pytest.mark.parametrize("mean", [P.MEAN_COMPAT, P.MEAN_CANONICAL])
def test_preproc_conv0_matches_restatement(K, h4, mean):
    rng = np.random.default_rng(h4)
    frames = rng.integers(0, 256, (3, h4, 400, 3), dtype=np.uint8)
    w = _rand((32, 3, 3, 3), 0.3, 11)
    b = _rand((32,), 0.2, 12)
    w27 = w.permute(2, 3, 1, 0).reshape(27, 32).contiguous().to(DEV)      # (kh,kw,cin) x cout
    out = K.preproc_conv0(torch.from_numpy(frames).to(DEV), w27, b.to(DEV), mean=mean)
    torch.cuda.synchronize()
    blobs = np.concatenate([P.blob_from_image(P.resize_linear_u8(f, 300, 300), mean) for f in frames])
    ref = F.relu(F.conv2d(torch.from_numpy(blobs), w, b, stride=2, padding=1)).permute(0, 2, 3, 1)
    err = (out.float().cpu() - ref).abs().max().item()
    assert err <= 4e-3 * max(1.0, ref.abs().max().item()), err   # fp16 im2col operands on the tensor-core path


@pytest.mark.parametrize("h4,forced", [(300, True), (1300, False)])
def test_preproc_conv0_cuda_core_fallback(K, h4, forced, monkeypatch):
    """K1b has two conv0 back ends: the tcgen05 one and an fp32 CUDA-core one.  The second is chosen when the source band
    of a very tall 400-wide frame does not leave room for the tensor-core staging (h4 = 1300: a 4:13 portrait frame) and
    by DVC_CONV0_CUDA_CORES=1; both must match the restatement."""
    if forced:
        monkeypatch.setenv("DVC_CONV0_CUDA_CORES", "1")
    rng = np.random.default_rng(h4)
    frames = rng.integers(0, 256, (2, h4, 400, 3), dtype=np.uint8)
    w = _rand((32, 3, 3, 3), 0.3, 11)
    b = _rand((32,), 0.2, 12)
    w27 = w.permute(2, 3, 1, 0).reshape(27, 32).contiguous().to(DEV)
    out = K.preproc_conv0(torch.from_numpy(frames).to(DEV), w27, b.to(DEV), mean=P.MEAN_COMPAT)
    torch.cuda.synchronize()
    blobs = np.concatenate([P.blob_from_image(P.resize_linear_u8(f, 300, 300), P.MEAN_COMPAT) for f in frames])
    ref = F.relu(F.conv2d(torch.from_numpy(blobs), w, b, stride=2, padding=1)).permute(0, 2, 3, 1)
    err = (out.float().cpu() - ref).abs().max().item()
    assert err <= 4e-3 * max(1.0, ref.abs().max().item()), err


def test_preproc_frame_index_gathers_ring_slots(K):
    rng = np.random.default_rng(5)
    ring = torch.from_numpy(rng.integers(0, 256, (6, 300, 400, 3), dtype=np.uint8)).to(DEV)
    w27 = _rand((27, 32), 0.3, 13).to(DEV)
    b = _rand((32,), 0.2, 14).to(DEV)
    idx = torch.tensor([4, 1, 5], dtype=torch.int32, device=DEV)
    a = K.preproc_conv0(ring, w27, b, frame_index=idx, B=3)
    bb = K.preproc_conv0(ring[[4, 1, 5]].contiguous(), w27, b)
    assert torch.equal(a, bb)


def test_priorbox_matches_numpy(K, graph):
    boxes, var = K.priorbox(graph.priors_spec, graph.input_hw, DEV)
    assert np.abs(boxes.cpu().numpy() - graph.priors).max() < 1e-6
    assert np.array_equal(var.cpu().numpy(), graph.variances)


@pytest.mark.parametrize("gain,seed", [(1.5, 0), (3.0, 1), (8.0, 2), (0.2, 3)])
def test_detection_output_matches_restatement(K, graph, gain, seed):
    rng = np.random.default_rng(seed)
    B, P_, C_ = 3, graph.num_priors, graph.num_classes
    loc = rng.standard_normal((B, P_, 4)).astype(np.float32)
    conf = (rng.standard_normal((B, P_, C_)) * gain).astype(np.float32)
    pri = torch.from_numpy(graph.priors).to(DEV)
    var = torch.from_numpy(graph.variances).to(DEV)
    prm = K.det_params(graph.det, P_, 400, 300)
    det, cnt, boxes_i, bcnt = K.detection_output(torch.from_numpy(loc.reshape(B, -1)).to(DEV),
                                                 torch.from_numpy(conf.reshape(B, -1)).to(DEV), pri, var, prm)
    torch.cuda.synchronize()
    det, cnt, boxes_i, bcnt = det.cpu().numpy(), cnt.cpu().numpy(), boxes_i.cpu().numpy(), bcnt.cpu().numpy()
    for b in range(B):
        ref = D.detection_output(loc[b], D.softmax(conf[b]), graph.priors, graph.variances, graph.det)
        assert cnt[b] == len(ref), (cnt[b], len(ref))
        got = det[b, : cnt[b]]
        assert np.all(det[b, cnt[b]:] == 0)
        m, oa, ob = D.match_detections(got, ref, 1e-5, 1e-5)
        assert m == len(ref), (m, len(ref), len(oa), len(ob))
        if len(ref):
            assert np.all(np.diff(got[:, 2]) <= 0)            # sorted by confidence, descending
            # same deterministic order as the restatement wherever scores are not within fp noise of a tie
            gaps = np.abs(np.diff(ref[:, 2])) > 1e-6
            solid = np.concatenate([[True], gaps]) & np.concatenate([gaps, [True]])
            assert np.array_equal(got[solid, 1], ref[solid, 1])
        rb, rn = D.app_filter(got, 400, 300)
        assert bcnt[b] == rn and np.array_equal(boxes_i[b, :rn], rb)


def test_annotate_matches_cv2(K):
    cv2 = pytest.importorskip("cv2")
    from distributedvolunteercomputing_b200.utils import annotate as A
    rng = np.random.default_rng(3)
    B, h, w = 4, 225, 400
    frames = rng.integers(0, 256, (B, h, w, 3), dtype=np.uint8)
    boxes = np.zeros((B, 100, 4), np.int32)
    counts = np.array([0, 1, 7, 100], np.int32)
    for b in range(B):
        for k in range(counts[b]):
            x0, x1 = sorted(rng.integers(-40, 440, 2)); y0, y1 = sorted(rng.integers(-40, 260, 2))
            boxes[b, k] = (x0, y0, x1, y1)
    rid = "127.0.0.1:5554"
    idm, ixy, lm, lxy = A.text_masks(rid, h, w)
    masks = K.TextMasks(torch.from_numpy(idm).to(DEV), ixy, torch.from_numpy(lm).to(DEV), lxy)
    out = K.annotate(torch.from_numpy(frames.copy()).to(DEV), torch.from_numpy(boxes).to(DEV),
                     torch.from_numpy(counts).to(DEV), masks).cpu().numpy()
    for b in range(B):
        ref = A.annotate_cpu(frames[b].copy(), boxes[b, : counts[b]], rid, A.label_text({"person": int(counts[b])}))
        assert np.array_equal(out[b], ref), (b, int((out[b] != ref).sum()))


def test_p2p_buffers_epoch_broadcast_and_weight_push(K):
    """P2P substrate on one GPU: cudaMalloc'ed exportable buffers, K14 epoch word, K13 multicast push."""
    import ctypes as C
    from distributedvolunteercomputing_b200.ops import lib as L
    from distributedvolunteercomputing_b200.parallel import p2p
    lib = L.load()
    bufs = [p2p.DevBuffer(4096, DEV) for _ in range(3)]
    assert len(bufs[0].handle()) == 64
    views = [b.tensor(torch.int32, (1024,)) for b in bufs]
    assert all(int(v.abs().sum()) == 0 for v in views)                     # zero-initialised
    ptrs = p2p.device_ptr_array([b.ptr + 16 for b in bufs], DEV)
    L.check(lib.dvc_epoch_broadcast(ptrs.data_ptr(), 3, 7, L.stream_ptr()), "epoch")
    src = torch.arange(1024, dtype=torch.int32, device=DEV)
    wptrs = p2p.device_ptr_array([b.ptr + 1024 for b in bufs], DEV)
    L.check(lib.dvc_push_multicast(src.data_ptr(), wptrs.data_ptr(), 3, 2048, L.stream_ptr()), "push")
    sink = torch.zeros(1, dtype=torch.int32, device=DEV)
    L.check(lib.dvc_peer_read(bufs[0].ptr, 4096, sink.data_ptr(), L.stream_ptr()), "peer_read")
    torch.cuda.synchronize()
    for v in views:
        assert int(v[4]) == 7                                              # epoch word landed
        assert torch.equal(v[256:768], src[:512])                          # 2048 bytes of weights landed
        assert int(v[:4].abs().sum()) == 0 and int(v[768:].abs().sum()) == 0
    for b in bufs:
        b.free()
