"""One launch of each hot kernel at flagship shapes (after one warm-up launch each) - the ncu --set full target.
Order: resize, preproc_conv0, dw(conv1/dw), gemm(conv1), dw(conv7/dw), gemm(conv7), gemm(conv13), gemm head(conv11_mbox), softmax, class_nms, merge_deliver, annotate (12 launches)."""
import sys
import numpy as np, torch
sys.path.insert(0, '.')
from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
from distributedvolunteercomputing_b200.ops import kernels as K
from distributedvolunteercomputing_b200.utils import annotate as A
dev = "cuda:0"
g = SSDGraph.random(seed=1, conf_gain=2.5)
rng = np.random.default_rng(0)
raw = torch.from_numpy(rng.integers(0, 256, (64, 300, 300, 3), dtype=np.uint8)).to(dev)
t_area = K.make_area_tables(300, 300, 400, 400, dev)
w27 = torch.randn((27, 32), device=dev) * 0.2; b32 = torch.randn((32,), device=dev) * 0.1
def rnd(*s): return torch.randn(s, device=dev).half()
x1 = rnd(64, 150, 150, 32); w1 = torch.randn((9, 32), device=dev); bb1 = torch.randn((32,), device=dev)
a1 = rnd(64 * 22500, 32); wt1 = rnd(64, 32); bias1 = torch.randn((64,), device=dev)
x7 = rnd(64, 19, 19, 512); w7 = torch.randn((9, 512), device=dev); bb7 = torch.randn((512,), device=dev)
a7 = rnd(64 * 361, 512); wt7 = rnd(512, 512) * 0.05; bias7 = torch.randn((512,), device=dev)
a13 = rnd(6400, 1024); wt13 = rnd(1024, 1024) * 0.03; bias13 = torch.randn((1024,), device=dev)
wh = rnd(80, 512) * 0.05; bh = torch.randn((80,), device=dev)
loc = torch.zeros((64, 1917 * 4), device=dev); conf = torch.zeros((64, 1917 * 21), device=dev)
locr = torch.randn((64, 1917 * 4), device=dev); confr = torch.randn((64, 1917 * 21), device=dev) * 2.0
pri = torch.from_numpy(g.priors).to(dev); var = torch.from_numpy(g.variances).to(dev)
prm = K.det_params(g.det, 1917, 400, 400)
idm, ixy, lm, lxy = A.text_masks("127.0.0.1:5554", 400, 400)
masks = K.TextMasks(torch.from_numpy(idm).to(dev), ixy, torch.from_numpy(lm).to(dev), lxy)
def once():
    f400 = K.resize_u8(raw, t_area)
    K.preproc_conv0(f400, w27, b32)
    K.dwconv3x3(x1, w1, bb1, 1, True)
    K.gemm_f16(a1, wt1, bias1, True)
    K.dwconv3x3(x7, w7, bb7, 1, True)
    K.gemm_f16(a7, wt7, bias7, True)
    K.gemm_f16(a13, wt13, bias13, True)
    K.gemm_head(a7, wh, bh, loc, conf, 361, 12, 63, 0, 21)
    det = K.detection_output(locr, confr, pri, var, prm)
    K.annotate(f400, det[2], det[3], masks)
    torch.cuda.synchronize()
once(); once()
print("ok")
