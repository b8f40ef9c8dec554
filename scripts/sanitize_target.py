"""Every native kernel once at small shapes - the compute-sanitizer (memcheck / racecheck / synccheck) target."""
import os, sys
import numpy as np, torch
sys.path.insert(0, '.')
from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
from distributedvolunteercomputing_b200.ops import kernels as K, lib as L
from distributedvolunteercomputing_b200.utils import annotate as A
dev = "cuda:0"
rng = np.random.default_rng(0)
g = SSDGraph.random(seed=1, conf_gain=2.5)
def rnd(*s): return torch.randn(s, device=dev).half()
# resize: linear (enlarging), area tables, integer ratio
for (h, w) in [(60, 80), (240, 427), (120, 800)]:
    fr = torch.from_numpy(rng.integers(0, 256, (2, h, w, 3), dtype=np.uint8)).to(dev)
    from distributedvolunteercomputing_b200.models import preprocess as P
    dh, dw = P.out_size_400(h, w)
    K.resize_u8(fr, K.make_area_tables(h, w, dh, dw, dev))
f400 = torch.from_numpy(rng.integers(0, 256, (2, 225, 400, 3), dtype=np.uint8)).to(dev)
w27 = torch.randn((27, 32), device=dev) * 0.2; b32 = torch.randn((32,), device=dev)
K.preproc_conv0(f400, w27, b32)
os.environ["DVC_CONV0_CUDA_CORES"] = "1"; K.preproc_conv0(f400, w27, b32); os.environ.pop("DVC_CONV0_CUDA_CORES")
for (B, H, W, C, s) in [(2, 19, 19, 64, 1), (2, 20, 21, 32, 2), (1, 7, 5, 8, 1)]:
    K.dwconv3x3(rnd(B, H, W, C), torch.randn((9, C), device=dev), torch.randn((C,), device=dev), s, True)
K.im2col3x3s2(rnd(2, 5, 5, 64))
for (M, N, Kd, bn) in [(300, 128, 64, 0), (700, 64, 32, 64), (513, 48, 96, 48), (260, 256, 320, 128), (1000, 512, 256, 256)]:
    K.gemm_f16(rnd(M, Kd), rnd(N, Kd), torch.randn((N,), device=dev), True, block_n=bn)
loc = torch.zeros((2, 1917 * 4), device=dev); conf = torch.zeros((2, 1917 * 21), device=dev)
K.gemm_head(rnd(2 * 361, 512), rnd(80, 512) * 0.05, torch.randn((80,), device=dev), loc, conf, 361, 12, 63, 0, 21)
K.priorbox(g.priors_spec, g.input_hw, dev)
pri = torch.from_numpy(g.priors).to(dev); var = torch.from_numpy(g.variances).to(dev)
prm = K.det_params(g.det, 1917, 400, 225)
locr = torch.randn((2, 1917 * 4), device=dev); confr = torch.randn((2, 1917 * 21), device=dev) * 2.0
det = K.detection_output(locr, confr, pri, var, prm)
import ctypes as C
o = [torch.zeros((2, 100, 7), device=dev), torch.zeros((2,), dtype=torch.int32, device=dev), torch.zeros((2, 100, 4), dtype=torch.int32, device=dev), torch.zeros((2,), dtype=torch.int32, device=dev)]
L.check(L.load().dvc_detection_output_static(locr.data_ptr(), confr.data_ptr(), pri.data_ptr(), var.data_ptr(), 2, C.byref(prm), None,
        o[0].data_ptr(), o[1].data_ptr(), o[2].data_ptr(), o[3].data_ptr(), None, 0, L.stream_ptr()), "static")
idm, ixy, lm, lxy = A.text_masks("127.0.0.1:5554", 225, 400)
K.annotate(f400, det[2], det[3], K.TextMasks(torch.from_numpy(idm).to(dev), ixy, torch.from_numpy(lm).to(dev), lxy))
torch.cuda.synchronize()
print("sanitize target ok")
