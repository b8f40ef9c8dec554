import sys, numpy as np, torch
sys.path.insert(0, '.')
from distributedvolunteercomputing_b200.ops import kernels as K
dev = "cuda:0"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, reps=9):
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return min(ts)
shapes = [("conv1", 64 * 22500, 64, 32), ("conv2", 64 * 5625, 128, 64), ("conv3", 64 * 5625, 128, 128), ("conv4", 64 * 1444, 256, 128),
          ("conv5", 64 * 1444, 256, 256), ("conv6", 64 * 361, 512, 256), ("conv7", 64 * 361, 512, 512), ("conv12", 6400, 1024, 512),
          ("conv13", 6400, 1024, 1024), ("conv14_1", 6400, 256, 1024), ("conv14_2", 1600, 512, 2304), ("conv15_1", 1600, 128, 512),
          ("conv15_2", 576, 256, 1152), ("conv16_1", 576, 128, 256), ("conv16_2", 256, 256, 1152), ("conv17_1", 256, 64, 256), ("conv17_2", 64, 128, 576),
          ("head11", 23104, 80, 512), ("head13", 6400, 160, 1024), ("head14", 1600, 160, 512), ("head15", 576, 160, 256)]
for name, M, N, Kd in shapes:
    a = torch.randn((M, Kd), device=dev).half(); w = torch.randn((N, Kd), device=dev).half() * 0.05; b = torch.randn((N,), device=dev)
    out = torch.empty((M, N), device=dev, dtype=torch.float16)
    res = []
    for bn in (0, 32, 64, 80, 128, 160, 256):
        if bn and N % bn: continue
        t = timeit(lambda: K.gemm_f16(a, w, b, True, block_n=bn, out=out))
        res.append(f"bn={bn or 'auto'}: {t:6.1f}us")
    print(f"{name:9s} M={M:8d} N={N:5d} K={Kd:5d}  " + "  ".join(res), flush=True)
