import os, sys, numpy as np, torch
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
shapes = [("conv5", 64 * 1444, 256, 256), ("conv6", 64 * 361, 512, 256), ("conv7", 64 * 361, 512, 512), ("conv12", 6400, 1024, 512),
          ("conv13", 6400, 1024, 1024), ("conv14_1", 6400, 256, 1024), ("conv14_2", 1600, 512, 2304), ("big", 8192, 1024, 4096)]
for name, M, N, Kd in shapes:
    a = torch.randn((M, Kd), device=dev).half(); w = torch.randn((N, Kd), device=dev).half() * 0.05; b = torch.randn((N,), device=dev)
    out = torch.empty((M, N), device=dev, dtype=torch.float16)
    ref = torch.relu(a.float() @ w.float().t() + b)
    res = []
    for cl in (1, 2, 4):
        os.environ["DVC_GEMM_CLUSTER"] = str(cl)
        for bn in (128, 256):
            if N % bn: continue
            out.zero_()
            t = timeit(lambda: K.gemm_f16(a, w, b, True, block_n=bn, out=out))
            err = (out.float() - ref).abs().max().item() / max(1.0, ref.abs().max().item())
            res.append(f"c{cl}/bn{bn}: {t:5.1f}us{'' if err < 3e-3 else ' ERR %.3g' % err}")
    print(f"{name:9s} M={M:6d} N={N:5d} K={Kd:5d}  " + "  ".join(res), flush=True)
