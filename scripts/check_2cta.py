"""CTA-pair (tcgen05 cta_group::2) GEMM: numerics vs fp32 torch and timing vs the single-CTA path."""
import os, sys
import numpy as np, torch
sys.path.insert(0, '.')
from distributedvolunteercomputing_b200.ops import kernels as K
dev = "cuda:0"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, reps=7):
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return min(ts)
shapes = [(256, 128, 64), (300, 128, 128), (1000, 256, 256), (23104, 512, 512), (6400, 1024, 1024), (6400, 256, 1024), (92416, 256, 256), (23104, 512, 256), (6400, 1024, 512), (23104, 512, 4096), (18944, 512, 512), (18944, 256, 4096)]
if len(sys.argv) > 1: shapes = shapes[int(sys.argv[1]):]
torch.manual_seed(0)
for M, N, Kd in shapes:
    a = torch.randn((M, Kd), device=dev).half(); w = (torch.randn((N, Kd), device=dev) * 0.05).half(); b = torch.randn((N,), device=dev)
    ref = torch.relu(a.float() @ w.float().t() + b)
    line = f"M={M} N={N} K={Kd}:"
    for bn in (128, 256):
        if N % bn: continue
        for mode in ("0", "1"):
            os.environ["DVC_GEMM_2CTA"] = mode
            out = K.gemm_f16(a, w, b, True, block_n=bn)
            torch.cuda.synchronize()
            err = (out.float() - ref).abs().max().item()
            t = timeit(lambda: K.gemm_f16(a, w, b, True, block_n=bn, out=out))
            line += f"  bn={bn} {'pair' if mode == '1' else 'solo'} err={err:.3g} {t:.1f}us"
    print(line, flush=True)
