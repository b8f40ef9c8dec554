"""Per backbone GEMM shape (batch 64): time of gemm_f16 (bias + ReLU fused) for every legal block_n against a bare
torch.matmul (cuBLAS) of the same shape; L2 flushed before every launch.  python scripts/sweep_gemm.py"""
import sys
import torch
sys.path.insert(0, '.')
from distributedvolunteercomputing_b200.ops import kernels as K
dev = "cuda:0"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, reps=9):
    ts = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]
shapes = [("conv1", 1440000, 64, 32), ("conv2", 360000, 128, 64), ("conv3", 360000, 128, 128), ("conv4", 92416, 256, 128),
          ("conv5", 92416, 256, 256), ("conv6", 23104, 512, 256), ("conv7-11", 23104, 512, 512), ("conv12", 6400, 1024, 512),
          ("conv13", 6400, 1024, 1024), ("conv14_1", 6400, 256, 1024), ("conv11_mbox", 23104, 80, 512), ("conv13_mbox", 6400, 160, 1024)]
print("| layer | M | N | K | cuBLAS matmul us | " + " | ".join(f"ours bn={bn}" for bn in (32, 64, 128, 256)) + " | default |")
print("|---|---|---|---|---|---|---|---|---|---|")
for name, M, N, Kd in shapes:
    a = torch.randn((M, Kd), device=dev).half(); w = (torch.randn((N, Kd), device=dev) * 0.05).half(); bias = torch.randn((N,), device=dev)
    tc = timeit(lambda: torch.matmul(a, w.t()))
    cols = []
    for bn in (32, 64, 128, 256):
        if N % bn or bn > N:
            cols.append("-"); continue
        try:
            cols.append(f"{timeit(lambda: K.gemm_f16(a, w, bias, True, block_n=bn)):.1f}")
        except Exception as e:
            cols.append("x")
    td = timeit(lambda: K.gemm_f16(a, w, bias, True))
    print(f"| {name} | {M} | {N} | {Kd} | {tc:.1f} | " + " | ".join(cols) + f" | {td:.1f} |", flush=True)
