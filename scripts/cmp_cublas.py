"""Pointwise / head GEMM shapes at batch 64: our fused tcgen05 kernel (bias+ReLU in the epilogue) vs cuBLAS fp16
(torch.matmul, no bias / ReLU) and vs cuBLAS + separate bias+ReLU.  CUDA events, L2 flushed between reps."""
import sys
import numpy as np, torch
sys.path.insert(0, '.')
from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
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
    return float(min(ts))
g = SSDGraph.random(seed=1)
B = 64
print("| layer | M | N | K | ours us (bias+ReLU fused) | cuBLAS matmul us | cuBLAS + bias+ReLU us | ours / (cuBLAS+epilogue) |")
print("|---|---|---|---|---|---|---|---|")
seen = set()
to, tc, tce = 0.0, 0.0, 0.0
for c in g.body[1:]:
    if c.depthwise or c.k != 1: continue
    Ho, Wo = c.out_hw
    M = B * Ho * Wo
    a = torch.randn((M, c.cin), device=dev).half(); w = torch.randn((c.cout, c.cin), device=dev).half(); b = torch.randn((c.cout,), device=dev)
    bh = b.half()
    out = torch.empty((M, c.cout), device=dev, dtype=torch.float16)
    out2 = torch.empty_like(out)
    wt = w.t()
    ours = timeit(lambda: K.gemm_f16(a, w, b, True, out=out))
    cub = timeit(lambda: torch.matmul(a, wt, out=out2))
    def cub_epi():
        torch.matmul(a, wt, out=out2); out2.add_(bh).relu_()
    cube = timeit(cub_epi)
    to += ours; tc += cub; tce += cube
    print(f"| {c.name} | {M} | {c.cout} | {c.cin} | {ours:.1f} | {cub:.1f} | {cube:.1f} | {ours / cube:.2f} |", flush=True)
    del a, out, out2
print(f"| total | | | | {to:.1f} | {tc:.1f} | {tce:.1f} | {to / tce:.2f} |")
