"""Depthwise 3x3 tile/CTA-shape sweep at batch 64 (env overrides read by dw_prepare): per layer, best config."""
import os, sys
import numpy as np, torch
sys.path.insert(0, '.')
from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
from distributedvolunteercomputing_b200.ops import kernels as K
dev = "cuda:0"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, reps=5):
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return float(min(ts))
g = SSDGraph.random(seed=1)
B = 64
cfgs = [(256, 56, 0), (128, 56, 0), (128, 48, 0), (128, 40, 0), (128, 34, 0), (64, 56, 0), (64, 40, 0), (64, 28, 0), (64, 20, 0),
        (96, 56, 0), (96, 36, 0), (128, 56, 128), (64, 28, 32), (64, 56, 128)]
seen = set()
tot_base, tot_best = 0.0, 0.0
for c in g.body[1:]:
    if not c.depthwise: continue
    H, W = c.in_hw; Ho, Wo = c.out_hw
    key = (H, W, c.cin, c.stride)
    x = torch.randn((B, H, W, c.cin), device=dev).half()
    w = torch.randn((9, c.cin), device=dev); b = torch.randn((c.cin,), device=dev)
    byts = x.numel() * 2 + B * Ho * Wo * c.cout * 2
    res = []
    for thr, kb, ct in cfgs:
        if ct and c.cin % ct: continue
        os.environ["DVC_DW_THREADS"] = str(thr); os.environ["DVC_DW_SMEM_KB"] = str(kb); os.environ["DVC_DW_CT"] = str(ct)
        try:
            t = timeit(lambda: K.dwconv3x3(x, w, b, c.stride, True))
        except Exception as e:
            t = float("inf")
        res.append((t, thr, kb, ct))
    base = res[0][0]
    best = min(res)
    tot_base += base; tot_best += best[0]
    print(f"{c.name:10s} s{c.stride} {H}x{W}x{c.cin}: base {base:6.1f} us ({byts/base/1e3:5.0f} GB/s)  best {best[0]:6.1f} us thr={best[1]} smem={best[2]}KB ct={best[3]}  all: "
          + " ".join(f"{t:.1f}" for t, *_ in res), flush=True)
    del x
print(f"total base {tot_base:.1f} us, best-per-layer {tot_best:.1f} us")
