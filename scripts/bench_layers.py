"""Per-op timing at the flagship shapes (batch 64, sub-batch 16): CUDA events, L2 flushed between reps.
Prints microseconds and achieved bandwidth / FLOP rate vs the measured peaks."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, '.')
from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
from distributedvolunteercomputing_b200.ops import kernels as K

dev = "cuda:0"
peaks = {"hbm_gbs": 6542.1, "bf16_tflops": 1668.0}
if os.path.exists("MEASURED_PEAKS.json"):
    peaks.update(json.load(open("MEASURED_PEAKS.json")))
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def timeit(fn, reps=7):
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return float(np.median(ts)), float(min(ts))

g = SSDGraph.random(seed=1)
B, SUB = 64, 16
rows = []
only = sys.argv[1] if len(sys.argv) > 1 else ""
for c in g.body[1:]:
    nb = SUB if g.body.index(c) <= [x.name for x in g.body].index("conv6") else B
    H, W = c.in_hw; Ho, Wo = c.out_hw
    if c.depthwise:
        if only and only != "dw": continue
        x = torch.randn((nb, H, W, c.cin), device=dev).half()
        w = torch.randn((9, c.cin), device=dev); b = torch.randn((c.cin,), device=dev)
        med, mn = timeit(lambda: K.dwconv3x3(x, w, b, c.stride, True))
        byts = x.numel() * 2 + nb * Ho * Wo * c.cout * 2
        rows.append((c.name, f"dw s{c.stride} {nb}x{H}x{W}x{c.cin}", med, mn, byts / mn / 1e3, None))
    elif c.k == 1:
        if only and only != "gemm": continue
        M = nb * Ho * Wo
        a = torch.randn((M, c.cin), device=dev).half(); w = torch.randn((c.cout, c.cin), device=dev).half(); b = torch.randn((c.cout,), device=dev)
        out = torch.empty((M, c.cout), device=dev, dtype=torch.float16)
        med, mn = timeit(lambda: K.gemm_f16(a, w, b, True, out=out))
        byts = a.numel() * 2 + w.numel() * 2 + out.numel() * 2
        rows.append((c.name, f"gemm M={M} N={c.cout} K={c.cin}", med, mn, byts / mn / 1e3, 2.0 * M * c.cout * c.cin / mn / 1e6))
for name, desc, med, mn, gbs, tf in rows:
    frac_bw = gbs / peaks["hbm_gbs"]
    s = f"{name:12s} {desc:34s} med {med:7.1f} us  min {mn:7.1f} us  {gbs:7.0f} GB/s ({frac_bw:4.0%} of measured HBM)"
    if tf is not None:
        s += f"  {tf:6.1f} TFLOP/s ({tf / peaks['bf16_tflops']:4.0%} of measured cuBLAS)"
    print(s)
print("total us (min):", round(sum(r[3] * (4 if 'x150x' in r[1] or 'x75x' in r[1] or 'x38x' in r[1] or r[0] in ('conv1','conv2','conv3','conv4','conv5','conv6','conv6/dw') else 1) for r in rows), 1))
