"""profiles/r2/launch_list_step.md (ncu durations of one 64-frame step) + the layer shapes of the model ->
profiles/r2/roofline_step.md: per launch the bytes it has to move at least (inputs + weights + outputs, each once), its
FLOPs, and what that is against the MEASURED peaks of this pod (MEASURED_PEAKS.json: copy bandwidth, cuBLAS bf16).
No GPU needed:  python scripts/roofline_step.py"""
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph

B = 64
peaks = {"hbm_gbs": 6541.5, "bf16_tflops": 1667.7}
try:
    peaks.update({k: v for k, v in json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).items() if k in peaks})
except OSError:
    pass
g = SSDGraph.random(seed=1)
heads = {h.tap: h for h in g.heads}

# (label, bytes, flops) in launch order: K1a, K1b, the body with each head right after its tap layer, K7-K9, K10
ops = [("K1a resize 300->400 (u8)", B * (300 * 300 + 400 * 400) * 3, 0),
       ("K1b resize 400->300 + conv0", B * (400 * 400 * 3 + 150 * 150 * 32 * 2), 2 * 27 * 32 * 150 * 150 * B)]
for c in g.body[1:]:
    hi, wi = c.in_hw
    ho, wo = c.out_hw
    if c.depthwise:
        ops.append((f"{c.name} dw3x3/s{c.stride} C={c.cin}", B * (hi * wi + ho * wo) * c.cin * 2, 2 * 9 * B * ho * wo * c.cin))
    elif c.k == 1:
        M = B * ho * wo
        ops.append((f"{c.name} 1x1 M={M} K={c.cin} N={c.cout}", (M * c.cin + c.cin * c.cout + M * c.cout) * 2, 2 * M * c.cin * c.cout))
    else:
        M = B * ho * wo
        ops.append((f"{c.name} 3x3/s2 implicit M={M} K={9 * c.cin} N={c.cout}",
                    (B * hi * wi * c.cin + 9 * c.cin * c.cout + M * c.cout) * 2, 2 * M * 9 * c.cin * c.cout))
    h = heads.get(c.top)
    if h is not None:
        M = B * h.grid_hw[0] * h.grid_hw[1]
        n = h.loc.cout + h.conf.cout
        ops.append((f"{c.name} heads M={M} K={c.cout} N={n}", (M * c.cout + c.cout * n) * 2 + M * n * 4, 2 * M * c.cout * n))
P, C = g.num_priors, g.num_classes
ops += [("K7 softmax", 2 * B * P * C * 4, 0), ("K8 per-class NMS", B * P * (C + 4) * 4, 0), ("K9 merge + deliver", B * 100 * 11 * 4, 0),
        ("K10 overlay", 0, 0)]

rows = []
for ln in open(os.path.join(ROOT, "profiles", "r2", "launch_list_step.md")):
    m = re.match(r"\| (\d+) \| (\S+?)(?:<.*>)? \| \(([\d, ]+)\) \| ([\d.]+) \|", ln)
    if m:
        rows.append((int(m.group(1)), m.group(2).replace("dvc::", ""), m.group(3), float(m.group(4))))
work = [r for r in rows if not r[1].startswith("mbx_")]
assert len(work) == len(ops), (len(work), len(ops))
out = ["# One 64-frame step against the measured peaks of this pod", "",
       f"Durations: `launch_list_step.md` (ncu, warm L2, serialised).  Peaks: `MEASURED_PEAKS.json` - copy bandwidth "
       f"{peaks['hbm_gbs']:.0f} GB/s, cuBLAS bf16 {peaks['bf16_tflops']:.0f} TFLOP/s (burst).  Bytes are the minimum a launch must move "
       "(activations in and out once, weights once); tensors that fit the 126 MB L2 may never reach HBM, so a figure above 100 % "
       "of the copy rate means the launch ran out of L2, not that the number is wrong.", "",
       "| # | kernel | op | us | MB | GB/s | % of copy bw | GFLOP | TFLOP/s | % of cuBLAS bf16 |", "|---|---|---|---|---|---|---|---|---|---|"]
tb = tf = tt = 0.0
for (i, name, grid, us), (label, nbytes, flops) in zip(work, ops):
    gbs = nbytes / us / 1e3
    tfl = flops / us / 1e6
    out.append(f"| {i} | {name} | {label} | {us:.1f} | {nbytes / 1e6:.1f} | {gbs:.0f} | {100 * gbs / peaks['hbm_gbs']:.0f} | "
               f"{flops / 1e9:.2f} | {tfl:.0f} | {100 * tfl / peaks['bf16_tflops']:.0f} |")
    tb += nbytes; tf += flops; tt += us
out += ["", f"Whole step: {tt:.0f} us serialised, {tb / 1e9:.2f} GB of compulsory traffic = {tb / tt / 1e3:.0f} GB/s "
            f"({100 * tb / tt / 1e3 / peaks['hbm_gbs']:.0f} % of the copy rate), {tf / 1e9:.0f} GFLOP = {tf / tt / 1e6:.0f} TFLOP/s "
            f"({100 * tf / tt / 1e6 / peaks['bf16_tflops']:.0f} % of cuBLAS bf16).  With four batches in flight the same work takes 661 us "
            f"per step (`bench_native_n1_final.json`): {tb / 661 / 1e3:.0f} GB/s = {100 * tb / 661 / 1e3 / peaks['hbm_gbs']:.0f} % of the copy rate, "
            f"{tf / 661 / 1e6:.0f} TFLOP/s."]
open(os.path.join(ROOT, "profiles", "r2", "roofline_step.md"), "w").write("\n".join(out) + "\n")
print("\n".join(out[-1:]))
print(len(work), "launches")
