"""Harness for the EXPERIMENTAL fused depthwise->pointwise kernel (csrc/dwpw_fused.cu).  Run under a timeout on a GPU:

    timeout 120 python scripts/check_dwpw_fused.py

Compares against the validated two-kernel path (dwconv3x3 -> gemm_f16) and an fp32 torch restatement, then times both
at the flagship shapes (batch 64, L2 flushed).  Small shapes first so a hang or a wrong descriptor shows up cheaply."""
import sys
import torch
import torch.nn.functional as F
sys.path.insert(0, '.')
from distributedvolunteercomputing_b200.ops import kernels as K
dev = "cuda:0"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, reps=7):
    ts = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return min(ts)


def case(B, H, W, C_in, C_out, time_it=False, stride=1):
    torch.manual_seed(C_in + C_out + H)
    x = torch.randn((B, H, W, C_in), device=dev).half()
    wdw = torch.randn((9, C_in), device=dev) * 0.3
    bdw = torch.randn((C_in,), device=dev) * 0.2
    wpw = (torch.randn((C_out, C_in), device=dev) / C_in ** 0.5).half()
    bpw = torch.randn((C_out,), device=dev) * 0.2
    mid = K.dwconv3x3(x, wdw, bdw, stride, True)
    Ho, Wo = mid.shape[1], mid.shape[2]
    two = K.gemm_f16(mid.view(-1, C_in), wpw, bpw, True).view(B, Ho, Wo, C_out)
    fused = K.dwpw_fused(x, wdw, bdw, wpw, bpw, stride)
    torch.cuda.synchronize()
    # fp32 restatement (the intermediate rounded to fp16 like both GPU paths do)
    w4 = wdw.t().reshape(C_in, 1, 3, 3)
    m32 = torch.relu(F.conv2d(x.float().permute(0, 3, 1, 2), w4, bdw, padding=1, stride=stride, groups=C_in)).half().float()
    ref = torch.relu(m32.permute(0, 2, 3, 1).reshape(-1, C_in) @ wpw.float().t() + bpw).view(B, Ho, Wo, C_out)
    e_two = (two.float() - ref).abs().max().item()
    e_fused = (fused.float() - ref).abs().max().item()
    line = f"B={B} {H}x{W} s{stride} {C_in}->{C_out}: max err two-kernel {e_two:.4f} fused {e_fused:.4f} bit-identical={torch.equal(two, fused)}"
    if time_it:
        t2 = timeit(lambda: K.gemm_f16(K.dwconv3x3(x, wdw, bdw, stride, True).view(-1, C_in), wpw, bpw, True))
        tf = timeit(lambda: K.dwpw_fused(x, wdw, bdw, wpw, bpw, stride))
        line += f"  two kernels {t2:.1f} us, fused {tf:.1f} us"
    print(line, flush=True)
    assert e_fused <= max(2 * e_two, 2e-2), "fused kernel disagrees with the fp32 restatement"


for shape in [(1, 8, 16, 64, 64), (1, 8, 16, 32, 64), (2, 19, 19, 128, 128), (1, 23, 37, 256, 256), (3, 10, 10, 64, 192)]:
    case(*shape)
for shape in [(1, 16, 32, 64, 128), (2, 19, 19, 128, 256), (1, 23, 37, 64, 64), (3, 150, 150, 64, 128)]:
    case(*shape, stride=2)
if "small" not in sys.argv:
    for shape in [(64, 150, 150, 32, 64), (64, 75, 75, 128, 128), (64, 38, 38, 256, 256)]:
        case(*shape, time_it=True)
    for shape in [(64, 150, 150, 64, 128), (64, 75, 75, 128, 256)]:
        case(*shape, time_it=True, stride=2)
print("ok")
