"""ncu target: one launch each of the fused depthwise->pointwise kernel and of the two-kernel path at the flagship
shapes (conv1, conv3, conv5 stride 1; conv2, conv4 stride 2; batch 64).  Run it under ncu:

    python scripts/prof_fused.py && ncu --set full --clock-control none --import-source on \
        -k regex:'dwpw_fused|dw3x3_tma|gemm_f16' -o gpurun_out/prof_fused python scripts/prof_fused.py
"""
import sys
import torch
sys.path.insert(0, '.')
from distributedvolunteercomputing_b200.ops import kernels as K
dev = "cuda:0"
shapes = [(64, 150, 150, 32, 64, 1), (64, 75, 75, 128, 128, 1), (64, 38, 38, 256, 256, 1), (64, 150, 150, 64, 128, 2), (64, 75, 75, 128, 256, 2)]
if len(sys.argv) > 1:
    shapes = [shapes[int(a)] for a in sys.argv[1:]]
for (B, H, W, C_in, C_out, s) in shapes:
    torch.manual_seed(0)
    x = torch.randn((B, H, W, C_in), device=dev).half()
    wdw = torch.randn((9, C_in), device=dev) * 0.3
    bdw = torch.randn((C_in,), device=dev) * 0.2
    wpw = (torch.randn((C_out, C_in), device=dev) / C_in ** 0.5).half()
    bpw = torch.randn((C_out,), device=dev) * 0.2
    torch.cuda.synchronize()
    mid = K.dwconv3x3(x, wdw, bdw, s, True)
    two = K.gemm_f16(mid.view(-1, C_in), wpw, bpw, True)
    fused = K.dwpw_fused(x, wdw, bdw, wpw, bpw, s)
    torch.cuda.synchronize()
print("ok")
