"""Two launches of the implicit 3x3/s2 convolution at the conv14_2 shape (64 x 10 x 10 x 256 -> 5 x 5 x 512): the
`ncu --set full` target for the im2col-mode TMA path of gemm_f16_kernel (scripts/gpu_profile2.sh profiles the second)."""
import sys
import torch
sys.path.insert(0, '.')
from distributedvolunteercomputing_b200.ops import kernels as K
dev = "cuda:0"
x = torch.randn((64, 10, 10, 256), device=dev).half()
wt = (torch.randn((512, 9 * 256), device=dev) * 0.03).half()
b = torch.randn((512,), device=dev)
for _ in range(2):
    y = K.conv3x3s2_f16(x, wt, b, True)
torch.cuda.synchronize()
ref = K.gemm_f16(K.im2col3x3s2(x), wt, b, True).view_as(y)
torch.cuda.synchronize()
assert torch.equal(y, ref)
print("ok")
