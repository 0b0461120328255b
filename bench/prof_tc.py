"""Representative tensor-core kernel launches for ncu: GEMM 8192^3 and 4096x2304x768, halo conv
(500x32x32, 64->128), per-tap conv (500x8x8, 256->512), wgrad (500x16x16, 128->128)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from commefficient_b200.ops import tc
d = "cuda"
A = torch.randn(8192, 8192, device=d).bfloat16(); B = torch.randn(8192, 8192, device=d).bfloat16()
a = torch.randn(4096, 768, device=d).bfloat16(); b = torch.randn(2304, 768, device=d).bfloat16()
x1 = torch.randn(500, 32, 32, 64, device=d).bfloat16(); w1 = torch.randn(128, 3, 3, 64, device=d).bfloat16()
x3 = torch.randn(500, 8, 8, 256, device=d).bfloat16(); w3 = torch.randn(512, 3, 3, 256, device=d).bfloat16()
xr = torch.randn(500, 16, 16, 128, device=d).bfloat16(); dz = torch.randn(500, 16, 16, 128, device=d).bfloat16()
dw = torch.zeros(128, 3, 3, 128, device=d)
def run():
    tc.gemm_tn(A, B); tc.gemm_tn(a, b); tc.conv3x3_nhwc(x1, w1, relu=True); tc.conv3x3_nhwc(x3, w3, relu=True)
    tc.conv3x3_wgrad_nhwc(dz, xr, dw)
for _ in range(3): run()
torch.cuda.synchronize(); torch.cuda.profiler.start(); run(); torch.cuda.synchronize(); torch.cuda.profiler.stop()
print("ok")
