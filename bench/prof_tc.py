"""Representative tensor-core kernel launches for ncu (profile between cudaProfilerStart/Stop):
GEMM 8192^3 and the GPT-2 c_attn shape (CTA-pair tile kernel), halo conv 500x32x32 64->128 (single
CTA) and 500x16x16 128->128 (pair), per-tap conv 500x8x8 256->512 (pair), conv wgrad 500x16x16
128->256 (pair) and 500x32x32 64->128 (single, fused taps), linear wgrad 16384x768x2304 (pair)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from commefficient_b200.ops import tc
d = "cuda"
A = torch.randn(8192, 8192, device=d).bfloat16(); B = torch.randn(8192, 8192, device=d).bfloat16()
a = torch.randn(16384, 768, device=d).bfloat16(); b = torch.randn(2304, 768, device=d).bfloat16()
x1 = torch.randn(500, 32, 32, 64, device=d).bfloat16(); w1 = torch.randn(128, 3, 3, 64, device=d).bfloat16()
x2 = torch.randn(500, 16, 16, 128, device=d).bfloat16(); w2 = torch.randn(128, 3, 3, 128, device=d).bfloat16()
x3 = torch.randn(500, 8, 8, 256, device=d).bfloat16(); w3 = torch.randn(512, 3, 3, 256, device=d).bfloat16()
dz2 = torch.randn(500, 16, 16, 256, device=d).bfloat16(); dw2 = torch.zeros(256, 3, 3, 128, device=d)
dz1 = torch.randn(500, 32, 32, 128, device=d).bfloat16(); dw1 = torch.zeros(128, 3, 3, 64, device=d)
ga = torch.randn(16384, 768, device=d).bfloat16(); gb = torch.randn(16384, 2304, device=d).bfloat16()
gc = torch.zeros(768, 2304, device=d)
def run():
    tc.gemm_tn(A, B); tc.gemm_tn(a, b)
    tc.conv3x3_nhwc(x1, w1, relu=True); tc.conv3x3_nhwc(x2, w2, relu=True); tc.conv3x3_nhwc(x3, w3, relu=True)
    tc.conv3x3_wgrad_nhwc(dz2, x2, dw2); tc.conv3x3_wgrad_nhwc(dz1, x1, dw1); tc.gemm_mn_acc(ga, gb, gc)
for _ in range(3): run()
torch.cuda.synchronize(); torch.cuda.profiler.start(); run(); torch.cuda.synchronize(); torch.cuda.profiler.stop()
print("ok")
