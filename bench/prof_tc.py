"""Two representative tile-kernel launches for ncu (GEMM 4096x2304x768, conv 500x32x32 64->128)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from commefficient_b200.ops import tc
a = torch.randn(4096, 768, device="cuda").bfloat16(); b = torch.randn(2304, 768, device="cuda").bfloat16()
x = torch.randn(500, 32, 32, 64, device="cuda").bfloat16(); w = torch.randn(128, 3, 3, 64, device="cuda").bfloat16()
for _ in range(3):
    tc.gemm_tn(a, b); tc.conv3x3_nhwc(x, w, relu=True)
torch.cuda.synchronize()
torch.cuda.profiler.start()
tc.gemm_tn(a, b); tc.conv3x3_nhwc(x, w, relu=True)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
