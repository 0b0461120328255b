import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from commefficient_b200.ops import tc
x = torch.randn(500, 32, 32, 3, device="cuda"); w = torch.randn(64, 3, 3, 3, device="cuda")
dz = torch.randn(500, 32, 32, 64, device="cuda").bfloat16(); dw = torch.zeros(64, 3, 3, 3, device="cuda")
for _ in range(2):
    tc.stem_conv(x, w); tc.stem_wgrad(dz, x, dw)
torch.cuda.synchronize(); torch.cuda.profiler.start()
tc.stem_conv(x, w); tc.stem_wgrad(dz, x, dw)
torch.cuda.synchronize(); torch.cuda.profiler.stop(); print("ok")
