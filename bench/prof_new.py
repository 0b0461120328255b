#!/usr/bin/env python
"""ncu target for the kernels added in round 2: tcgen05 attention (forward / backward, GPT-2 bench shape) and the
TF32 (kind::tf32) convolution forward / weight-gradient kernels on the ResNet-9 shapes.  Few launches, no timing."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from commefficient_b200.ops import tc, tfm  # noqa: E402

B, T, H = 128, 128, 12
qkv = torch.randn(B * T, 3 * H * 64, device="cuda").bfloat16().requires_grad_(True)
dout = torch.randn(B * T, H * 64, device="cuda").bfloat16()
state = tfm.DropoutState("cuda", 1)
for _ in range(2):
    o = tfm.attention(qkv, B, T, H, 0.1, 1, state)
    torch.autograd.grad(o, qkv, dout)
tc.set_precision("tf32")
for (n, h, cin, cout) in ((500, 32, 64, 128), (500, 16, 128, 256), (500, 8, 256, 512)):
    x = torch.randn(n, h, h, cin, device="cuda")
    w = torch.randn(cout, 3, 3, cin, device="cuda") / (3 * cin ** 0.5)
    dz = torch.randn(n, h, h, cout, device="cuda")
    dw = torch.zeros(cout, 3, 3, cin, device="cuda")
    for _ in range(2):
        tc.conv3x3_nhwc(x, w, relu=True)
        tc.conv3x3_wgrad_nhwc(dz, x, dw)
torch.cuda.synchronize()
print("ok")
