#!/usr/bin/env python
"""ncu target: the bandwidth-bound transformer kernels at the GPT-2 bench shape (16 384 rows x 768 / 3072)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from commefficient_b200.ops import tc, tfm  # noqa: E402

rows, E = 16384, 768
resid = torch.randn(rows, E, device="cuda", requires_grad=True)
branch = torch.randn(rows, E, device="cuda").bfloat16().requires_grad_(True)
w = torch.ones(E, device="cuda", requires_grad=True)
b = torch.zeros(E, device="cuda", requires_grad=True)
state = tfm.DropoutState("cuda", 1)
x = torch.randn(rows, 4 * E, device="cuda").bfloat16().requires_grad_(True)
dy = torch.randn(rows, E, device="cuda").bfloat16()
dr = torch.randn(rows, E, device="cuda")
for _ in range(3):
    nr, y = tfm.add_layer_norm(resid, branch, w, b, 1e-5, 0.1, 3, state)
    torch.autograd.backward([nr, y], [dr, dy])
    g = tfm.gelu_tanh(x)
    g.backward(torch.ones_like(g))
    tc.colsum_acc(x.detach(), torch.zeros(4 * E, device="cuda"))
# round-2 additions: GELU inside the MLP GEMM epilogues, segment BatchNorm on fp32 activations
a = (torch.randn(8192, E, device="cuda") * 0.5).bfloat16()
wfc = (torch.randn(4 * E, E, device="cuda") * 0.05).bfloat16()
bias = torch.zeros(4 * E, device="cuda")
gdy = (torch.randn(8192, E, device="cuda") * 0.5).bfloat16()
wpr = (torch.randn(4 * E, E, device="cuda") * 0.05).bfloat16()
xbn = torch.randn(128, 64, 32, 32, device="cuda").contiguous(memory_format=torch.channels_last).requires_grad_(True)
seg = torch.repeat_interleave(torch.arange(16, device="cuda"), 8)
cnt = torch.full((16,), 8.0, device="cuda")
gam = torch.ones(64, device="cuda", requires_grad=True)
bet = torch.zeros(64, device="cuda", requires_grad=True)
for _ in range(2):
    pre, act = tc.gemm_tn_gelu(a, wfc, bias)
    tc.gemm_tn_gelu_grad(gdy, wpr, pre)
    ybn, _m, _v = tfm.segment_batch_norm(xbn, gam, bet, seg, cnt, 16, 1e-5)
    ybn.backward(torch.ones_like(ybn))
torch.cuda.synchronize()
print("ok")
