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
torch.cuda.synchronize()
print("ok")
