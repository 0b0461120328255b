#!/usr/bin/env python
"""Per-kernel timing of the tcgen05 attention kernels against torch SDPA (cuDNN / flash) on the GPT-2 shapes.
CUDA events, warm-up, an L2 flush between iterations."""
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from commefficient_b200.ops import tfm  # noqa: E402


def timeit(fn, iters=20):
    flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device="cuda")
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    out = []
    for B, T, H, p in ((128, 128, 12, 0.1), (128, 128, 12, 0.0), (16, 1024, 12, 0.1), (64, 256, 12, 0.1)):
        E = H * 64
        qkv = torch.randn(B * T, 3 * E, device="cuda").bfloat16().requires_grad_(True)
        dout = torch.randn(B * T, E, device="cuda").bfloat16()
        state = tfm.DropoutState("cuda", 1)
        o = tfm.attention(qkv, B, T, H, p, 1, state)
        t_f = timeit(lambda: tfm.attention(qkv, B, T, H, p, 1, state))
        t_b = timeit(lambda: torch.autograd.grad(o, qkv, dout, retain_graph=True))
        q, k, v = (t.view(B, T, H, 64).transpose(1, 2) for t in qkv.detach().view(B * T, 3, E).unbind(1))
        q, k, v = (t.contiguous().requires_grad_(True) for t in (q, k, v))
        o2 = F.scaled_dot_product_attention(q, k, v, is_causal=True, dropout_p=p)
        d2 = dout.view(B, T, H, 64).transpose(1, 2).contiguous()
        l_f = timeit(lambda: F.scaled_dot_product_attention(q, k, v, is_causal=True, dropout_p=p))
        l_b = timeit(lambda: torch.autograd.grad(o2, (q, k, v), d2, retain_graph=True))
        flops_f = 4 * B * H * T * T * 64 / 2            # causal: half the tiles
        out.append({"B": B, "T": T, "H": H, "p_drop": p, "ours_fwd_us": t_f, "ours_bwd_us": t_b, "sdpa_fwd_us": l_f,
                    "sdpa_bwd_us": l_b, "ours_fwd_tflops": flops_f / t_f / 1e6, "ours_bwd_tflops": 2.5 * flops_f / t_b / 1e6})
        print(json.dumps(out[-1]), flush=True)


if __name__ == "__main__":
    main()
