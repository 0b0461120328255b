#!/usr/bin/env python
"""tcgen05 tile kernels vs cuBLAS / cuDNN on the shapes the models use (1 GPU, CUDA events)."""
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from commefficient_b200.ops import tc  # noqa: E402


def timed(fn, iters=20, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3


def main():
    dev = "cuda"
    rows = []
    for M, N, K in [(16384, 2304, 768), (16384, 768, 768), (16384, 3072, 768), (16384, 768, 3072), (8192, 8192, 8192),
                    (4096, 50264, 768), (16384, 768, 50264)]:
        a = torch.randn(M, K, device=dev).bfloat16()
        b = torch.randn(N, K, device=dev).bfloat16()
        t_ours = timed(lambda: tc.gemm_tn(a, b))
        t_lib = timed(lambda: a @ b.t())
        fl = 2.0 * M * N * K
        rows.append({"op": "gemm_tn", "shape": [M, N, K], "ours_us": t_ours, "cublas_us": t_lib,
                     "ours_tflops": fl / t_ours / 1e6, "cublas_tflops": fl / t_lib / 1e6})
    n = 500
    for h, cin, cout in [(32, 64, 128), (16, 128, 128), (16, 128, 256), (8, 256, 512), (4, 512, 512)]:
        x = torch.randn(n, h, h, cin, device=dev).bfloat16()
        w = torch.randn(cout, 3, 3, cin, device=dev).bfloat16()
        xc = x.permute(0, 3, 1, 2)          # channels_last view
        wc = w.permute(0, 3, 1, 2)
        torch.backends.cudnn.benchmark = True
        t_ours = timed(lambda: tc.conv3x3_nhwc(x, w, relu=True))
        t_lib = timed(lambda: F.relu(F.conv2d(xc, wc, padding=1)))
        fl = 2.0 * n * h * h * cout * 9 * cin
        rows.append({"op": "conv3x3+relu", "shape": [n, h, cin, cout], "ours_us": t_ours, "cudnn_us": t_lib,
                     "ours_tflops": fl / t_ours / 1e6, "cudnn_tflops": fl / t_lib / 1e6})
    for h, cin, cout in [(32, 64, 128), (16, 128, 128), (16, 128, 256), (8, 256, 512), (4, 512, 512)]:
        x = torch.randn(n, h, h, cin, device=dev).bfloat16()
        dz = torch.randn(n, h, h, cout, device=dev).bfloat16()
        dw = torch.zeros(cout, 3, 3, cin, device=dev)
        xc, dzc = x.permute(0, 3, 1, 2), dz.permute(0, 3, 1, 2)
        t_ours = timed(lambda: tc.conv3x3_wgrad_nhwc(dz, x, dw))
        t_lib = timed(lambda: torch.nn.grad.conv2d_weight(xc, (cout, cin, 3, 3), dzc, padding=1))
        fl = 2.0 * n * h * h * cout * 9 * cin
        rows.append({"op": "conv3x3 wgrad", "shape": [n, h, cin, cout], "ours_us": t_ours, "cudnn_us": t_lib,
                     "ours_tflops": fl / t_ours / 1e6, "cudnn_tflops": fl / t_lib / 1e6})
    for K, M, N in [(16384, 768, 2304), (16384, 3072, 768), (16384, 768, 3072)]:
        a = torch.randn(K, M, device=dev).bfloat16()
        b = torch.randn(K, N, device=dev).bfloat16()
        c = torch.zeros(M, N, device=dev)
        t_ours = timed(lambda: tc.gemm_mn_acc(a, b, c))
        t_lib = timed(lambda: torch.mm(a.t(), b))
        fl = 2.0 * M * N * K
        rows.append({"op": "gemm_mn_acc (linear wgrad)", "shape": [K, M, N], "ours_us": t_ours, "cublas_us": t_lib,
                     "ours_tflops": fl / t_ours / 1e6, "cublas_tflops": fl / t_lib / 1e6})
    print(json.dumps(rows, indent=1))


if __name__ == "__main__":
    main()
