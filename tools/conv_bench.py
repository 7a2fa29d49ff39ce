#!/usr/bin/env python
"""Implicit-GEMM Conv2D (TMA im2col + tcgen05, csrc/kernels/conv_igemm.cu) against cuDNN on ResNet-50's convolution
shapes at batch 256, bf16: forward, input gradient, weight gradient. CUDA-event timing of 20 back-to-back launches after
warm-up; every tensor is far larger than L2 or rotated. Prints a markdown table (profiles/conv_igemm_vs_cudnn.md).

    python tools/conv_bench.py [--quick]"""
import argparse
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tfdl.ops import conv as C  # noqa: E402

SHAPES = [  # N, H, W, Cin, Cout, k, stride   (Keras ResNet-50 stages at 224x224)
    (256, 56, 56, 64, 64, 3, 1), (256, 28, 28, 128, 128, 3, 1), (256, 14, 14, 256, 256, 3, 1), (256, 7, 7, 512, 512, 3, 1),
    (256, 56, 56, 256, 128, 1, 2), (256, 28, 28, 512, 256, 1, 2), (256, 56, 56, 256, 512, 1, 2),
]


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3  # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda")
    torch.backends.cudnn.benchmark = True
    print("| shape (N,H,W,Cin,Cout,k,s) | op | tcgen05 implicit GEMM us | TFLOP/s | cuDNN us | ratio (cuDNN / ours) |")
    print("|---|---|---|---|---|---|")
    for (N, H, W, Ci, Co, k, s) in (SHAPES[:2] if a.quick else SHAPES):
        pad = (k - 1) // 2
        pads = (pad, pad, pad, pad)
        x = torch.randn((N, H, W, Ci), device=dev).bfloat16()
        w = (torch.randn((Co, k, k, Ci), device=dev) / (k * k * Ci) ** 0.5).bfloat16()
        P, Q = (H + 2 * pad - k) // s + 1, (W + 2 * pad - k) // s + 1
        dy = torch.randn((N, P, Q, Co), device=dev).bfloat16()
        flops = 2.0 * N * P * Q * Co * Ci * k * k
        xc = x.permute(0, 3, 1, 2)  # NHWC memory viewed as NCHW channels_last
        wc = w.permute(0, 3, 1, 2)
        dyc = dy.permute(0, 3, 1, 2)
        rows = [("forward", lambda: C.conv2d_fwd(x, w, None, (s, s), pads), lambda: F.conv2d(xc, wc, None, s, pad)),
                ("wgrad", lambda: C.conv2d_wgrad(x, dy, tuple(w.shape), (s, s), pads),
                 lambda: torch.ops.aten.convolution_backward(dyc, xc, wc, None, (s, s), (pad, pad), (1, 1), False, (0, 0), 1,
                                                            (False, True, False)))]
        if s == 1:
            rows.insert(1, ("dgrad", lambda: C.conv2d_dgrad(dy, w, (N, H, W, Ci), (s, s), pads),
                            lambda: torch.ops.aten.convolution_backward(dyc, xc, wc, None, (s, s), (pad, pad), (1, 1), False,
                                                                       (0, 0), 1, (True, False, False))))
        for name, ours, lib in rows:
            t1, t2 = timed(ours), timed(lib)
            print(f"| {N},{H},{W},{Ci},{Co},{k},{s} | {name} | {t1:.1f} | {flops / t1 / 1e6:.0f} | {t2:.1f} | {t2 / t1:.2f} |", flush=True)


if __name__ == "__main__":
    main()
