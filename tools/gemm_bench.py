#!/usr/bin/env python
"""tcgen05 TMA GEMM (ops/gemm.py:gemm_tma) vs cuBLAS (torch.matmul) on layer-sized shapes.

Each measurement is 20 back-to-back launches captured in one CUDA graph (kernel time, no host in the loop),
timed with CUDA events. Roofline column: fraction of the cuBLAS bf16 throughput in MEASURED_PEAKS.json.
Prints a markdown table (copied to profiles/gemm_tma_vs_cublas.md)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from tfdl.ops.gemm import gemm_tma  # noqa: E402

PEAK = 1624.3
try:
    pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    PEAK = float(pk.get("bf16_tflops", pk.get("cublas_bf16_tflops", PEAK)))
except Exception:
    pass


def graph_time(fn, iters=20, reps=5):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn()
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(iters):
                fn()
        g.replay()
        s.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(reps):
            g.replay()
        e1.record(s)
    s.synchronize()
    return e0.elapsed_time(e1) / (iters * reps) * 1e-3


def main():
    print("| M | N | K | variant | tfdl us | cuBLAS us | tfdl TFLOP/s | cuBLAS TFLOP/s | tfdl / cuBLAS | tfdl / measured peak |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    shapes = [(8192, 8192, 8192), (4096, 4096, 4096), (8192, 768, 768), (8192, 2304, 768), (8192, 3072, 768),
              (8192, 768, 3072), (32768, 1000, 2048)]
    for M, N, K in shapes:
        a = torch.randn(M, K, device="cuda").bfloat16()
        b = torch.randn(N, K, device="cuda").bfloat16()
        at, bt = a.t().contiguous(), b.t().contiguous()
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        fl = 2.0 * M * N * K
        for name, fn, ref in [
            ("NT (fwd)", lambda: gemm_tma(a, b, out=out), lambda: torch.matmul(a, b.t(), out=out)),
            ("NN (dgrad)", lambda: gemm_tma(a, bt, b_mn=True, out=out), lambda: torch.matmul(a, bt, out=out)),
            ("TN (wgrad)", lambda: gemm_tma(at, bt, a_mn=True, b_mn=True, out=out), lambda: torch.matmul(at.t(), bt, out=out)),
        ]:
            t1, t2 = graph_time(fn), graph_time(ref)
            print(f"| {M} | {N} | {K} | {name} | {t1 * 1e6:.1f} | {t2 * 1e6:.1f} | {fl / t1 / 1e12:.0f} | {fl / t2 / 1e12:.0f} "
                  f"| {t2 / t1:.2f} | {fl / t1 / 1e12 / PEAK:.2f} |", flush=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(200):
        gemm_tma(a, b, out=out)
    print(f"\nhost cost of one gemm_tma call (ctypes + 3 tensor-map encodes + launch): {(time.perf_counter() - t0) / 200 * 1e6:.1f} us")


if __name__ == "__main__":
    main()
