"""tcgen05 TMA GEMM vs cuBLAS (torch.matmul) on layer-sized shapes; roofline fraction of MEASURED_PEAKS.json."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tfdl.ops.gemm import gemm_tma
peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json"))) if os.path.exists("MEASURED_PEAKS.json") else {"bf16_tflops": 1590.0}
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e-3
print("| M | N | K | variant | tfdl TFLOP/s | cuBLAS TFLOP/s | tfdl / cuBLAS | tfdl / measured peak |")
print("|---|---|---|---|---|---|---|---|")
for M, N, K in [(8192, 8192, 8192), (4096, 4096, 4096), (8192, 768, 768), (8192, 3072, 768), (8192, 768, 3072), (32768, 1000, 2048)]:
    a = torch.randn(M, K, device="cuda").bfloat16(); b = torch.randn(N, K, device="cuda").bfloat16()
    at, bt = a.t().contiguous(), b.t().contiguous()
    fl = 2.0 * M * N * K
    for name, fn, ref in [("NT (fwd)", lambda: gemm_tma(a, b), lambda: a @ b.t()),
                          ("NN (dgrad)", lambda: gemm_tma(a, bt, b_mn=True), lambda: a @ bt),
                          ("TN (wgrad)", lambda: gemm_tma(at, bt, a_mn=True, b_mn=True), lambda: at.t() @ bt)]:
        t1, t2 = t(fn), t(ref)
        print(f"| {M} | {N} | {K} | {name} | {fl/t1/1e12:.0f} | {fl/t2/1e12:.0f} | {t2/t1:.2f} | {fl/t1/1e12/peaks['bf16_tflops']:.2f} |")
