import sys, time, torch
sys.path.insert(0, "/root/repo")
from tfdl.ops.gemm import gemm_tma
def graph_time(fn, iters=20, reps=5):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn(); s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(iters): fn()
        g.replay(); s.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(reps): g.replay()
        e1.record(s)
    s.synchronize()
    return e0.elapsed_time(e1) / (iters * reps) * 1e3
for M, N, K in [(8192, 768, 768), (8192, 3072, 768), (8192, 768, 3072), (8192, 2304, 768), (4096,4096,4096)]:
    a = torch.randn(M, K, device="cuda").bfloat16(); b = torch.randn(N, K, device="cuda").bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    t1 = graph_time(lambda: gemm_tma(a, b, out=out)); t2 = graph_time(lambda: torch.matmul(a, b.t(), out=out))
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(200): gemm_tma(a, b, out=out)
    host = (time.perf_counter() - t0) / 200 * 1e6; torch.cuda.synchronize()
    fl = 2.0 * M * N * K
    print(f"{M}x{N}x{K}: tfdl {t1:.1f} us ({fl/t1/1e6:.0f} TF)  cublas {t2:.1f} us ({fl/t2/1e6:.0f} TF)  host/call {host:.1f} us")
