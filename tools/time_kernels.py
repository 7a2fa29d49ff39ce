"""GPU-side time of each step kernel: 50 copies of one kernel captured in a CUDA graph (no CPU in the
loop, warm caches), replayed and timed with CUDA events. Includes the serialisation gap between
dependent launches, i.e. what the kernel costs inside the step graph."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tfdl.ops import _lib, cnn
from tfdl.ops.structs import make_dims

B = int(os.environ.get("B", "64"))
d = make_dims(B)
dev = "cuda"
torch.manual_seed(0)
L = _lib.kernels()
p = _lib.ptr
x = torch.rand(B, 28, 28, device=dev)
w1, b1 = torch.randn(32, 9, device=dev) * 0.3, torch.zeros(32, device=dev)
w2, b2 = (torch.randn(64, 3, 3, 32, device=dev) * 0.1).bfloat16(), torch.zeros(64, device=dev)
wd, bd = (torch.randn(128, 1600, device=dev) * 0.05).bfloat16(), torch.zeros(128, device=dev)
wo, bo = torch.randn(10, 128, device=dev) * 0.1, torch.zeros(10, device=dev)
y = torch.randint(0, 10, (B,), device=dev, dtype=torch.int32)
a1, i1 = cnn.conv1_fwd(x, w1, b1, d)
a2, i2 = cnn.conv2_fwd(a1, w2, b2, d)
acc = torch.zeros(B, 128, device=dev)
dh = torch.zeros(B, 128, dtype=torch.bfloat16, device=dev)
gw2, gb2, gb1 = torch.zeros(10, 128, device=dev), torch.zeros(10, device=dev), torch.zeros(128, device=dev)
met, so = torch.zeros(3, device=dev), torch.zeros(2, device=dev)
dy2 = torch.zeros(B, 13, 13, 64, dtype=torch.bfloat16, device=dev)
dy2[:, :10, :10] = (torch.randn(B, 10, 10, 64, device=dev) * 0.1).bfloat16()
g1 = torch.zeros(B, 13, 13, 32, dtype=torch.bfloat16, device=dev)
gW1 = torch.zeros(128, 1600, device=dev)
gwc, gbc = torch.zeros(64, 288, device=dev), torch.zeros(64, device=dev)
gw1c, gb1c = torch.zeros(32, 9, device=dev), torch.zeros(32, device=dev)
empty = torch.zeros(4, device=dev)
D = C.byref(d)
st = torch.cuda.Stream()
S = lambda: C.c_void_p(st.cuda_stream)  # noqa: E731

kernels = {
    "memset_16B": lambda: empty.zero_(),
    "conv1_fwd": lambda: L.tfdl_cnn_conv1_fwd(p(x), p(w1), p(b1), p(a1), p(i1), None, D, S()),
    "conv2_fwd": lambda: L.tfdl_cnn_conv2_fwd(p(a1), p(w2), p(b2), p(a2), p(i2), D, S()),
    "dense1_fwd": lambda: L.tfdl_gemm_bf16_gather_ex(p(wd), p(a2), p(acc), None, None, 128, B, 1600, 1600, 1600, 128, 0, 0, 0, 25, 0, 64, 1, S()),
    "head": lambda: L.tfdl_cnn_head(p(acc), p(bd), p(wo), p(bo), p(y), p(dh), p(gw2), p(gb2), p(gb1), p(met), p(so), D, S()),
    "dense1_wgrad": lambda: L.tfdl_gemm_bf16_gather_ex(p(dh), p(a2), p(gW1), None, None, 128, 1600, B, 128, 1600, 1600, 1, 1, 0, 1, 64, 0, 0, S()),
    "dense1_dgrad": lambda: L.tfdl_cnn_dense1_dgrad(p(dh), p(wd), p(a2), p(i2), p(dy2), D, S()),
    "conv2_wgrad": lambda: L.tfdl_cnn_conv2_wgrad(p(a1), p(dy2), p(gwc), p(gbc), D, S()),
    "conv2_dgrad": lambda: L.tfdl_cnn_conv2_dgrad(p(dy2), p(w2), p(a1), p(g1), D, S()),
    "conv2_fwd_tma": lambda: L.tfdl_cnn_conv2_fwd_tma(p(a1), p(w2), p(b2), p(a2), p(i2), D, S()),
    "conv2_wgrad_tma": lambda: L.tfdl_cnn_conv2_wgrad_tma(p(a1), p(dy2), p(gwc), p(gbc), D, S()),
    "conv2_dgrad_tma": lambda: L.tfdl_cnn_conv2_dgrad_tma(p(dy2), p(w2), p(a1), p(g1), D, S()),
    "conv12_fwd_tma": lambda: L.tfdl_cnn_conv12_fwd_tma(p(x), p(w1), p(b1), p(w2), p(b2), p(a1), p(i1), p(a2), p(i2), None, D, S()),
    "conv21_bwd_tma": lambda: L.tfdl_cnn_conv21_bwd_tma(p(dy2), p(w2), p(a1), p(i1), p(x), p(gw1c), p(gb1c), None, D, S()),
    "conv1_wgrad": lambda: L.tfdl_cnn_conv1_wgrad(p(x), p(g1), p(i1), p(gw1c), p(gb1c), D, S()),
}
N = 50
total = 0.0
for name, fn in kernels.items():
    with torch.cuda.stream(st):
        fn()
        st.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for _ in range(N):
                fn()
        for _ in range(3):
            g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(10):
            g.replay()
        e1.record(st)
    st.synchronize()
    us = 1000 * e0.elapsed_time(e1) / (10 * N)
    if name != "memset_16B" and name not in ("dense1_wgrad", "conv2_wgrad", "conv2_wgrad_tma", "conv2_fwd", "conv2_dgrad"):
        total += us
    print(f"{name:14s} {us:7.2f} us per launch (in-graph, chained)")
print(f"critical-path kernels sum: {total:.1f} us (+ all-reduce)")
acc2 = [torch.zeros(B, 128, device=dev) for _ in range(2)]
sync = torch.zeros(4, dtype=torch.int32, device=dev)
w2h = wo.bfloat16().contiguous()
def fused():
    return L.tfdl_cnn_dense_block(p(a2), p(i2), p(wd), p(w2h), p(bd), p(bo), p(y), p(acc2[0]), p(acc2[1]), p(sync), p(dy2),
                                  p(gW1), p(gb1), p(gw2), p(gb2), p(met), p(so), None, D, S())
with torch.cuda.stream(st):
    fused(); st.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=st):
        for _ in range(N):
            fused()
            acc2[0].zero_()
    for _ in range(3): g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(10): g.replay()
    e1.record(st)
st.synchronize()
print(f"dense_block    {1000 * e0.elapsed_time(e1) / (10 * N) - 0.8:7.2f} us per launch (in-graph, chained; minus a memset node)")
