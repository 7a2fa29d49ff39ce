"""Phase breakdown of conv2_fwd from in-kernel clock64 stamps (library built with -DTFDL_PROFILE_PHASES)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tfdl.ops import _lib, cnn
from tfdl.ops.structs import make_dims
d = make_dims(64)
x = torch.rand(64, 28, 28, device="cuda"); w1 = torch.randn(32, 9, device="cuda"); b1 = torch.zeros(32, device="cuda")
w2 = (torch.randn(64, 3, 3, 32, device="cuda") * 0.1).bfloat16(); b2 = torch.zeros(64, device="cuda")
a1, _ = cnn.conv1_fwd(x, w1, b1, d)
for _ in range(5):
    cnn.conv2_fwd(a1, w2, b2, d)
torch.cuda.synchronize()
buf = (C.c_longlong * (256 * 8))()
assert _lib.kernels().tfdl_debug_read_phases(buf) == 0, "rebuild with TFDL_NVCC_EXTRA=-DTFDL_PROFILE_PHASES"
t = np.array(buf[:], dtype=np.int64).reshape(256, 8)[:50]
names = ["start", "tile_begin", "gather issued", "staging_done", "mma issued", "mma done", "epilogue", "tile_end"]
dt = np.diff(t, axis=1)
print("phase                cycles(median over CTAs)")
for i, n in enumerate(names[1:]):
    print(f"{n:18s} {np.median(dt[:, i]):8.0f}")
print("total", np.median(t[:, 7] - t[:, 0]), "cycles; CTA start spread", t[:, 0].max() - t[:, 0].min())
