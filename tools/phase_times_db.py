"""Phase breakdown of the fused dense block (library built with -DTFDL_PROFILE_PHASES)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tfdl.ops import _lib, cnn
from tfdl.ops.structs import make_dims
B = 64; d = make_dims(B)
a2 = torch.relu(torch.randn(B, 1600, device="cuda")).bfloat16(); idx2 = torch.randint(0, 4, (B, 1600), device="cuda", dtype=torch.uint8)
w1 = (torch.randn(128, 1600, device="cuda") * 0.03).bfloat16(); w2 = (torch.randn(10, 128, device="cuda") * 0.2).bfloat16()
b1 = torch.zeros(128, device="cuda"); b2 = torch.zeros(10, device="cuda"); y = torch.randint(0, 10, (B,), device="cuda", dtype=torch.int32)
sync = torch.zeros(4, dtype=torch.int32, device="cuda")
for _ in range(5):
    cnn.dense_block(a2, idx2, w1, w2, b1, b2, y, d, sync=sync)
torch.cuda.synchronize()
buf = (C.c_longlong * (64 * 16))()
assert _lib.kernels().tfdl_debug_read_phases_db(buf) == 0, "rebuild with TFDL_NVCC_EXTRA=-DTFDL_PROFILE_PHASES"
t = np.array(buf[:], dtype=np.int64).reshape(64, 16)[:25, :15]
names = ["alloc+init", "tma+fwd mma", "RED epilogue", "grid barrier", "build H", "logits mma", "softmax+DL", "dH mma",
         "dH epilogue", "bwd mmas", "dgrad epilogue", "wgrad epilogue", "cta0 extras", "dealloc"]
dt = np.diff(t, axis=1)
for i, n in enumerate(names):
    print(f"{n:16s} median {np.median(dt[:, i]):7.0f}   max {dt[:, i].max():7.0f}   cta0 {dt[0, i]:7.0f}")
print("total median", np.median(t[:, 14] - t[:, 0]), "max", (t[:, 14] - t[:, 0]).max())
