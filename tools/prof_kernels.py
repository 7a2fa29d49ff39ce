"""Run each CNN step kernel a few times on random data (target for `ncu -k regex:<name>`)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tfdl.ops import cnn
from tfdl.ops.structs import make_dims
from tfdl.ops.gemm import gemm_gather

B = int(os.environ.get("B", "64"))
d = make_dims(B)
torch.manual_seed(0)
dev = "cuda"
x = torch.rand(B, 28, 28, device=dev)
w1 = torch.randn(32, 9, device=dev) * 0.3
b1 = torch.zeros(32, device=dev)
w2 = (torch.randn(64, 3, 3, 32, device=dev) * 0.1).bfloat16()
b2 = torch.zeros(64, device=dev)
wd = (torch.randn(128, 1600, device=dev) * 0.05).bfloat16()
bd = torch.zeros(128, device=dev)
wo = torch.randn(10, 128, device=dev) * 0.1
bo = torch.zeros(10, device=dev)
y = torch.randint(0, 10, (B,), device=dev, dtype=torch.int32)
for it in range(int(os.environ.get("ITERS", "5"))):
    a1, i1 = cnn.conv1_fwd(x, w1, b1, d)
    a2, i2 = cnn.conv2_fwd(a1, w2, b2, d)
    acc = gemm_gather(wd, a2.view(B, 1600), k_split=25)  # [128, B] (not transposed here)
    acc = acc.t().contiguous()
    dh, gw2, gb2, gb1, met, so = cnn.head(acc, bd, wo, bo, y, d)
    dy2 = cnn.dense1_dgrad(dh, wd, a2.view(B, 1600), i2.view(B, 1600), d)
    gwc, gbc = cnn.conv2_wgrad(a1, dy2, d)
    g1 = cnn.conv2_dgrad(dy2, w2, a1, d)
    gw1, gb1c = cnn.conv1_wgrad(x, g1, i1, d)
torch.cuda.synchronize()
print("ok", float(so[0]))

