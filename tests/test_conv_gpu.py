"""Implicit-GEMM Conv2D (TMA im2col + tcgen05) and the pooling kernels against a float64 CPU oracle.

Oracle: torch CPU convolution / pooling in float64 on the bf16-rounded inputs, so the only differences are the fp32
accumulation order and the final bf16 rounding of the outputs (relative 2^-8) - the bands below are derived from that,
not tuned. Shapes cover the reference CNN's second convolution with other filter counts and 'same' padding
(tf_dist_example.py:41-44) and the convolution classes of Keras ResNet-50 (3x3/s1 same, 1x1/s2, 3x3/s2, 7x7/s2)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

BF16_EPS = 2.0 ** -8


def _keras_pads(size, k, s, padding):
    if padding == "valid":
        return 0, 0
    out = -(-size // s)
    tot = max((out - 1) * s + k - size, 0)
    return tot // 2, tot - tot // 2


CASES = [
    # N, H, W, C, K, R, S, stride, padding
    (4, 14, 14, 64, 64, 3, 3, 1, "same"),
    (2, 13, 13, 32, 64, 3, 3, 1, "valid"),     # the reference CNN's conv2
    (2, 13, 13, 48, 48, 3, 3, 1, "same"),      # filters=48: channel blocks with a zero-filled tail
    (3, 10, 12, 128, 72, 3, 3, 1, "same"),     # K tail, non-square image
    (2, 28, 28, 128, 256, 1, 1, 2, "valid"),   # ResNet shortcut / first 1x1 of a stage
    (2, 16, 16, 64, 128, 3, 3, 2, "same"),     # strided 3x3 (forward + wgrad)
    (2, 32, 32, 8, 64, 7, 7, 2, "same"),       # 7x7/s2 with the smallest channel block
    (8, 7, 7, 512, 512, 3, 3, 1, "same"),      # deep layer: many K blocks, pixels wrap over images inside one tile
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "x".join(str(v) for v in c))
def test_conv2d_fwd_dgrad_wgrad_vs_fp64(case):
    from tfdl.ops import conv as C

    N, H, W, Ci, K, R, S, st, padding = case
    g = torch.Generator().manual_seed(hash(case) % (1 << 31))
    x = torch.randn((N, H, W, Ci), generator=g).bfloat16()
    w = (torch.randn((K, R, S, Ci), generator=g) / np.sqrt(R * S * Ci)).bfloat16()
    b = torch.randn((K,), generator=g)
    pt, pb = _keras_pads(H, R, st, padding)
    pl, pr = _keras_pads(W, S, st, padding)
    pads = (pt, pl, pb, pr)
    # ---- oracle (float64, CPU) ----
    x64 = x.double().permute(0, 3, 1, 2).requires_grad_(True)
    w64 = w.double().permute(0, 3, 1, 2).requires_grad_(True)
    y64 = F.conv2d(F.pad(x64, (pl, pr, pt, pb)), w64, b.double(), stride=st)
    dy = torch.randn(y64.shape, generator=g).bfloat16()
    y64.backward(dy.double())
    y_ref = y64.detach().permute(0, 2, 3, 1)
    dx_ref = x64.grad.permute(0, 2, 3, 1)
    dw_ref = w64.grad.permute(0, 2, 3, 1)
    # ---- kernels ----
    dev = torch.device("cuda")
    xd, wd, bd = x.to(dev), w.to(dev), b.to(dev)
    dyd = dy.permute(0, 2, 3, 1).contiguous().to(dev)
    y = C.conv2d_fwd(xd, wd, bd, (st, st), pads)
    torch.cuda.synchronize()
    assert y.shape == y_ref.shape
    scale = float(y_ref.abs().max())
    torch.testing.assert_close(y.double().cpu(), y_ref, rtol=2 * BF16_EPS, atol=2 * BF16_EPS * scale * 0.05 + 1e-6)
    yr = C.conv2d_fwd(xd, wd, bd, (st, st), pads, act="relu")
    torch.testing.assert_close(yr.double().cpu(), y_ref.clamp_min(0), rtol=2 * BF16_EPS, atol=2 * BF16_EPS * scale * 0.05 + 1e-6)
    dw = C.conv2d_wgrad(xd, dyd, tuple(w.shape), (st, st), pads)
    torch.cuda.synchronize()
    sw = float(dw_ref.abs().max())
    torch.testing.assert_close(dw.double().cpu(), dw_ref, rtol=1e-4, atol=2e-5 * sw)  # fp32 sums of exact bf16 products
    if st == 1 or (R, S) == (1, 1):
        dx = C.conv2d_dgrad(dyd, wd, (N, H, W, Ci), (st, st), pads)
        torch.cuda.synchronize()
        sx = float(dx_ref.abs().max())
        torch.testing.assert_close(dx.double().cpu(), dx_ref, rtol=2 * BF16_EPS, atol=2 * BF16_EPS * sx * 0.05 + 1e-6)


def test_conv2d_autograd_and_first_layer():
    from tfdl.ops import conv as C

    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(3)
    # first layer of the reference CNN: 1 input channel, 3x3 valid (tf_dist_example.py:41) through the patch-matrix path
    x = torch.rand((16, 28, 28, 1), generator=g).bfloat16()
    w = (torch.randn((32, 3, 3, 1), generator=g) / 3).bfloat16()
    b = torch.randn((32,), generator=g)
    w64 = w.double().permute(0, 3, 1, 2).requires_grad_(True)
    y64 = torch.relu(F.conv2d(x.double().permute(0, 3, 1, 2), w64, b.double()))
    dy = torch.randn(y64.shape, generator=g).bfloat16()
    y64.backward(dy.double())
    wd = w.float().to(dev).requires_grad_(True)
    bd = b.to(dev).requires_grad_(True)
    y = C.conv2d_small_cin(x.to(dev), wd, bd, (1, 1), (0, 0, 0, 0), "relu")
    y.backward(dy.permute(0, 2, 3, 1).contiguous().to(dev))
    torch.cuda.synchronize()
    ref = y64.detach().permute(0, 2, 3, 1)
    torch.testing.assert_close(y.double().cpu(), ref, rtol=2 * BF16_EPS, atol=1e-3 * float(ref.abs().max()))
    dw_ref = w64.grad.permute(0, 2, 3, 1)
    torch.testing.assert_close(wd.grad.double().cpu(), dw_ref, rtol=1e-2, atol=1e-2 * float(dw_ref.abs().max()))
    # autograd wrapper of the implicit-GEMM path (bias + relu + all three gradients)
    x2 = torch.randn((4, 12, 12, 64), generator=g).bfloat16()
    w2 = (torch.randn((64, 3, 3, 64), generator=g) / 24).bfloat16()
    x64 = x2.double().permute(0, 3, 1, 2).requires_grad_(True)
    w264 = w2.double().permute(0, 3, 1, 2).requires_grad_(True)
    b2 = torch.randn((64,), generator=g)
    b64 = b2.double().requires_grad_(True)
    y2 = torch.relu(F.conv2d(x64, w264, b64, padding=1))
    dy2 = torch.randn(y2.shape, generator=g).bfloat16()
    y2.backward(dy2.double())
    xd = x2.to(dev).requires_grad_(True)
    wd2 = w2.float().to(dev).requires_grad_(True)
    bd2 = b2.to(dev).requires_grad_(True)
    out = C.conv2d(xd, wd2, bd2, (1, 1), (1, 1, 1, 1), "relu")
    out.backward(dy2.permute(0, 2, 3, 1).contiguous().to(dev))
    torch.cuda.synchronize()
    for got, want in ((xd.grad, x64.grad.permute(0, 2, 3, 1)), (wd2.grad, w264.grad.permute(0, 2, 3, 1)), (bd2.grad, b64.grad)):
        torch.testing.assert_close(got.double().cpu(), want, rtol=2e-2, atol=1e-2 * float(want.abs().max()))


POOLS = [
    # N, H, W, C, window, stride, padding
    (4, 26, 26, 32, 2, 2, "valid"),   # the reference CNN's first pool
    (4, 11, 11, 64, 2, 2, "valid"),   # second pool: last row / column dropped
    (2, 112, 112, 64, 3, 2, "same"),  # ResNet-50's 3x3/s2 pool (overlapping windows)
    (2, 9, 7, 48, 3, 1, "same"),
]


@pytest.mark.parametrize("case", POOLS, ids=lambda c: "x".join(str(v) for v in c))
def test_maxpool_vs_fp64(case):
    from tfdl.ops import pool as P

    N, H, W, C, k, s, padding = case
    g = torch.Generator().manual_seed(11)
    x = torch.randn((N, H, W, C), generator=g).bfloat16()
    pt, pb = _keras_pads(H, k, s, padding)
    pl, pr = _keras_pads(W, k, s, padding)
    x64 = x.double().permute(0, 3, 1, 2).requires_grad_(True)
    y64 = F.max_pool2d(F.pad(x64, (pl, pr, pt, pb), value=float("-inf")), k, s)
    dy = torch.randn(y64.shape, generator=g).bfloat16()
    y64.backward(dy.double())
    dev = torch.device("cuda")
    xd = x.to(dev).requires_grad_(True)
    y = P.max_pool(xd, (k, k), (s, s), (pt, pl), (y64.shape[2], y64.shape[3]))
    y.backward(dy.permute(0, 2, 3, 1).contiguous().to(dev))
    torch.cuda.synchronize()
    assert torch.equal(y.double().cpu(), y64.detach().permute(0, 2, 3, 1))  # a selection: exact
    # overlapping windows add up to 4 bf16 gradients in fp32 and round once
    torch.testing.assert_close(xd.grad.double().cpu(), x64.grad.permute(0, 2, 3, 1), rtol=2 * BF16_EPS, atol=1e-6)


@pytest.mark.parametrize("shape", [(8, 7, 7, 2048), (4, 5, 5, 64), (3, 3, 4, 48)])
def test_global_avgpool_vs_fp64(shape):
    from tfdl.ops import pool as P

    g = torch.Generator().manual_seed(5)
    x = torch.randn(shape, generator=g).bfloat16()
    dy = torch.randn((shape[0], shape[3]), generator=g).bfloat16()
    dev = torch.device("cuda")
    xd = x.to(dev).requires_grad_(True)
    y = P.global_avg_pool(xd)
    y.backward(dy.to(dev))
    torch.cuda.synchronize()
    ref = x.double().mean(dim=(1, 2))
    torch.testing.assert_close(y.double().cpu(), ref, rtol=2 * BF16_EPS, atol=1e-3)
    dref = (dy.double() / (shape[1] * shape[2]))[:, None, None, :].expand(shape)
    torch.testing.assert_close(xd.grad.double().cpu(), dref, rtol=2 * BF16_EPS, atol=1e-6)
