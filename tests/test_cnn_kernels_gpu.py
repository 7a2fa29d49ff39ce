"""Every CNN step kernel (csrc/kernels/cnn.cu) against a plain PyTorch fp32 reference of the same op.
Inputs that the kernels consume as bf16 are rounded to bf16 first, so the comparison isolates the
kernel arithmetic (fp32 accumulation in TMEM) from the storage precision."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

B = 64


def _dims(b=B):
    from tfdl.ops.structs import make_dims

    return make_dims(b)


def _bf(t):
    return t.bfloat16()


def test_conv1_fwd_matches_conv_bias_pool_relu():
    from tfdl.ops import cnn

    d = _dims()
    torch.manual_seed(0)
    x = torch.rand(B, 28, 28, device="cuda")
    w = torch.randn(32, 9, device="cuda") * 0.3
    b = torch.randn(32, device="cuda") * 0.1
    a1, idx = cnn.conv1_fwd(x, w, b, d)
    y = F.conv2d(x[:, None], w.view(32, 1, 3, 3), b)  # [B,32,26,26]
    ref, ridx = F.max_pool2d(y, 2, return_indices=True)
    ref = torch.relu(ref).permute(0, 2, 3, 1)
    torch.testing.assert_close(a1.float(), ref, rtol=1e-2, atol=1e-2)
    # argmax: the value at the reported window position equals the window maximum
    yw = y.permute(0, 2, 3, 1).reshape(B, 13, 2, 13, 2, 32).permute(0, 1, 3, 5, 2, 4).reshape(B, 13, 13, 32, 4)
    picked = torch.gather(yw, 4, idx.long()[..., None])[..., 0]
    torch.testing.assert_close(picked, yw.max(-1).values, rtol=0, atol=1e-5)


@pytest.mark.parametrize("impl", ["gather", "tma"])
def test_conv2_fwd_tcgen05_matches_reference(impl):
    from tfdl.ops import cnn

    d = _dims()
    torch.manual_seed(1)
    a1 = _bf(torch.relu(torch.randn(B, 13, 13, 32, device="cuda")))
    w = _bf(torch.randn(64, 3, 3, 32, device="cuda") * 0.1)
    b = torch.randn(64, device="cuda") * 0.1
    a2, idx = cnn.conv2_fwd(a1, w, b, d, impl=impl)
    y = F.conv2d(a1.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), b)  # [B,64,11,11]
    ref = torch.relu(F.max_pool2d(y, 2)).permute(0, 2, 3, 1)  # floor: row/col 10 dropped
    torch.testing.assert_close(a2.float(), ref, rtol=2e-2, atol=2e-2)
    yw = y[:, :, :10, :10].permute(0, 2, 3, 1).reshape(B, 5, 2, 5, 2, 64).permute(0, 1, 3, 5, 2, 4).reshape(B, 5, 5, 64, 4)
    picked = torch.gather(yw, 4, idx.long()[..., None])[..., 0]
    torch.testing.assert_close(picked, yw.max(-1).values, rtol=0, atol=2e-3)


def test_head_loss_and_small_gradients():
    from tfdl.ops import cnn

    d = _dims()
    torch.manual_seed(2)
    acc = torch.randn(B, 128, device="cuda")
    b1 = torch.randn(128, device="cuda") * 0.1
    w2 = (torch.randn(10, 128, device="cuda") * 0.2).requires_grad_()
    b2 = (torch.randn(10, device="cuda") * 0.1).requires_grad_()
    y = torch.randint(0, 10, (B,), device="cuda", dtype=torch.int32)
    pre = (acc + b1).clone().requires_grad_()
    logits = torch.relu(pre) @ w2.t() + b2
    loss = F.cross_entropy(logits, y.long(), reduction="mean")
    loss.backward()
    acc_in = acc.clone()
    dh, g_w2, g_b2, g_b1, metrics, step_out = cnn.head(acc_in, b1, w2.detach(), b2.detach(), y, d)
    assert torch.count_nonzero(acc_in) == 0  # scratch handed back clean
    torch.testing.assert_close(step_out[0], loss.detach(), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(metrics[0] / B, loss.detach(), rtol=1e-4, atol=1e-5)
    assert metrics[2].item() == B
    assert metrics[1].item() == (logits.argmax(-1) == y.long()).sum().item()
    torch.testing.assert_close(g_w2, w2.grad, rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(g_b2, b2.grad, rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(g_b1, pre.grad.sum(0), rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(dh.float(), pre.grad, rtol=1e-2, atol=1e-5)


def test_dense1_dgrad_unpool_epilogue():
    from tfdl.ops import cnn

    d = _dims()
    torch.manual_seed(3)
    dh = _bf(torch.randn(B, 128, device="cuda") * 0.1)
    w1 = _bf(torch.randn(128, 1600, device="cuda") * 0.05)
    a2 = _bf(torch.relu(torch.randn(B, 1600, device="cuda")))
    idx2 = torch.randint(0, 4, (B, 1600), device="cuda", dtype=torch.uint8)
    dy2 = cnn.dense1_dgrad(dh, w1, a2, idx2, d)
    da2 = (dh.float() @ w1.float()).view(B, 5, 5, 64)
    a2v, iv = a2.view(B, 5, 5, 64), idx2.view(B, 5, 5, 64)
    ref = torch.zeros(B, 10, 10, 64, device="cuda")
    for q in range(4):
        ref[:, (q >> 1)::2, (q & 1)::2, :] = torch.where((iv == q) & (a2v > 0), da2, torch.zeros_like(da2))
    assert dy2.shape == (B, 13, 13, 64)  # stored on the conv2 input grid; the margin stays zero
    torch.testing.assert_close(dy2[:, :10, :10].float(), ref, rtol=2e-2, atol=2e-3)
    assert torch.count_nonzero(dy2[:, 10:]) == 0 and torch.count_nonzero(dy2[:, :, 10:]) == 0


def _sparse_dy2():
    """(dy2 on the 13x13 grid as the kernels expect it, the dense 10x10 core for the reference)"""
    core = torch.randn(B, 10, 10, 64, device="cuda") * 0.1
    core = _bf(core * (torch.rand(B, 10, 10, 64, device="cuda") < 0.25))
    grid = torch.zeros(B, 13, 13, 64, dtype=torch.bfloat16, device="cuda")
    grid[:, :10, :10] = core
    return grid, core


@pytest.mark.parametrize("impl", ["gather", "tma"])
def test_conv2_wgrad_split_reduction(impl):
    from tfdl.ops import cnn

    d = _dims()
    torch.manual_seed(4)
    a1 = _bf(torch.relu(torch.randn(B, 13, 13, 32, device="cuda")))
    grid, dy2 = _sparse_dy2()
    g_w, g_b = cnn.conv2_wgrad(a1, grid, d, impl=impl)
    xin = a1.float().permute(0, 3, 1, 2)[:, :, :12, :12].contiguous()
    gw_ref = torch.nn.grad.conv2d_weight(xin, (64, 32, 3, 3), dy2.float().permute(0, 3, 1, 2).contiguous())
    torch.testing.assert_close(g_w, gw_ref.permute(0, 2, 3, 1), rtol=5e-3, atol=5e-3)
    torch.testing.assert_close(g_b, dy2.float().sum((0, 1, 2)), rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("impl", ["gather", "tma"])
def test_conv2_dgrad_relu_masked(impl):
    from tfdl.ops import cnn

    d = _dims()
    torch.manual_seed(5)
    a1 = _bf(torch.relu(torch.randn(B, 13, 13, 32, device="cuda")))
    w = _bf(torch.randn(64, 3, 3, 32, device="cuda") * 0.1)
    grid, dy2 = _sparse_dy2()
    g1 = cnn.conv2_dgrad(grid, w, a1, d, impl=impl)
    gi = torch.nn.grad.conv2d_input((B, 32, 12, 12), w.float().permute(0, 3, 1, 2).contiguous(),
                                    dy2.float().permute(0, 3, 1, 2).contiguous())
    ref = torch.zeros(B, 13, 13, 32, device="cuda")
    ref[:, :12, :12, :] = gi.permute(0, 2, 3, 1)
    ref = ref * (a1.float() > 0)
    torch.testing.assert_close(g1.float(), ref, rtol=2e-2, atol=2e-3)


def test_conv1_wgrad_through_argmax():
    from tfdl.ops import cnn

    d = _dims()
    torch.manual_seed(6)
    x = torch.rand(B, 28, 28, device="cuda")
    g1 = _bf(torch.randn(B, 13, 13, 32, device="cuda") * 0.1)
    idx1 = torch.randint(0, 4, (B, 13, 13, 32), device="cuda", dtype=torch.uint8)
    g_w, g_b = cnn.conv1_wgrad(x, g1, idx1, d)
    dy1 = torch.zeros(B, 26, 26, 32, device="cuda")
    for q in range(4):
        dy1[:, (q >> 1)::2, (q & 1)::2, :] = torch.where(idx1 == q, g1.float(), torch.zeros_like(g1.float()))
    gw_ref = torch.nn.grad.conv2d_weight(x[:, None], (32, 1, 3, 3), dy1.permute(0, 3, 1, 2).contiguous())
    torch.testing.assert_close(g_w, gw_ref.view(32, 9), rtol=2e-3, atol=2e-3)
    torch.testing.assert_close(g_b, g1.float().sum((0, 1, 2)), rtol=1e-3, atol=1e-3)


def test_fused_allreduce_sgd_single_gpu_matches_sgd():
    """world = 1: the fused kernel is a plain (momentum) SGD + bf16 shadow refresh + grad clear."""
    from tfdl.ops.allreduce import P2PFusedAllReduce, SymmetricBuffer

    dev = torch.device("cuda", 0)
    ops = P2PFusedAllReduce(dev)
    n = 225280
    params, grads = ops.alloc_arena(n, dev), ops.alloc_arena(n, dev)
    torch.manual_seed(7)
    params.copy_(torch.randn(n, device=dev))
    st = ops.fused_state(params, grads, momentum=True)
    st.set_hyper(0.1, 0.9, False, 1.0)
    w_ref, v_ref = params.clone(), torch.zeros(n, device=dev)
    for step in range(4):
        par = step & 1
        g = torch.randn(n, device=dev)
        st.grads[par].tensor.copy_(g)
        st.step(par)
        v_ref = 0.9 * v_ref - 0.1 * g
        w_ref = w_ref + v_ref
        torch.cuda.synchronize()
        torch.testing.assert_close(params, w_ref, rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(st.shadow.tensor.float(), params.bfloat16().float(), rtol=0, atol=0)
        assert torch.count_nonzero(st.grads[par ^ 1].tensor) == 0
        assert st.counters[0].item() == step + 1


@pytest.mark.parametrize("b", [64, 32, 128])
def test_fused_dense_block_matches_fp32_reference(b):
    """Dense(128,relu)+Dense(10)+xent forward AND backward in one kernel vs torch autograd."""
    from tfdl.ops import cnn

    d = _dims(b)
    torch.manual_seed(8)
    a2 = _bf(torch.relu(torch.randn(b, 1600, device="cuda")))
    idx2 = torch.randint(0, 4, (b, 1600), device="cuda", dtype=torch.uint8)
    w1 = _bf(torch.randn(128, 1600, device="cuda") * 0.03)
    w2 = _bf(torch.randn(10, 128, device="cuda") * 0.2)
    b1 = torch.randn(128, device="cuda") * 0.1
    b2 = torch.randn(10, device="cuda") * 0.1
    y = torch.randint(0, 10, (b,), device="cuda", dtype=torch.int32)
    sync = torch.zeros(4, dtype=torch.int32, device="cuda")
    for rep in range(2):  # twice: the barrier counter and the accumulator hand-over must survive reuse
        out = cnn.dense_block(a2, idx2, w1, w2, b1, b2, y, d, sync=sync)
        torch.cuda.synchronize()
        A2 = a2.float().requires_grad_()
        W1, W2 = w1.float().requires_grad_(), w2.float().requires_grad_()
        B1, B2 = b1.clone().requires_grad_(), b2.clone().requires_grad_()
        pre = A2 @ W1.t() + B1
        pre.retain_grad()
        # the kernel feeds h to the second layer in bf16
        h = torch.relu(pre)
        hq = h + (h.detach().bfloat16().float() - h.detach())
        logits = hq @ W2.t() + B2
        loss = F.cross_entropy(logits, y.long())
        loss.backward()
        torch.testing.assert_close(out["step_out"][0], loss.detach(), rtol=2e-3, atol=2e-3)
        assert out["metrics"][2].item() == b
        assert abs(out["metrics"][1].item() - (logits.argmax(-1) == y.long()).sum().item()) <= 1
        torch.testing.assert_close(out["g_b2"], B2.grad, rtol=2e-2, atol=2e-4)
        torch.testing.assert_close(out["g_w2"], W2.grad, rtol=3e-2, atol=3e-4)
        torch.testing.assert_close(out["g_b1"], B1.grad, rtol=3e-2, atol=3e-4)
        torch.testing.assert_close(out["dh"].float(), pre.grad, rtol=3e-2, atol=1e-4)
        torch.testing.assert_close(out["g_w1"], W1.grad, rtol=5e-2, atol=6e-4)
        da2 = A2.grad.view(b, 5, 5, 64)
        a2v, iv = a2.view(b, 5, 5, 64), idx2.view(b, 5, 5, 64)
        ref = torch.zeros(b, 10, 10, 64, device="cuda")
        for q in range(4):
            ref[:, (q >> 1)::2, (q & 1)::2, :] = torch.where((iv == q) & (a2v > 0), da2, torch.zeros_like(da2))
        torch.testing.assert_close(out["dy2"][:, :10, :10].float(), ref, rtol=5e-2, atol=6e-4)
        assert torch.count_nonzero(out["acc_next"]) == 0  # the other parity's accumulator was cleared
        assert sync[0].item() == 25 * (rep + 1)


def test_fused_conv_kernels_match_the_unfused_ones():
    """conv1 fused into conv2 forward, and conv1 wgrad fused into conv2 dgrad, against the separately
    validated kernels (bitwise for the forward, RED-order tolerance for the weight gradient)."""
    from tfdl.ops import cnn

    d = _dims()
    torch.manual_seed(9)
    x = torch.rand(B, 28, 28, device="cuda")
    w1 = torch.randn(32, 9, device="cuda") * 0.3
    b1 = torch.randn(32, device="cuda") * 0.1
    w2 = _bf(torch.randn(64, 3, 3, 32, device="cuda") * 0.1)
    b2 = torch.randn(64, device="cuda") * 0.1
    a1r, i1r = cnn.conv1_fwd(x, w1, b1, d)
    a2r, i2r = cnn.conv2_fwd(a1r, w2, b2, d, impl="tma")
    a1, i1, a2, i2 = cnn.conv12_fwd(x, w1, b1, w2, b2, d)
    assert torch.equal(a1, a1r) and torch.equal(i1, i1r)
    assert torch.equal(a2, a2r) and torch.equal(i2, i2r)
    grid, _ = _sparse_dy2()
    g1r = cnn.conv2_dgrad(grid, w2, a1, d, impl="tma")
    gwr, gbr = cnn.conv1_wgrad(x, g1r, i1, d)
    gw, gb, g1 = cnn.conv21_bwd(grid, w2, a1, i1, x, d)
    assert torch.equal(g1, g1r)
    torch.testing.assert_close(gw, gwr, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(gb, gbr, rtol=1e-4, atol=1e-4)
