"""tcgen05 GEMM kernels vs a plain PyTorch fp32 reference of the same op."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(a, b, a_mn, b_mn):
    A = a.float().t() if a_mn else a.float()
    B = b.float().t() if b_mn else b.float()
    return A @ B.t()


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (64, 128, 1600), (200, 48, 288), (1600, 128, 64), (384, 64, 512)])
def test_gather_gemm_all_majors(M, N, K, a_mn, b_mn):
    from tfdl.ops.gemm import gemm_gather

    torch.manual_seed(0)
    a = torch.randn((K, M) if a_mn else (M, K), device="cuda").bfloat16()
    b = torch.randn((K, N) if b_mn else (N, K), device="cuda").bfloat16()
    out = gemm_gather(a, b, a_mn=a_mn, b_mn=b_mn)
    ref = _ref(a, b, a_mn, b_mn)
    torch.testing.assert_close(out, ref, rtol=2e-3, atol=2e-2 * (K ** 0.5) / 8)


def test_gather_gemm_bias_relu_bf16_out():
    from tfdl.ops.gemm import gemm_gather

    torch.manual_seed(1)
    a = torch.randn(64, 1600, device="cuda").bfloat16()
    b = torch.randn(128, 1600, device="cuda").bfloat16()
    bias = torch.randn(128, device="cuda")
    out = gemm_gather(a, b, bias=bias, relu=True, out_dtype=torch.bfloat16)
    ref = torch.relu(a.float() @ b.float().t() + bias)
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=0.3)


@pytest.mark.parametrize("k_split", [5, 25])
def test_gather_gemm_split_k(k_split):
    from tfdl.ops.gemm import gemm_gather

    torch.manual_seed(2)
    a = torch.randn(64, 1600, device="cuda").bfloat16()
    b = torch.randn(128, 1600, device="cuda").bfloat16()
    out = gemm_gather(a, b, k_split=k_split)
    torch.testing.assert_close(out, a.float() @ b.float().t(), rtol=2e-3, atol=0.1)


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 256, 512), (1000, 200, 1608), (4096, 4096, 1024),
                                   (64, 3072, 768), (8192, 768, 3072), (300, 64, 128)])
def test_tma_gemm(M, N, K):
    from tfdl.ops.gemm import gemm_tma

    torch.manual_seed(3)
    a = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    b = (torch.randn(N, K, device="cuda") * 0.5).bfloat16()
    out = gemm_tma(a, b, out_dtype=torch.float32)
    ref = a.float() @ b.float().t()
    torch.testing.assert_close(out, ref, rtol=2e-3, atol=2e-2 * (K ** 0.5) / 8)


def test_tma_gemm_bias_act_bf16():
    from tfdl.ops.gemm import gemm_tma

    torch.manual_seed(4)
    a = (torch.randn(512, 768, device="cuda") * 0.3).bfloat16()
    b = (torch.randn(3072, 768, device="cuda") * 0.3).bfloat16()
    bias = torch.randn(3072, device="cuda")
    for act, f in (("relu", torch.relu), ("gelu", lambda x: torch.nn.functional.gelu(x, approximate="tanh"))):
        out = gemm_tma(a, b, bias=bias, act=act)
        ref = f(a.float() @ b.float().t() + bias)
        torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=0.15)


@pytest.mark.parametrize("B,N,K,relu", [(64, 128, 1600, True), (16, 10 * 16, 128, False), (256, 256, 512, True)])
def test_dense_forward_splitk_scratch_roundtrip(B, N, K, relu):
    from tfdl.ops.dense import dense_forward

    torch.manual_seed(11)
    x = (torch.randn(B, K, device="cuda") * 0.5).bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
    bias = torch.randn(N, device="cuda")
    scratch = torch.zeros(B * N, device="cuda")
    for _ in range(2):  # the scratch comes back zeroed, so it can be reused as is
        out = dense_forward(x, w, bias, relu=relu, scratch=scratch)
        ref = x.float() @ w.float().t() + bias
        ref = torch.relu(ref) if relu else ref
        torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=0.1)
        assert torch.count_nonzero(scratch) == 0


@pytest.mark.parametrize("a_mn,b_mn", [(False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(256, 256, 512), (1000, 200, 1608), (768, 3072, 4096), (128, 64, 64)])
def test_tma_gemm_mn_major_operands(M, N, K, a_mn, b_mn):
    from tfdl.ops.gemm import gemm_tma

    torch.manual_seed(5)
    a = (torch.randn((K, M) if a_mn else (M, K), device="cuda") * 0.5).bfloat16()
    b = (torch.randn((K, N) if b_mn else (N, K), device="cuda") * 0.5).bfloat16()
    out = gemm_tma(a, b, a_mn=a_mn, b_mn=b_mn, out_dtype=torch.float32)
    torch.testing.assert_close(out, _ref(a, b, a_mn, b_mn), rtol=2e-3, atol=2e-2 * (K ** 0.5) / 8)


@pytest.mark.parametrize("act", [None, "relu", "gelu"])
def test_linear_autograd_all_three_gemms_on_tcgen05(act):
    from tfdl.ops.gemm import linear

    torch.manual_seed(6)
    x = (torch.randn(4, 128, 768, device="cuda") * 0.5).requires_grad_()
    w = (torch.randn(3072, 768, device="cuda") * 0.03).requires_grad_()
    b = torch.randn(3072, device="cuda").requires_grad_()
    y = linear(x, w, b, act)
    gy = torch.randn_like(y.float()) * 0.1
    y.float().backward(gy)
    xr, wr, br = [t.detach().clone().requires_grad_() for t in (x, w, b)]
    pre = torch.nn.functional.linear(xr.bfloat16().float(), wr.bfloat16().float(), br)
    yr = {None: lambda t: t, "relu": torch.relu, "gelu": lambda t: torch.nn.functional.gelu(t, approximate="tanh")}[act](pre)
    yr.backward(gy)
    torch.testing.assert_close(y.float(), yr, rtol=2e-2, atol=5e-2)
    torch.testing.assert_close(x.grad, xr.grad, rtol=5e-2, atol=2e-2)
    torch.testing.assert_close(w.grad, wr.grad, rtol=5e-2, atol=5e-2)
    torch.testing.assert_close(b.grad, br.grad, rtol=2e-2, atol=5e-2)


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(4096, 1024, 256), (4000, 1000, 200), (8192, 768, 64)])
def test_tma_gemm_cta_pair_kernel(M, N, K, a_mn, b_mn):
    """Shapes with >= 120 128x256 tiles go to the cta_group::2 kernel (256 x 256 tile per CTA pair): all four
    operand-major combinations, ragged M / N / K edges."""
    from tfdl.ops.gemm import gemm_tma

    torch.manual_seed(8)
    a = (torch.randn((K, M) if a_mn else (M, K), device="cuda") * 0.5).bfloat16()
    b = (torch.randn((K, N) if b_mn else (N, K), device="cuda") * 0.5).bfloat16()
    out = gemm_tma(a, b, a_mn=a_mn, b_mn=b_mn, out_dtype=torch.float32)
    torch.testing.assert_close(out, _ref(a, b, a_mn, b_mn), rtol=2e-3, atol=2e-2 * (K ** 0.5) / 8)
    bias = torch.randn(N, device="cuda")
    out16 = gemm_tma(a, b, a_mn=a_mn, b_mn=b_mn, bias=bias, act="gelu")
    ref = torch.nn.functional.gelu(_ref(a, b, a_mn, b_mn) + bias, approximate="tanh")
    torch.testing.assert_close(out16.float(), ref, rtol=2e-2, atol=0.15)


@pytest.mark.parametrize("M,N,K", [(768, 768, 8192), (768, 3072, 4096), (128, 512, 2048)])
def test_tma_gemm_split_k_reduce_add(M, N, K):
    """Few output tiles + long reduction (weight gradients): K is split across CTAs, partial tiles accumulate with
    TMA reduce-add into the zero-filled fp32 output."""
    from tfdl.ops.gemm import auto_ksplit, gemm_tma

    torch.manual_seed(9)
    a = (torch.randn(K, M, device="cuda") * 0.5).bfloat16()
    b = (torch.randn(K, N, device="cuda") * 0.5).bfloat16()
    assert auto_ksplit(768, 768, 8192) > 1 and auto_ksplit(8192, 8192, 8192) == 1
    ref = _ref(a, b, True, True)
    for ks in (None, 3, 7):
        out = gemm_tma(a, b, a_mn=True, b_mn=True, out_dtype=torch.float32, k_split=ks)
        torch.testing.assert_close(out, ref, rtol=2e-3, atol=2e-2 * (K ** 0.5) / 8)
    bias = torch.randn(N, device="cuda")
    out = gemm_tma(a, b, a_mn=True, b_mn=True, out_dtype=torch.float32, bias=bias, k_split=4)
    torch.testing.assert_close(out, ref + bias, rtol=2e-3, atol=2e-2 * (K ** 0.5) / 8)


def test_tied_decoder_unaligned_vocab():
    from tfdl.ops.gemm import tied_decoder

    torch.manual_seed(10)
    V, H, T = 1003, 256, 512
    h = (torch.randn(2, T // 2, H, device="cuda") * 0.5).bfloat16().requires_grad_()
    w = (torch.randn(V, H, device="cuda") * 0.05).requires_grad_()
    b = torch.randn(V, device="cuda").requires_grad_()
    y = tied_decoder(h, w, b)
    assert y.shape == (2, T // 2, V)
    gy = torch.randn(2, T // 2, V, device="cuda") * 0.1
    y.float().backward(gy)
    hr, wr, br = h.detach().float().requires_grad_(), w.detach().bfloat16().float().requires_grad_(), b.detach().clone().requires_grad_()
    yr = torch.nn.functional.linear(hr, wr, br)
    yr.backward(gy.bfloat16().float())
    torch.testing.assert_close(y.float(), yr, rtol=2e-2, atol=5e-2)
    torch.testing.assert_close(h.grad.float(), hr.grad, rtol=5e-2, atol=2e-2)
    torch.testing.assert_close(w.grad, wr.grad, rtol=5e-2, atol=5e-2)
    torch.testing.assert_close(b.grad, br.grad, rtol=2e-2, atol=5e-2)


@pytest.mark.parametrize("H", [256, 768, 1024])
@pytest.mark.parametrize("with_res", [False, True])
def test_native_layernorm_fwd_bwd(H, with_res):
    from tfdl.ops.layernorm import layer_norm, supported

    torch.manual_seed(12)
    T = 1000
    x = (torch.randn(4, T // 4, H, device="cuda") * 2 + 0.5).bfloat16().requires_grad_()
    r = (torch.randn(4, T // 4, H, device="cuda")).bfloat16().requires_grad_() if with_res else None
    base = torch.randn(2 * H + 1, device="cuda")
    g = (base[1:H + 1] * 0.1 + 1).requires_grad_()      # deliberately only 4-byte aligned, like an arena view
    b = base[H + 1:].clone().requires_grad_()
    assert supported(x)
    y = layer_norm(x, g, b, 1e-5, r)
    gy = torch.randn_like(y.float()) * 0.1
    y.float().backward(gy)
    xr = x.detach().float().requires_grad_()
    rr = r.detach().float().requires_grad_() if with_res else None
    gr, br = g.detach().clone().requires_grad_(), b.detach().clone().requires_grad_()
    z = (xr + rr).bfloat16().float() if with_res else xr
    yr = torch.nn.functional.layer_norm(z, (H,), gr, br, 1e-5)
    yr.backward(gy.bfloat16().float())
    torch.testing.assert_close(y.float(), yr, rtol=2e-2, atol=3e-2)
    torch.testing.assert_close(x.grad.float(), xr.grad, rtol=3e-2, atol=3e-3)
    if with_res:
        torch.testing.assert_close(r.grad.float(), rr.grad, rtol=3e-2, atol=3e-3)
    torch.testing.assert_close(g.grad, gr.grad, rtol=2e-2, atol=5e-2)
    torch.testing.assert_close(b.grad, br.grad, rtol=2e-2, atol=5e-2)


@pytest.mark.parametrize("C,relu,with_res", [(64, True, False), (256, True, True), (2048, False, False), (8, True, True)])
def test_native_batchnorm_fwd_bwd(C, relu, with_res):
    from tfdl.ops.batchnorm import batch_norm_infer, batch_norm_train, supported

    torch.manual_seed(13)
    x = (torch.randn(6, 7, 9, C, device="cuda") * 1.5 + 0.3).bfloat16().requires_grad_()
    r = torch.randn(6, 7, 9, C, device="cuda").bfloat16().requires_grad_() if with_res else None
    g = (torch.randn(C, device="cuda") * 0.2 + 1).requires_grad_()
    b = (torch.randn(C, device="cuda") * 0.2).requires_grad_()
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    assert supported(x)
    in_bias = torch.randn(C, device="cuda")  # a skipped conv bias: no effect on y, shifts the running mean
    y = batch_norm_train(x, g, b, rm, rv, 0.01, 1e-3, relu, r, in_bias)
    gy = torch.randn_like(y.float()) * 0.1
    y.float().backward(gy)
    xr = x.detach().float().requires_grad_()
    rr = r.detach().float().requires_grad_() if with_res else None
    gr, br = g.detach().clone().requires_grad_(), b.detach().clone().requires_grad_()
    rm2, rv2 = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    yr = torch.nn.functional.batch_norm(xr.permute(0, 3, 1, 2), rm2, rv2, gr, br, True, 0.01, 1e-3).permute(0, 2, 3, 1)
    if with_res:
        yr = yr + rr
    if relu:
        yr = torch.relu(yr)
    yr.backward(gy.bfloat16().float())
    torch.testing.assert_close(y.float(), yr, rtol=2e-2, atol=3e-2)
    torch.testing.assert_close(rm, rm2 + 0.01 * in_bias, rtol=1e-3, atol=1e-4)
    rm -= 0.01 * in_bias
    torch.testing.assert_close(rv, rv2, rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(x.grad.float(), xr.grad, rtol=5e-2, atol=5e-3)
    if with_res:
        torch.testing.assert_close(r.grad.float(), rr.grad, rtol=5e-2, atol=5e-3)
    torch.testing.assert_close(g.grad, gr.grad, rtol=3e-2, atol=5e-2)
    torch.testing.assert_close(b.grad, br.grad, rtol=3e-2, atol=5e-2)
    yi = batch_norm_infer(x.detach(), g.detach(), b.detach(), rm, rv, 1e-3, relu, r.detach() if with_res else None)
    yir = torch.nn.functional.batch_norm(x.detach().float().permute(0, 3, 1, 2), rm, rv, g.detach(), b.detach(), False, 0.0,
                                         1e-3).permute(0, 2, 3, 1)
    if with_res:
        yir = yir + r.detach().float()
    if relu:
        yir = torch.relu(yir)
    torch.testing.assert_close(yi.float(), yir, rtol=2e-2, atol=3e-2)


def test_conv1x1_on_tcgen05_gemm_matches_cudnn_path(monkeypatch):
    """keras.layers.Conv2D(1x1, stride 1) under mixed_bfloat16 on CUDA is a TMA + tcgen05 GEMM (forward, dgrad, wgrad)."""
    import tfdl
    from tfdl import keras

    x = (torch.randn(8, 14, 14, 256, device="cuda") * 0.5).bfloat16()
    gy = torch.randn(8, 14, 14, 64, device="cuda") * 0.1
    outs = {}
    for mode in ("gemm", "cudnn"):
        monkeypatch.setenv("TFDL_CONV1X1", mode)
        keras.clear_session()
        keras.layers.set_init_seed(21)
        keras.mixed_precision.set_global_policy("mixed_bfloat16")
        try:
            s = tfdl.distribute.MirroredStrategy(devices=["gpu:0"])
            with s.scope():
                m = keras.Sequential([keras.layers.Conv2D(64, 1, activation="relu", input_shape=(14, 14, 256))])
                m.compile(loss="mse", optimizer=keras.optimizers.SGD(0.1))
            eng = m._make_engine()
            ctx = eng._ctx(0, True)
            xin = x.clone().requires_grad_()
            y = m._forward(xin, ctx)
            y.float().backward(gy)
            outs[mode] = (y.float().detach(), xin.grad.float(), m.arenas[0].grads.clone())
        finally:
            keras.mixed_precision.set_global_policy("float32")
    for a, b in zip(outs["gemm"], outs["cudnn"]):
        torch.testing.assert_close(a, b, rtol=3e-2, atol=3e-2)
    assert float(outs["gemm"][2].abs().sum()) > 0


@pytest.mark.parametrize("T,N", [(8192, 768), (1000, 3072), (77, 8), (4096, 30528)])
def test_colsum_bias_gradient_kernel(T, N):
    from tfdl.ops.gemm import colsum

    torch.manual_seed(14)
    x = (torch.randn(T, N, device="cuda") * 0.5).bfloat16()
    torch.testing.assert_close(colsum(x), x.float().sum(0), rtol=1e-3, atol=1e-2)
    if N >= 16:  # a column slice of a wider buffer (row pitch > N)
        torch.testing.assert_close(colsum(x[:, :N - 8]), x[:, :N - 8].float().sum(0), rtol=1e-3, atol=1e-2)


@pytest.mark.parametrize("B,heads", [(3, 12), (2, 2)])
def test_fused_attention_fwd_bwd(B, heads):
    from tfdl.ops.attention import attention, supported

    torch.manual_seed(15)
    H = heads * 64
    qkv = (torch.randn(B, 128, 3 * H, device="cuda") * 0.8).bfloat16().requires_grad_()
    assert supported(qkv, heads)
    y = attention(qkv, heads)
    gy = torch.randn(B, 128, H, device="cuda") * 0.1
    y.float().backward(gy)
    ref_in = qkv.detach().float().requires_grad_()
    q, k, v = (t.view(B, 128, heads, 64).transpose(1, 2) for t in ref_in.split(H, -1))
    yr = torch.nn.functional.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(B, 128, H)
    yr.backward(gy.bfloat16().float())
    torch.testing.assert_close(y.float(), yr, rtol=3e-2, atol=2e-2)
    torch.testing.assert_close(qkv.grad.float(), ref_in.grad, rtol=5e-2, atol=5e-3)


@pytest.mark.parametrize("T,V,pad", [(300, 1003, True), (64, 512, False), (128, 30522, True)])
def test_fused_softmax_xent_fwd_bwd(T, V, pad):
    from tfdl.ops.xent import sparse_softmax_xent, supported

    torch.manual_seed(16)
    Vp = (V + 7) // 8 * 8 if pad else V
    buf = torch.full((T, Vp), -30000.0, device="cuda").bfloat16()
    buf[:, :V] = (torch.randn(T, V, device="cuda") * 3).bfloat16()
    logits = buf[:, :V].view(2, T // 2, V).detach().requires_grad_()
    y = torch.randint(0, V, (2, T // 2), device="cuda")
    y[0, :5] = -100
    assert supported(logits)
    loss = sparse_softmax_xent(logits, y, -100)
    g = torch.rand(T, device="cuda")
    loss.backward(g)
    ref_in = buf[:, :V].float().detach().requires_grad_()
    lr = torch.nn.functional.cross_entropy(ref_in, y.reshape(-1), reduction="none", ignore_index=-100)
    lr.backward(g)
    torch.testing.assert_close(loss, lr, rtol=1e-3, atol=1e-3)
    torch.testing.assert_close(logits.grad.float().reshape(T, V), ref_in.grad, rtol=2e-2, atol=2e-3)
