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
