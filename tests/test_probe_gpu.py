"""Hardware facts the conv kernels rely on (csrc/kernels/probe.cu), checked against plain fp32 matmuls:
a swizzled operand can be read at ANY row offset by moving the descriptor start address (base-offset
field left 0), for K-major and MN-major operands alike; LBO = 0 replicates the first 64 MN rows into
MMA rows 64..127; a 4-D tensor map with permuted strides emits the no-swizzle core-matrix layout."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from tfdl.ops import _lib

    return _lib.kernels(), _lib.ptr


@pytest.mark.parametrize("kc", [64, 32])
@pytest.mark.parametrize("shift", [0, 1, 3, 8, 13, 100])
def test_kmajor_swizzled_operand_read_at_any_row_shift(kc, shift):
    L, p = _lib()
    torch.manual_seed(0)
    A = torch.randn(256, kc, device="cuda").bfloat16()
    B = torch.randn(64, kc, device="cuda").bfloat16()
    D = torch.zeros(128, 64, device="cuda")
    assert L.tfdl_probe_shifted_umma(p(A), p(B), p(D), 256, 64, kc, shift, 0, None) == 0
    torch.cuda.synchronize()
    torch.testing.assert_close(D, A[shift:shift + 128].float() @ B.float().t(), rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("lbo_mode", [0, 1])
@pytest.mark.parametrize("shift", [0, 1, 13, 28])
def test_mnmajor_swizzled_operands_with_k_shift_and_lbo0_replication(lbo_mode, shift):
    L, p = _lib()
    torch.manual_seed(1)
    krows = 144
    A = torch.randn(krows, 128 if lbo_mode == 0 else 64, device="cuda").bfloat16()
    B = torch.randn(krows + 64, 32, device="cuda").bfloat16()
    D = torch.zeros(128, 32, device="cuda")
    assert L.tfdl_probe_mn_umma(p(A), p(B), p(D), krows, shift, lbo_mode, None) == 0
    torch.cuda.synchronize()
    Af = A.float() if lbo_mode == 0 else torch.cat([A.float(), A.float()], 1)
    torch.testing.assert_close(D, Af.t() @ B[shift:shift + krows].float(), rtol=1e-3, atol=1e-2)


def test_tma_permuted_strides_write_core_matrix_layout():
    L, p = _lib()
    rows, K = 64, 288
    W = torch.randn(rows, K, device="cuda").bfloat16()
    out = torch.zeros(rows * K, dtype=torch.bfloat16, device="cuda")
    assert L.tfdl_probe_tma_coremajor(p(W), p(out), rows, K, None) == 0
    torch.cuda.synchronize()
    assert torch.equal(out, W.view(rows // 8, 8, K // 8, 8).permute(0, 2, 1, 3).reshape(-1))
