"""tfdl_dense_fwd_splitk_static: the Dense forward for hosts that cannot allocate device memory
(a process whose GPU is entirely owned by its own pool, a loop that cannot afford cudaMalloc's
device-wide synchronisation, a plain C caller). Its split-K accumulator is an array inside the
library, so the test drives it the way such a host would - straight through the C ABI with raw
pointers - and checks it against the caller-scratch entry point on two calls in a row, since every
call reuses the same built-in array. Collected last on purpose (file name sorts after all others)."""
import ctypes as C
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                   "tensorflow-distributed-learning_b200", "_native", "libtfdl_kernels.so")
STATIC_FLOATS = 65536


def _lib():
    lib = C.CDLL(LIB)
    lib.tfdl_last_error.restype = C.c_char_p
    vp = C.c_void_p
    lib.tfdl_dense_fwd_splitk.argtypes = [vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp]
    lib.tfdl_dense_fwd_splitk_static.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp]
    return lib


def test_static_scratch_entry_point_matches_caller_scratch_twice_in_a_row():
    lib = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for call, (B, N, K, relu, seed) in enumerate([(64, 128, 1600, 1, 0), (32, 256, 512, 0, 1)]):
        torch.manual_seed(seed)
        x = (torch.randn(B, K, device=dev) * 0.5).bfloat16()
        w = (torch.randn(N, K, device=dev) * 0.1).bfloat16()
        bias = torch.randn(N, device=dev)
        out_ref = torch.empty(B, N, dtype=torch.bfloat16, device=dev)
        out_new = torch.empty(B, N, dtype=torch.bfloat16, device=dev)
        scratch = torch.zeros(B * N, device=dev)
        p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
        rc = lib.tfdl_dense_fwd_splitk(p(x), p(w), p(bias), p(out_ref), p(scratch), B, N, K, relu, 0, stream)
        assert rc == 0, lib.tfdl_last_error()
        if call == 1:
            # leave the whole built-in array dirty (NaN bit pattern), as an aborted earlier call could: the entry point
            # must clear the part it uses itself
            lib.tfdl_static_scratch_fill.argtypes = [C.c_float, C.c_int, C.c_void_p]
            nan = C.c_float.from_buffer_copy((0xFFFFFFFF).to_bytes(4, "little"))
            assert lib.tfdl_static_scratch_fill(nan, STATIC_FLOATS, stream) == 0, lib.tfdl_last_error()
        rc = lib.tfdl_dense_fwd_splitk_static(p(x), p(w), p(bias), p(out_new), B, N, K, relu, 0, stream)
        assert rc == 0, lib.tfdl_last_error()
        torch.cuda.synchronize()
        # same tolerance as the other GEMM tests; the two differ only in RED summation order
        torch.testing.assert_close(out_new.float(), out_ref.float(), rtol=2e-2, atol=0.1)
        ref = x.float() @ w.float().t() + bias
        ref = torch.relu(ref) if relu else ref
        torch.testing.assert_close(out_new.float(), ref, rtol=2e-2, atol=0.1)


def test_static_scratch_rejects_oversized_problems_with_a_message():
    lib = _lib()
    dev = torch.device("cuda", 0)
    B, N, K = 256, 512, 64  # B*N = 131072 > 65536
    assert B * N > STATIC_FLOATS
    x = torch.zeros(B, K, dtype=torch.bfloat16, device=dev)
    w = torch.zeros(N, K, dtype=torch.bfloat16, device=dev)
    out = torch.full((B, N), 7.0, dtype=torch.bfloat16, device=dev)
    rc = lib.tfdl_dense_fwd_splitk_static(C.c_void_p(x.data_ptr()), C.c_void_p(w.data_ptr()), None,
                                          C.c_void_p(out.data_ptr()), B, N, K, 0, 0, None)
    assert rc != 0
    msg = lib.tfdl_last_error().decode()
    assert "exceeds the built-in scratch" in msg and "65536" in msg
    torch.cuda.synchronize()
    assert torch.all(out.float() == 7.0)  # nothing was launched
