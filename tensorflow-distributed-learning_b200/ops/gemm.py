"""Dense-layer GEMMs on tcgen05 (ref: tf.keras.layers.Dense, /root/reference/tf_dist_example.py:46-47)."""
from __future__ import annotations

import ctypes

import torch

from . import _lib

_ACT = {None: 0, "none": 0, "linear": 0, "relu": 1, "gelu": 2}


def gemm_gather(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
                bias: torch.Tensor | None = None, relu: bool = False, k_split: int = 1,
                out_dtype: torch.dtype = torch.float32, out: torch.Tensor | None = None) -> torch.Tensor:
    """D[M,N] = A[M,K] @ B[N,K]^T on tcgen05 with cp.async-gathered operands.

    ``a_mn`` / ``b_mn``: the operand is stored transposed, i.e. ``a`` is [K, M] / ``b`` is [K, N].
    ``k_split`` > 1 accumulates partial products atomically into a zeroed fp32 output.
    """
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.is_cuda and b.is_cuda
    a = a.contiguous()
    b = b.contiguous()
    if a_mn:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_mn:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    assert K == Kb, (a.shape, b.shape)
    if out is None:
        out = (torch.zeros if k_split > 1 else torch.empty)((M, N), device=a.device, dtype=out_dtype)
    d32 = out if out.dtype == torch.float32 else None
    d16 = out if out.dtype == torch.bfloat16 else None
    rc = _lib.kernels().tfdl_gemm_bf16_gather(
        _lib.ptr(a), _lib.ptr(b), _lib.ptr(d32), _lib.ptr(d16), _lib.ptr(bias),
        M, N, K, a.stride(0), b.stride(0), out.stride(0), int(a_mn), int(b_mn), int(relu), int(k_split),
        _lib.cur_stream())
    _lib.check(rc, "tfdl_gemm_bf16_gather")
    return out


def gemm_tma(a: torch.Tensor, b: torch.Tensor, *, bias: torch.Tensor | None = None, act: str | None = None,
             out_dtype: torch.dtype = torch.bfloat16, out: torch.Tensor | None = None) -> torch.Tensor:
    """D[M,N] = act(A[M,K] @ B[N,K]^T + bias) on the persistent TMA + tcgen05 kernel."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.is_cuda and b.is_cuda
    a = a.contiguous()
    b = b.contiguous()
    M, K = a.shape
    N, Kb = b.shape
    assert K == Kb
    if out is None:
        out = torch.empty((M, N), device=a.device, dtype=out_dtype)
    d32 = out if out.dtype == torch.float32 else None
    d16 = out if out.dtype == torch.bfloat16 else None
    rc = _lib.kernels().tfdl_gemm_bf16_tma(
        _lib.ptr(a), _lib.ptr(b), _lib.ptr(d32), _lib.ptr(d16), _lib.ptr(bias),
        M, N, K, a.stride(0), b.stride(0), out.stride(0), _ACT[act], _lib.cur_stream())
    _lib.check(rc, "tfdl_gemm_bf16_tma")
    return out
