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


def gemm_tma(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
             bias: torch.Tensor | None = None, act: str | None = None, out_dtype: torch.dtype = torch.bfloat16,
             out: torch.Tensor | None = None) -> torch.Tensor:
    """D[M,N] = act(A[M,K] @ B[N,K]^T + bias) on the persistent TMA + tcgen05 kernel.

    ``a_mn`` / ``b_mn``: the operand is stored transposed (``a`` is [K, M] / ``b`` is [K, N]) and is read as
    an MN-major tensor-core operand - no transpose copy is made."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.is_cuda and b.is_cuda
    a = a.contiguous()
    b = b.contiguous()
    K, M = a.shape if a_mn else a.shape[::-1]
    Kb, N = b.shape if b_mn else b.shape[::-1]
    assert K == Kb, (a.shape, b.shape)
    if out is None:
        out = torch.empty((M, N), device=a.device, dtype=out_dtype)
    d32 = out if out.dtype == torch.float32 else None
    d16 = out if out.dtype == torch.bfloat16 else None
    rc = _lib.kernels().tfdl_gemm_bf16_tma_ex(
        _lib.ptr(a), _lib.ptr(b), _lib.ptr(d32), _lib.ptr(d16), _lib.ptr(bias),
        M, N, K, a.stride(0), b.stride(0), out.stride(0), int(a_mn), int(b_mn), _ACT[act], _lib.cur_stream())
    _lib.check(rc, "tfdl_gemm_bf16_tma_ex")
    return out


class _LinearFn(torch.autograd.Function):
    """y = act(x W^T + b) with forward, input-gradient and weight-gradient GEMMs all on tcgen05."""

    @staticmethod
    def forward(ctx, x, w, bias, act):
        xb = x if x.dtype == torch.bfloat16 else x.bfloat16()
        wb = w if w.dtype == torch.bfloat16 else w.bfloat16()
        x2 = xb.reshape(-1, xb.shape[-1])
        pre_act = act in ("relu", "gelu")
        y = gemm_tma(x2, wb, bias=None if bias is None else bias.float(), act=None if pre_act else act)
        ctx.act = act
        if pre_act:
            ctx.save_for_backward(x2, wb, y)
            out = torch.relu(y) if act == "relu" else torch.nn.functional.gelu(y, approximate="tanh")
        else:
            ctx.save_for_backward(x2, wb, None)
            out = y
        ctx.has_bias, ctx.x_shape, ctx.w_dtype = bias is not None, x.shape, w.dtype
        return out.reshape(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, wb, pre = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if ctx.act == "relu":
            dy2 = dy2 * (pre > 0)
        elif ctx.act == "gelu":
            dy2 = torch.ops.aten.gelu_backward(dy2.to(pre.dtype), pre, approximate="tanh")
        dyb = dy2.bfloat16().contiguous()
        dx = gemm_tma(dyb, wb, b_mn=True)                                      # [T,out] x W[out,in] (MN-major)
        dw = gemm_tma(dyb, x2, a_mn=True, b_mn=True, out_dtype=torch.float32)  # dy^T x, both MN-major
        db = dyb.sum(0, dtype=torch.float32) if ctx.has_bias else None
        return dx.reshape(ctx.x_shape), dw.to(ctx.w_dtype), db, None


def linear(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None, act: str | None = None) -> torch.Tensor:
    """Dense layer on the tensor core with autograd (bf16 operands, fp32 accumulation, fp32 weight gradient).
    Needs every GEMM dimension (tokens, in, out) to be a multiple of 8."""
    return _LinearFn.apply(x, w, bias, act)
