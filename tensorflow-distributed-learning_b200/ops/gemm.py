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


def auto_ksplit(M: int, N: int, K: int) -> int:
    """Split-K factor for problems whose output has too few tiles to fill 148 SMs but whose reduction is long
    (weight gradients: M, N = layer widths, K = tokens). 1 = no split."""
    tiles = -(-M // 128) * -(-N // 256)
    kb = -(-K // 64)
    if tiles >= 60 or kb < 16:
        return 1
    return max(1, min(148 // max(tiles, 1), kb // 8, 16))


def gemm_tma(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
             bias: torch.Tensor | None = None, act: str | None = None, out_dtype: torch.dtype = torch.bfloat16,
             out: torch.Tensor | None = None, k_split: int | None = None) -> torch.Tensor:
    """D[M,N] = act(A[M,K] @ B[N,K]^T + bias) on the persistent TMA + tcgen05 kernels (1-CTA or CTA-pair).

    ``a_mn`` / ``b_mn``: the operand is stored transposed (``a`` is [K, M] / ``b`` is [K, N]) and is read as
    an MN-major tensor-core operand - no transpose copy is made. ``out`` may be a column slice of a wider buffer
    (its row stride is the output pitch). ``k_split``: None = automatic for fp32 outputs without activation
    (the output is then zero-filled here and the splits accumulate with TMA reduce-add)."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.is_cuda and b.is_cuda
    if a.stride(-1) != 1 or a.stride(0) % 8:
        a = a.contiguous()
    if b.stride(-1) != 1 or b.stride(0) % 8:
        b = b.contiguous()
    K, M = a.shape if a_mn else a.shape[::-1]
    Kb, N = b.shape if b_mn else b.shape[::-1]
    assert K == Kb, (a.shape, b.shape)
    if out is None:
        out = torch.empty((M, N), device=a.device, dtype=out_dtype)
    assert out.shape == (M, N) and out.stride(1) == 1
    d32 = out if out.dtype == torch.float32 else None
    d16 = out if out.dtype == torch.bfloat16 else None
    ks = 1
    if d32 is not None and act in (None, "none", "linear") and out.stride(0) % 4 == 0:
        ks = auto_ksplit(M, N, K) if k_split is None else int(k_split)
    if ks > 1:
        out.zero_()
    rc = _lib.kernels().tfdl_gemm_bf16_tma_splitk(
        _lib.ptr(a), _lib.ptr(b), _lib.ptr(d32), _lib.ptr(d16), _lib.ptr(bias),
        M, N, K, a.stride(0), b.stride(0), out.stride(0), int(a_mn), int(b_mn), _ACT[act], ks, _lib.cur_stream())
    _lib.check(rc, "tfdl_gemm_bf16_tma_splitk")
    return out


def colsum(x: torch.Tensor) -> torch.Tensor:
    """fp32 column sums of a bf16 [T, N] matrix (a Dense layer's bias gradient) on the native streaming kernel;
    falls back to torch for shapes it does not take."""
    T, N = x.shape
    if x.is_cuda and x.dtype == torch.bfloat16 and N % 8 == 0 and N <= 8192 and x.stride(1) == 1 and x.stride(0) % 8 == 0 \
            and x.data_ptr() % 16 == 0:
        out = torch.zeros(N, device=x.device, dtype=torch.float32)
        _lib.check(_lib.kernels().tfdl_colsum_bf16(_lib.ptr(x), _lib.ptr(out), ctypes.c_longlong(T), N, x.stride(0),
                                                   _lib.cur_stream()), "tfdl_colsum_bf16")
        return out
    return x.sum(0, dtype=torch.float32)


class _LinearFn(torch.autograd.Function):
    """y = act(x W^T + b) with forward, input-gradient and weight-gradient GEMMs all on tcgen05.

    w_lp: optional bf16 copy of W that the caller already has (the shadow arena the fused all-reduce keeps current), so no
    cast runs here. sink: optional (fp32 gradient view in the arena, callback): the weight gradient is then written
    straight into the arena by the GEMM epilogue and autograd gets None - no separate dW tensor, no accumulate pass."""

    @staticmethod
    def forward(ctx, x, w, bias, act, w_lp=None, sink=None):
        xb = x if x.dtype == torch.bfloat16 else x.bfloat16()
        wb = w_lp if w_lp is not None else (w if w.dtype == torch.bfloat16 else w.bfloat16())
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
        ctx.has_bias, ctx.x_shape, ctx.w_dtype, ctx.sink = bias is not None, x.shape, w.dtype, sink
        ctx.need_dx = x.requires_grad
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
        dx = gemm_tma(dyb, wb, b_mn=True).reshape(ctx.x_shape) if ctx.need_dx else None  # [T,out] x W[out,in] (MN-major)
        db = colsum(dyb) if ctx.has_bias else None
        if ctx.sink is not None:
            view, done = ctx.sink
            gemm_tma(dyb, x2, a_mn=True, b_mn=True, out=view.view(wb.shape[0], -1))  # dy^T x into the gradient arena
            done()
            return dx, None, db, None, None, None
        dw = gemm_tma(dyb, x2, a_mn=True, b_mn=True, out_dtype=torch.float32)  # dy^T x, both MN-major
        return dx, dw.to(ctx.w_dtype), db, None, None, None


def linear(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None, act: str | None = None,
           w_lp: torch.Tensor | None = None, sink=None) -> torch.Tensor:
    """Dense layer on the tensor core with autograd (bf16 operands, fp32 accumulation, fp32 weight gradient).
    Needs every GEMM dimension (tokens, in, out) to be a multiple of 8."""
    return _LinearFn.apply(x, w, bias, act, w_lp, sink)


class _TiedDecoderFn(torch.autograd.Function):
    """logits = h W^T + b for an output width that is not a multiple of 8 (BERT's vocabulary: 30522).

    Row pitches that are not 16-byte multiples push library GEMMs onto slow legacy kernels and rule out TMA. Here the
    weight, the bias and the logits live in buffers padded to the next multiple of 8 columns: pad rows of W are zero
    and pad bias entries are -30000, so the pad logits vanish from any softmax; the caller sees the [T, V] column
    slice. Forward, dgrad and wgrad run on the tcgen05 CTA-pair kernel."""

    @staticmethod
    def forward(ctx, h, w, bias):
        V, H = w.shape
        Vp = (V + 7) // 8 * 8
        h2 = (h if h.dtype == torch.bfloat16 else h.bfloat16()).reshape(-1, H)
        wp = torch.zeros((Vp, H), device=w.device, dtype=torch.bfloat16)
        wp[:V] = w
        bp = torch.full((Vp,), -30000.0, device=w.device, dtype=torch.float32)
        if bias is not None:
            bp[:V] = bias
        else:
            bp[:V] = 0
        out = gemm_tma(h2, wp, bias=bp)
        ctx.save_for_backward(h2, wp)
        ctx.meta = (V, Vp, h.shape, w.dtype, bias is not None)
        return out[:, :V].reshape(*h.shape[:-1], V) if Vp == V else out[:, :V].unflatten(0, h.shape[:-1])

    @staticmethod
    def backward(ctx, dy):
        h2, wp = ctx.saved_tensors
        V, Vp, h_shape, w_dtype, has_bias = ctx.meta
        T = h2.shape[0]
        d2 = dy.view(T, V) if dy.dim() == 2 or dy.stride(-1) == 1 and dy.stride(-2) == Vp else None
        if (Vp != V and d2 is not None and d2.dtype == torch.bfloat16 and d2.stride() == (Vp, 1)
                and d2.data_ptr() % 16 == 0 and getattr(d2, "_base", None) is not None
                and d2._base.numel() == T * Vp):
            # the gradient already IS a column slice of a padded bf16 buffer with zero pad columns (ops/xent.py): use it
            dyp = torch.as_strided(d2, (T, Vp), (Vp, 1))
        else:
            dyp = torch.zeros((T, Vp), device=dy.device, dtype=torch.bfloat16)
            dyp[:, :V] = dy.reshape(-1, V)
        dh = gemm_tma(dyp, wp, b_mn=True)
        dw = gemm_tma(dyp, h2, a_mn=True, b_mn=True, out_dtype=torch.float32)[:V]
        db = colsum(dyp)[:V] if has_bias else None
        return dh.reshape(h_shape), dw.to(w_dtype), db


def tied_decoder(h: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None) -> torch.Tensor:
    """h [..., H] x w [V, H]^T + bias [V] -> [..., V] (bf16) on tcgen05; V need not be a multiple of 8."""
    return _TiedDecoderFn.apply(h, w, bias)
