"""Conv2D on tcgen05 (ref: tf.keras.layers.Conv2D, /root/reference/tf_dist_example.py:41,43).

Forward, input gradient and weight gradient run on the implicit-GEMM kernel of csrc/kernels/conv_igemm.cu (TMA im2col
loads, no im2col buffer). Activations are NHWC bf16, weights are the OHWI storage of the layer's kernel variable, the
weight gradient is fp32. Channel counts must be multiples of 8; layers that are not (the 1- or 3-channel first layer of
an image model) go through `conv2d_small_cin`, which builds the K = R*S*Cin patch matrix once and runs the GEMM kernels
on it.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib
from .gemm import colsum, gemm_tma

_ACT = {None: 0, "linear": 0, "relu": 1}


def out_size(size: int, k: int, s: int, pad_lo: int, pad_hi: int) -> int:
    return (size + pad_lo + pad_hi - k) // s + 1


def supported(x: torch.Tensor, cout: int) -> bool:
    return bool(x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 4 and x.shape[-1] % 8 == 0 and cout % 8 == 0)


def conv2d_fwd(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], strides: Tuple[int, int],
               pads: Tuple[int, int, int, int], act: Optional[str] = None) -> torch.Tensor:
    """x [N,H,W,C] bf16, w [K,R,S,C] bf16, bias fp32 [K] -> y [N,P,Q,K] bf16. pads = (top, left, bottom, right)."""
    N, H, W, C = x.shape
    K, R, S, _ = w.shape
    pt, pl, pb, pr = pads
    P, Q = out_size(H, R, strides[0], pt, pb), out_size(W, S, strides[1], pl, pr)
    y = torch.empty((N, P, Q, K), device=x.device, dtype=torch.bfloat16)
    _lib.check(_lib.kernels().tfdl_conv2d_fwd(_lib.ptr(x), _lib.ptr(w), _lib.ptr(bias), _lib.ptr(y), N, H, W, C, K, R, S,
                                              strides[0], strides[1], pt, pl, pb, pr, _ACT[act], _lib.cur_stream()),
               "conv2d_fwd")
    return y


def conv2d_dgrad(dy: torch.Tensor, w: torch.Tensor, x_shape: Tuple[int, ...], strides: Tuple[int, int],
                 pads: Tuple[int, int, int, int]) -> torch.Tensor:
    N, H, W, C = x_shape
    K, R, S, _ = w.shape
    pt, pl, pb, pr = pads
    lib = _lib.kernels()
    if strides == (1, 1):
        dx = torch.empty(x_shape, device=dy.device, dtype=torch.bfloat16)
        _lib.check(lib.tfdl_conv2d_dgrad(_lib.ptr(dy), _lib.ptr(w), _lib.ptr(dx), N, H, W, C, K, R, S, pt, pl, pb, pr,
                                         _lib.cur_stream()), "conv2d_dgrad")
        return dx
    if (R, S) == (1, 1) and pads == (0, 0, 0, 0):
        # strided 1x1: a GEMM on the sub-sampled grid, then a zero-filling scatter back to full resolution
        P, Q = dy.shape[1], dy.shape[2]
        sub = torch.empty((N, P, Q, C), device=dy.device, dtype=torch.bfloat16)
        _lib.check(lib.tfdl_conv2d_dgrad(_lib.ptr(dy), _lib.ptr(w), _lib.ptr(sub), N, P, Q, C, K, 1, 1, 0, 0, 0, 0,
                                         _lib.cur_stream()), "conv2d_dgrad")
        dx = torch.empty(x_shape, device=dy.device, dtype=torch.bfloat16)
        _lib.check(lib.tfdl_scatter_strided_nhwc(_lib.ptr(sub), _lib.ptr(dx), N, H, W, C, P, Q, strides[0], strides[1],
                                                 _lib.cur_stream()), "scatter_strided_nhwc")
        return dx
    # strided RxS input gradient (not used by the reference CNN or Keras ResNet-50): library fallback
    wt = w.permute(0, 3, 1, 2).float()
    dx = torch.nn.grad.conv2d_input((N, C, H + pt + pb, W + pl + pr), wt, dy.permute(0, 3, 1, 2).float(), stride=strides)
    return dx[:, :, pt:pt + H, pl:pl + W].permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)


def conv2d_wgrad(x: torch.Tensor, dy: torch.Tensor, w_shape: Tuple[int, ...], strides: Tuple[int, int],
                 pads: Tuple[int, int, int, int], out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fp32 dw [K,R,S,C]; accumulated into `out` when given (it must then hold the running gradient)."""
    N, H, W, C = x.shape
    K, R, S, _ = w_shape
    pt, pl, pb, pr = pads
    dw = out if out is not None else torch.zeros(w_shape, device=x.device, dtype=torch.float32)
    _lib.check(_lib.kernels().tfdl_conv2d_wgrad(_lib.ptr(x), _lib.ptr(dy), _lib.ptr(dw), N, H, W, C, K, R, S, strides[0],
                                                strides[1], pt, pl, pb, pr, _lib.cur_stream()), "conv2d_wgrad")
    return dw


class _Conv2dFn(torch.autograd.Function):
    """w_lp / sink as in ops/gemm.py:_LinearFn: a ready bf16 copy of the kernel, and a gradient-arena view that the
    weight-gradient kernel accumulates into directly (TMA reduce-add) instead of returning a tensor to autograd."""

    @staticmethod
    def forward(ctx, x, w, bias, strides, pads, act, w_lp=None, sink=None):
        xb = x if x.dtype == torch.bfloat16 else x.bfloat16()
        xb = xb.contiguous()
        wb = w_lp if w_lp is not None else (w if w.dtype == torch.bfloat16 else w.bfloat16()).contiguous()
        y = conv2d_fwd(xb, wb, None if bias is None else bias.float(), strides, pads, act)
        ctx.save_for_backward(xb, wb, y if act == "relu" else None)
        ctx.meta = (strides, pads, act, bias is not None, w.dtype, x.requires_grad, sink)
        return y

    @staticmethod
    def backward(ctx, dy):
        xb, wb, y = ctx.saved_tensors
        strides, pads, act, has_bias, w_dtype, need_dx, sink = ctx.meta
        dy = dy if dy.dtype == torch.bfloat16 else dy.bfloat16()
        if act == "relu":
            dy = dy * (y > 0)
        dy = dy.contiguous()
        dx = conv2d_dgrad(dy, wb, tuple(xb.shape), strides, pads) if need_dx else None
        db = colsum(dy.reshape(-1, dy.shape[-1])) if has_bias else None
        if sink is not None:
            view, done = sink
            conv2d_wgrad(xb, dy, tuple(wb.shape), strides, pads, out=view)
            done()
            return dx, None, db, None, None, None, None, None
        dw = conv2d_wgrad(xb, dy, tuple(wb.shape), strides, pads)
        return dx, dw.to(w_dtype), db, None, None, None, None, None


def conv2d(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], strides: Tuple[int, int],
           pads: Tuple[int, int, int, int], act: Optional[str] = None, w_lp: Optional[torch.Tensor] = None,
           sink=None) -> torch.Tensor:
    """Conv2D with autograd on the implicit-GEMM kernels. x NHWC (cast to bf16), w OHWI fp32/bf16."""
    return _Conv2dFn.apply(x, w, bias, tuple(strides), tuple(pads), act, w_lp, sink)


# ---- first layers: C not a multiple of 8 (1 or 3 input channels) -------------------------------------------------------
def _patches(x: torch.Tensor, R: int, S: int, strides: Tuple[int, int], pads: Tuple[int, int, int, int], kp: int) -> torch.Tensor:
    """[N*P*Q, kp] bf16 patch matrix (columns (r, s, c), zero-padded to kp) from an fp32 or bf16 NHWC input: one native
    gather kernel with the cast fused in (csrc/kernels/pool.cu:im2col_small_kernel)."""
    N, H, W, C = x.shape
    pt, pl, pb, pr = pads
    P, Q = out_size(H, R, strides[0], pt, pb), out_size(W, S, strides[1], pl, pr)
    if x.dtype not in (torch.float32, torch.bfloat16):
        x = x.float()
    x = x.contiguous()
    out = torch.empty((N * P * Q, kp), device=x.device, dtype=torch.bfloat16)
    _lib.check(_lib.kernels().tfdl_im2col_small(_lib.ptr(x), int(x.dtype == torch.float32), _lib.ptr(out), N, H, W, C, R, S,
                                                strides[0], strides[1], pt, pl, P, Q, kp, _lib.cur_stream()),
               "im2col_small")
    return out


class _ConvSmallCinFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, strides, pads, act):
        K, R, S, C = w.shape
        kp = (R * S * C + 7) // 8 * 8
        cols = _patches(x, R, S, strides, pads, kp)
        wm = torch.zeros((K, kp), device=w.device, dtype=torch.bfloat16)
        wm[:, :R * S * C] = w.reshape(K, R * S * C)
        y = gemm_tma(cols, wm, bias=None if bias is None else bias.float(), act=act)
        N, H, W, _ = x.shape
        P, Q = out_size(H, R, strides[0], pads[0], pads[2]), out_size(W, S, strides[1], pads[1], pads[3])
        ctx.save_for_backward(cols, y if act == "relu" else None)
        ctx.meta = (tuple(w.shape), act, bias is not None, w.dtype)
        return y.view(N, P, Q, K)

    @staticmethod
    def backward(ctx, dy):
        cols, y = ctx.saved_tensors
        w_shape, act, has_bias, w_dtype = ctx.meta
        K, R, S, C = w_shape
        dy2 = (dy if dy.dtype == torch.bfloat16 else dy.bfloat16()).reshape(-1, K)
        if act == "relu":
            dy2 = dy2 * (y > 0)
        dy2 = dy2.contiguous()
        dwm = gemm_tma(dy2, cols, a_mn=True, b_mn=True, out_dtype=torch.float32)  # [K, kp] = dy^T cols
        dw = dwm[:, :R * S * C].reshape(w_shape)
        db = colsum(dy2) if has_bias else None
        return None, dw.to(w_dtype), db, None, None, None


def conv2d_small_cin(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], strides: Tuple[int, int],
                     pads: Tuple[int, int, int, int], act: Optional[str] = None) -> torch.Tensor:
    """Conv2D for an input with few channels (no input gradient: it is the model's first layer)."""
    return _ConvSmallCinFn.apply(x, w, bias, tuple(strides), tuple(pads), act)
