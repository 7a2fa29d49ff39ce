"""Max pooling and global average pooling for NHWC bf16 activations on the native streaming kernels
(csrc/kernels/pool.cu; ref: tf.keras.layers.MaxPooling2D, /root/reference/tf_dist_example.py:42,44)."""
from __future__ import annotations

from typing import Tuple

import torch

from . import _lib


def supported(x: torch.Tensor) -> bool:
    return bool(x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 4 and x.shape[-1] % 8 == 0)


class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, window, strides, pads, out_hw):
        x = x.contiguous()
        N, H, W, C = x.shape
        P, Q = out_hw
        y = torch.empty((N, P, Q, C), device=x.device, dtype=torch.bfloat16)
        idx = torch.empty((N, P, Q, C), device=x.device, dtype=torch.uint8)
        _lib.check(_lib.kernels().tfdl_maxpool_fwd(_lib.ptr(x), _lib.ptr(y), _lib.ptr(idx), N, H, W, C, window[0], window[1],
                                                   strides[0], strides[1], pads[0], pads[1], P, Q, _lib.cur_stream()),
                   "maxpool_fwd")
        ctx.save_for_backward(idx)
        ctx.meta = (tuple(x.shape), window, strides, pads, out_hw)
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        (N, H, W, C), window, strides, pads, (P, Q) = ctx.meta
        dy = (dy if dy.dtype == torch.bfloat16 else dy.bfloat16()).contiguous()
        dx = torch.empty((N, H, W, C), device=dy.device, dtype=torch.bfloat16)
        _lib.check(_lib.kernels().tfdl_maxpool_bwd(_lib.ptr(dy), _lib.ptr(idx), _lib.ptr(dx), N, H, W, C, window[0], window[1],
                                                   strides[0], strides[1], pads[0], pads[1], P, Q, _lib.cur_stream()),
                   "maxpool_bwd")
        return dx, None, None, None, None


def max_pool(x: torch.Tensor, window: Tuple[int, int], strides: Tuple[int, int], pads: Tuple[int, int],
             out_hw: Tuple[int, int]) -> torch.Tensor:
    """x [N,H,W,C] bf16 -> [N,P,Q,C]; pads = (top, left): padded positions never win (Keras 'same')."""
    return _MaxPoolFn.apply(x, tuple(window), tuple(strides), tuple(pads), tuple(out_hw))


class _GlobalAvgPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = x.contiguous()
        N, H, W, C = x.shape
        y = torch.empty((N, C), device=x.device, dtype=torch.bfloat16)
        _lib.check(_lib.kernels().tfdl_global_avgpool_fwd(_lib.ptr(x), _lib.ptr(y), N, H * W, C, _lib.cur_stream()),
                   "global_avgpool_fwd")
        ctx.shape = (N, H, W, C)
        return y

    @staticmethod
    def backward(ctx, dy):
        N, H, W, C = ctx.shape
        dy = (dy if dy.dtype == torch.bfloat16 else dy.bfloat16()).contiguous()
        dx = torch.empty((N, H, W, C), device=dy.device, dtype=torch.bfloat16)
        _lib.check(_lib.kernels().tfdl_global_avgpool_bwd(_lib.ptr(dy), _lib.ptr(dx), N, H * W, C, _lib.cur_stream()),
                   "global_avgpool_bwd")
        return dx


def global_avg_pool(x: torch.Tensor) -> torch.Tensor:
    return _GlobalAvgPoolFn.apply(x)
