"""LayerNorm on the native streaming kernels (csrc/kernels/layernorm.cu): bf16 activations, fp32 gamma / beta /
statistics, optional fused residual add. Used by keras.layers.LayerNormalization under the mixed_bfloat16 policy on
CUDA (BASELINE.json config 4; the reference repo itself has no transformer)."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib


def supported(x: torch.Tensor) -> bool:
    return bool(x.is_cuda and x.dtype == torch.bfloat16 and x.shape[-1] % 256 == 0 and 1 <= x.shape[-1] // 256 <= 4)


class _LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, gamma, beta, eps):
        H = x.shape[-1]
        x2 = x.reshape(-1, H).contiguous()
        T = x2.shape[0]
        r2 = residual.reshape(-1, H).contiguous() if residual is not None else None
        g, b = gamma.float().contiguous(), beta.float().contiguous()
        y = torch.empty_like(x2)
        z = torch.empty_like(x2) if r2 is not None else x2
        mean = torch.empty(T, device=x.device, dtype=torch.float32)
        rstd = torch.empty(T, device=x.device, dtype=torch.float32)
        _lib.check(_lib.kernels().tfdl_layernorm_fwd(_lib.ptr(x2), _lib.ptr(r2), _lib.ptr(g), _lib.ptr(b), _lib.ptr(y),
                                                     _lib.ptr(z) if r2 is not None else None, _lib.ptr(mean),
                                                     _lib.ptr(rstd), T, H, C.c_float(eps), _lib.cur_stream()),
                   "tfdl_layernorm_fwd")
        ctx.save_for_backward(z, mean, rstd, g)
        ctx.meta = (x.shape, residual is not None, gamma.dtype, beta.dtype)
        return y.reshape(x.shape)

    @staticmethod
    def backward(ctx, dy):
        z, mean, rstd, g = ctx.saved_tensors
        shape, has_res, gdt, bdt = ctx.meta
        T, H = z.shape
        dy2 = (dy if dy.dtype == torch.bfloat16 else dy.bfloat16()).reshape(T, H).contiguous()
        dz = torch.empty_like(z)
        dgb = torch.zeros(2, H, device=z.device, dtype=torch.float32)
        _lib.check(_lib.kernels().tfdl_layernorm_bwd(_lib.ptr(dy2), _lib.ptr(z), _lib.ptr(mean), _lib.ptr(rstd), _lib.ptr(g),
                                                     _lib.ptr(dz), _lib.ptr(dgb[0]), _lib.ptr(dgb[1]), T, H,
                                                     _lib.cur_stream()), "tfdl_layernorm_bwd")
        dz = dz.reshape(shape)
        return dz, (dz if has_res else None), dgb[0].to(gdt), dgb[1].to(bdt), None


def layer_norm(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float = 1e-3,
               residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y = LayerNorm(x + residual) over the last axis; x (and residual) bf16 CUDA tensors, see `supported`."""
    return _LayerNormFn.apply(x, residual, gamma, beta, float(eps))
