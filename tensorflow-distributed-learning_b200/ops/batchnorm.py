"""BatchNorm (+ fused ReLU / residual add) on the native streaming kernels (csrc/kernels/batchnorm.cu): NHWC bf16
activations, fp32 parameters and statistics. Used by keras.layers.BatchNormalization under the mixed_bfloat16 policy on
CUDA - the ResNet-50 config of BASELINE.json; the reference repo itself has no normalisation layer."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib


def supported(x: torch.Tensor) -> bool:
    c = x.shape[-1]
    return bool(x.is_cuda and x.dtype == torch.bfloat16 and x.dim() >= 2 and c % 8 == 0 and 8 <= c <= 2048
                and ((c // 8) & (c // 8 - 1)) == 0)


class _BatchNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, gamma, beta, running_mean, running_var, momentum_new, eps, relu, in_bias):
        Cn = x.shape[-1]
        x = x.contiguous()
        M = x.numel() // Cn
        r = residual.contiguous() if residual is not None else None
        g, b = gamma.float(), beta.float()
        dev = x.device
        sums = torch.empty(2 * Cn, device=dev, dtype=torch.float32)
        stats = torch.empty(2, Cn, device=dev, dtype=torch.float32)
        y = torch.empty_like(x)
        _lib.check(_lib.kernels().tfdl_batchnorm_fwd(
            _lib.ptr(x), _lib.ptr(r), _lib.ptr(in_bias), _lib.ptr(g), _lib.ptr(b), _lib.ptr(running_mean), _lib.ptr(running_var),
            _lib.ptr(sums), _lib.ptr(stats[0]), _lib.ptr(stats[1]), _lib.ptr(y), C.c_longlong(M), Cn, C.c_float(eps),
            C.c_float(momentum_new), int(relu), 1, _lib.cur_stream()), "tfdl_batchnorm_fwd")
        ctx.save_for_backward(x, y if relu else None, stats, g)
        ctx.meta = (bool(relu), residual is not None, gamma.dtype, beta.dtype)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, stats, g = ctx.saved_tensors
        relu, has_res, gdt, bdt = ctx.meta
        Cn = x.shape[-1]
        M = x.numel() // Cn
        dy = (dy if dy.dtype == torch.bfloat16 else dy.bfloat16()).contiguous()
        dx = torch.empty_like(x)
        dres = torch.empty_like(x) if has_res else None
        dgb = torch.empty(2, Cn, device=x.device, dtype=torch.float32)
        sums = torch.empty(2 * Cn, device=x.device, dtype=torch.float32)
        _lib.check(_lib.kernels().tfdl_batchnorm_bwd(
            _lib.ptr(dy), _lib.ptr(x), _lib.ptr(y), _lib.ptr(stats[0]), _lib.ptr(stats[1]), _lib.ptr(g), _lib.ptr(sums),
            _lib.ptr(dx), _lib.ptr(dres), _lib.ptr(dgb[0]), _lib.ptr(dgb[1]), C.c_longlong(M), Cn, int(relu),
            _lib.cur_stream()), "tfdl_batchnorm_bwd")
        return dx, dres, dgb[0].to(gdt), dgb[1].to(bdt), None, None, None, None, None, None


def batch_norm_train(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, running_mean: torch.Tensor,
                     running_var: torch.Tensor, momentum_new: float, eps: float, relu: bool = False,
                     residual: Optional[torch.Tensor] = None, in_bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y = [relu](BN_batch(x + in_bias) + residual), channels last. ``in_bias`` (a conv bias the producer skipped) costs
    nothing: batch statistics absorb it, it only moves the running mean, and its gradient is exactly zero. The running statistics are updated in place by the kernel
    (new = (1 - momentum_new) * old + momentum_new * batch; the variance with Bessel's correction, as TF does)."""
    ib = in_bias.detach().float() if in_bias is not None else None
    return _BatchNormFn.apply(x, residual, gamma, beta, running_mean, running_var, float(momentum_new), float(eps), bool(relu), ib)


def batch_norm_infer(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, running_mean: torch.Tensor,
                     running_var: torch.Tensor, eps: float, relu: bool = False,
                     residual: Optional[torch.Tensor] = None, in_bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    Cn = x.shape[-1]
    x = x.contiguous()
    y = torch.empty_like(x)
    r = residual.contiguous() if residual is not None else None
    _lib.check(_lib.kernels().tfdl_batchnorm_fwd(
        _lib.ptr(x), _lib.ptr(r), _lib.ptr(in_bias.float() if in_bias is not None else None), _lib.ptr(gamma.float()),
        _lib.ptr(beta.float()), _lib.ptr(running_mean.float()),
        _lib.ptr(running_var.float()), None, None, None, _lib.ptr(y), C.c_longlong(x.numel() // Cn), Cn, C.c_float(eps),
        C.c_float(0.0), int(relu), 0, _lib.cur_stream()), "tfdl_batchnorm_fwd")
    return y
