"""Fused self-attention on tcgen05 (csrc/kernels/attention.cu): S = 128 tokens, head dim 64, no mask - BERT-base at the
sequence length of BASELINE.json config 4. Input is the fused query|key|value projection [B, S, 3H]; the
kernels cut the per-head tiles out of it with TMA, so there is no head split, transpose or concat anywhere."""
from __future__ import annotations

import torch

from . import _lib


def supported(qkv: torch.Tensor, heads: int) -> bool:
    return bool(qkv.is_cuda and qkv.dtype == torch.bfloat16 and qkv.dim() == 3 and qkv.shape[1] == 128
                and qkv.shape[2] == 3 * heads * 64)


class _AttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, heads):
        B, S, H3 = qkv.shape
        H = H3 // 3
        qkv = qkv.contiguous()
        out = torch.empty((B, S, H), device=qkv.device, dtype=torch.bfloat16)
        lse2 = torch.empty((B * heads, S), device=qkv.device, dtype=torch.float32)
        _lib.check(_lib.kernels().tfdl_attention_fwd(_lib.ptr(qkv), _lib.ptr(out), _lib.ptr(lse2), B, S, H, heads,
                                                     _lib.cur_stream()), "tfdl_attention_fwd")
        ctx.save_for_backward(qkv, lse2)
        ctx.heads = heads
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, lse2 = ctx.saved_tensors
        B, S, H3 = qkv.shape
        do = (dout if dout.dtype == torch.bfloat16 else dout.bfloat16()).contiguous()
        dqkv = torch.empty_like(qkv)
        _lib.check(_lib.kernels().tfdl_attention_bwd(_lib.ptr(qkv), _lib.ptr(do), _lib.ptr(lse2), _lib.ptr(dqkv), B, S, H3 // 3,
                                                     ctx.heads, _lib.cur_stream()), "tfdl_attention_bwd")
        return dqkv, None


def attention(qkv: torch.Tensor, heads: int) -> torch.Tensor:
    """softmax(Q K^T / 8) V per head. qkv: bf16 [B, 128, 3 * heads * 64] -> bf16 [B, 128, heads * 64]."""
    return _AttentionFn.apply(qkv, heads)
