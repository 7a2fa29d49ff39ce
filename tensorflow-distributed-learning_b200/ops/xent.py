"""Sparse softmax cross-entropy on bf16 logits (csrc/kernels/softmax_xent.cu): the loss of the reference
(SparseCategoricalCrossentropy(from_logits=True), tf_dist_example.py:50) for wide vocabularies, where the eager
formulation spends ~9 passes over a 1 GB fp32 logits tensor."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib


def _padded_base(logits2d: torch.Tensor):
    """(buffer, V, ld) when `logits2d` [T, V] is contiguous, or a column slice of a contiguous [T, ld] buffer whose pad
    columns the producer filled with a very negative value (ops/gemm.py:tied_decoder)."""
    T, V = logits2d.shape
    ld = logits2d.stride(0)
    if logits2d.stride(1) != 1 or ld < V or ld % 8 or logits2d.data_ptr() % 16:
        return None
    return torch.as_strided(logits2d, (T, ld), (ld, 1)) if ld != V else logits2d, V, ld


def supported(logits: torch.Tensor) -> bool:
    if not (logits.is_cuda and logits.dtype == torch.bfloat16 and logits.dim() >= 2):
        return False
    V = logits.shape[-1]
    try:
        l2 = logits.view(-1, V)  # must be a view: a copy would lose the padded layout (and cost a pass)
    except RuntimeError:
        return False
    pb = _padded_base(l2)
    return pb is not None and bool(_lib.kernels().tfdl_softmax_xent_supported(V, pb[2]))


class _XentFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, ignore_index):
        V = logits.shape[-1]
        l2 = logits.view(-1, V)
        buf, V, ld = _padded_base(l2)
        T = l2.shape[0]
        y = labels.reshape(-1).to(torch.int64).contiguous()
        loss = torch.empty(T, device=logits.device, dtype=torch.float32)
        probs = torch.empty((T, ld), device=logits.device, dtype=torch.bfloat16)
        _lib.check(_lib.kernels().tfdl_softmax_xent_fwd(_lib.ptr(buf), _lib.ptr(y), _lib.ptr(loss), _lib.ptr(probs),
                                                        C.c_longlong(T), V, ld, C.c_longlong(ignore_index),
                                                        _lib.cur_stream()), "tfdl_softmax_xent_fwd")
        ctx.save_for_backward(probs, y)
        ctx.meta = (V, ld, ignore_index, logits.shape)
        return loss

    @staticmethod
    def backward(ctx, gloss):
        probs, y = ctx.saved_tensors
        V, ld, ignore_index, shape = ctx.meta
        T = probs.shape[0]
        g = gloss.reshape(-1).float().contiguous()
        # in place: the probabilities become dlogits (bf16, pad columns stay exactly zero)
        _lib.check(_lib.kernels().tfdl_softmax_xent_bwd(_lib.ptr(probs), _lib.ptr(y), _lib.ptr(g), C.c_longlong(T), V, ld,
                                                        C.c_longlong(ignore_index), _lib.cur_stream()), "tfdl_softmax_xent_bwd")
        return probs[:, :V].view(shape), None, None


def sparse_softmax_xent(logits: torch.Tensor, labels: torch.Tensor, ignore_index: int = -(1 << 62)) -> torch.Tensor:
    """Per-example loss [T] (fp32) of bf16 logits [..., V]; rows whose label equals ``ignore_index`` give 0 and no
    gradient. The gradient comes back as a column slice of a padded bf16 buffer, which tied_decoder's backward uses
    in place."""
    return _XentFn.apply(logits, labels, int(ignore_index))
