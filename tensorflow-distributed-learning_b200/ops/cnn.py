"""Python entry points of the CNN step kernels (csrc/kernels/cnn.cu), one per kernel, used by the
numerics tests and by anyone who wants a single fused op. The training path does not come through
here: keras/native_engine.py hands the whole step to the C++ program executor."""
from __future__ import annotations

import ctypes as C
from typing import Tuple

import torch

from . import _lib
from .structs import CnnDims, make_dims

_s = _lib.cur_stream
_p = _lib.ptr


def conv1_fwd(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor, d: CnnDims) -> Tuple[torch.Tensor, torch.Tensor]:
    """x [B,H0,W0] fp32, w [C1,9] fp32 (OHWI with I=1), b [C1] -> a1 bf16 [B,P1h,P1w,C1], idx1 uint8."""
    a1 = torch.empty((d.B, d.P1h, d.P1w, d.C1), dtype=torch.bfloat16, device=x.device)
    idx = torch.empty((d.B, d.P1h, d.P1w, d.C1), dtype=torch.uint8, device=x.device)
    _lib.check(_lib.kernels().tfdl_cnn_conv1_fwd(_p(x), _p(w), _p(b), _p(a1), _p(idx), None, C.byref(d), _s()), "conv1_fwd")
    return a1, idx


def conv2_fwd(a1: torch.Tensor, w2h: torch.Tensor, b2: torch.Tensor, d: CnnDims,
              impl: str = "gather") -> Tuple[torch.Tensor, torch.Tensor]:
    """a1 bf16 [B,P1h,P1w,C1], w2h bf16 [C2,3,3,C1], b2 fp32 -> a2 bf16 [B,P2h,P2w,C2], idx2 uint8.
    impl: "gather" (cp.async im2col tile) or "tma" (im2col-free, shifted descriptors)."""
    a2 = torch.empty((d.B, d.P2h, d.P2w, d.C2), dtype=torch.bfloat16, device=a1.device)
    idx = torch.empty((d.B, d.P2h, d.P2w, d.C2), dtype=torch.uint8, device=a1.device)
    fn = _lib.kernels().tfdl_cnn_conv2_fwd_tma if impl == "tma" else _lib.kernels().tfdl_cnn_conv2_fwd
    _lib.check(fn(_p(a1), _p(w2h), _p(b2), _p(a2), _p(idx), C.byref(d), _s()), "conv2_fwd")
    return a2, idx


def head(acc: torch.Tensor, b1: torch.Tensor, w2: torch.Tensor, b2: torch.Tensor, labels: torch.Tensor, d: CnnDims):
    """acc fp32 [B,D1] (zeroed on return), returns dh bf16, g_w2, g_b2, g_b1, metrics[3], step_out[2]."""
    dev = acc.device
    dh = torch.empty((d.B, d.D1), dtype=torch.bfloat16, device=dev)
    g_w2 = torch.zeros((d.NC, d.D1), device=dev)  # the kernel accumulates into these
    g_b2 = torch.zeros((d.NC,), device=dev)
    g_b1 = torch.zeros((d.D1,), device=dev)
    metrics = torch.zeros(3, device=dev)
    step_out = torch.zeros(2, device=dev)
    _lib.check(_lib.kernels().tfdl_cnn_head(_p(acc), _p(b1), _p(w2), _p(b2), _p(labels), _p(dh), _p(g_w2), _p(g_b2),
                                            _p(g_b1), _p(metrics), _p(step_out), C.byref(d), _s()), "head")
    return dh, g_w2, g_b2, g_b1, metrics, step_out


def dense1_dgrad(dh: torch.Tensor, w1h: torch.Tensor, a2: torch.Tensor, idx2: torch.Tensor, d: CnnDims) -> torch.Tensor:
    dy2 = torch.zeros((d.B, d.P1h, d.P1w, d.C2), dtype=torch.bfloat16, device=dh.device)  # conv2 input grid
    _lib.check(_lib.kernels().tfdl_cnn_dense1_dgrad(_p(dh), _p(w1h), _p(a2), _p(idx2), _p(dy2), C.byref(d), _s()),
               "dense1_dgrad")
    return dy2


def conv2_wgrad(a1: torch.Tensor, dy2: torch.Tensor, d: CnnDims, impl: str = "gather") -> Tuple[torch.Tensor, torch.Tensor]:
    """dy2 bf16 on the conv2 input grid [B,P1h,P1w,C2] (zero outside [0:2P2h, 0:2P2w])."""
    g_w = torch.zeros((d.C2, 3, 3, d.C1), device=a1.device)
    g_b = torch.zeros((d.C2,), device=a1.device)
    fn = _lib.kernels().tfdl_cnn_conv2_wgrad_tma if impl == "tma" else _lib.kernels().tfdl_cnn_conv2_wgrad
    _lib.check(fn(_p(a1), _p(dy2), _p(g_w), _p(g_b), C.byref(d), _s()), "conv2_wgrad")
    return g_w, g_b


def conv2_dgrad(dy2: torch.Tensor, w2h: torch.Tensor, a1: torch.Tensor, d: CnnDims, impl: str = "gather") -> torch.Tensor:
    g1 = torch.empty((d.B, d.P1h, d.P1w, d.C1), dtype=torch.bfloat16, device=a1.device)
    fn = _lib.kernels().tfdl_cnn_conv2_dgrad_tma if impl == "tma" else _lib.kernels().tfdl_cnn_conv2_dgrad
    _lib.check(fn(_p(dy2), _p(w2h), _p(a1), _p(g1), C.byref(d), _s()), "conv2_dgrad")
    return g1


def conv1_wgrad(x: torch.Tensor, g1: torch.Tensor, idx1: torch.Tensor, d: CnnDims) -> Tuple[torch.Tensor, torch.Tensor]:
    g_w = torch.zeros((d.C1, 9), device=x.device)
    g_b = torch.zeros((d.C1,), device=x.device)
    _lib.check(_lib.kernels().tfdl_cnn_conv1_wgrad(_p(x), _p(g1), _p(idx1), _p(g_w), _p(g_b), C.byref(d), _s()),
               "conv1_wgrad")
    return g_w, g_b


def conv12_fwd(x: torch.Tensor, w1: torch.Tensor, b1: torch.Tensor, w2h: torch.Tensor, b2: torch.Tensor, d: CnnDims):
    """Fused conv1+pool+relu -> conv2+pool+relu forward. Returns (a1, idx1, a2, idx2)."""
    dev = x.device
    a1 = torch.empty((d.B, d.P1h, d.P1w, d.C1), dtype=torch.bfloat16, device=dev)
    i1 = torch.empty((d.B, d.P1h, d.P1w, d.C1), dtype=torch.uint8, device=dev)
    a2 = torch.empty((d.B, d.P2h, d.P2w, d.C2), dtype=torch.bfloat16, device=dev)
    i2 = torch.empty((d.B, d.P2h, d.P2w, d.C2), dtype=torch.uint8, device=dev)
    _lib.check(_lib.kernels().tfdl_cnn_conv12_fwd_tma(_p(x), _p(w1), _p(b1), _p(w2h), _p(b2), _p(a1), _p(i1), _p(a2), _p(i2),
                                                      None, C.byref(d), _s()), "conv12_fwd_tma")
    return a1, i1, a2, i2


def conv21_bwd(dy2: torch.Tensor, w2h: torch.Tensor, a1: torch.Tensor, idx1: torch.Tensor, x: torch.Tensor, d: CnnDims):
    """Fused conv2 dgrad (+ReLU mask) -> conv1 wgrad. Returns (g_w1c [C1,9], g_b1c [C1], g1 bf16)."""
    dev = x.device
    g_w = torch.zeros((d.C1, 9), device=dev)
    g_b = torch.zeros((d.C1,), device=dev)
    g1 = torch.zeros((d.B, d.P1h, d.P1w, d.C1), dtype=torch.bfloat16, device=dev)
    _lib.check(_lib.kernels().tfdl_cnn_conv21_bwd_tma(_p(dy2), _p(w2h), _p(a1), _p(idx1), _p(x), _p(g_w), _p(g_b), _p(g1),
                                                      C.byref(d), _s()), "conv21_bwd_tma")
    return g_w, g_b, g1


def dense_block(a2: torch.Tensor, idx2: torch.Tensor, w1h: torch.Tensor, w2h: torch.Tensor, b1: torch.Tensor,
                b2: torch.Tensor, labels: torch.Tensor, d: CnnDims, acc: torch.Tensor = None,
                acc_next: torch.Tensor = None, sync: torch.Tensor = None):
    """Fused Dense(128,relu)+Dense(NC)+softmax-xent forward/backward (csrc/kernels/dense_block.cu).
    Returns dict(dy2, g_w1, g_b1, g_w2, g_b2, metrics, step_out, dh)."""
    dev = a2.device
    acc = torch.zeros(d.B, d.D1, device=dev) if acc is None else acc
    acc_next = torch.ones(d.B, d.D1, device=dev) if acc_next is None else acc_next
    sync = torch.zeros(4, dtype=torch.int32, device=dev) if sync is None else sync
    out = dict(dy2=torch.zeros((d.B, d.P1h, d.P1w, d.C2), dtype=torch.bfloat16, device=dev),
               g_w1=torch.zeros(d.D1, d.F, device=dev), g_b1=torch.zeros(d.D1, device=dev),
               g_w2=torch.zeros(d.NC, d.D1, device=dev), g_b2=torch.zeros(d.NC, device=dev),
               metrics=torch.zeros(3, device=dev), step_out=torch.zeros(2, device=dev),
               dh=torch.zeros(d.B, d.D1, dtype=torch.bfloat16, device=dev), acc=acc, acc_next=acc_next, sync=sync)
    _lib.check(_lib.kernels().tfdl_cnn_dense_block(
        _p(a2), _p(idx2), _p(w1h), _p(w2h), _p(b1), _p(b2), _p(labels), _p(acc), _p(acc_next), _p(sync), _p(out["dy2"]),
        _p(out["g_w1"]), _p(out["g_b1"]), _p(out["g_w2"]), _p(out["g_b2"]), _p(out["metrics"]), _p(out["step_out"]),
        _p(out["dh"]), C.byref(d), _s()), "dense_block")
    return out


__all__ = ["conv12_fwd", "conv21_bwd", "dense_block", "conv1_fwd", "conv2_fwd", "head", "dense1_dgrad", "conv2_wgrad", "conv2_dgrad", "conv1_wgrad", "make_dims",
           "CnnDims"]
