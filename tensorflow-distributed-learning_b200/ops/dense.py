"""Small-batch Dense forward on tcgen05 with split-K (ref: tf.keras.layers.Dense, tf_dist_example.py:46-47)."""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib


def dense_forward(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False,
                  scratch: Optional[torch.Tensor] = None, k_split: int = 0) -> torch.Tensor:
    """out[B,N] (bf16) = act(x[B,K] @ w[N,K]^T + bias). `scratch`: fp32 [B*N] zeros (reused across calls;
    the kernel hands it back zeroed). Allocated here when not given."""
    assert x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and x.is_cuda
    x, w = x.contiguous(), w.contiguous()
    B, K = x.shape
    N = w.shape[0]
    if scratch is None:
        scratch = torch.zeros(B * N, dtype=torch.float32, device=x.device)
    out = torch.empty((B, N), dtype=torch.bfloat16, device=x.device)
    rc = _lib.kernels().tfdl_dense_fwd_splitk(_lib.ptr(x), _lib.ptr(w), _lib.ptr(bias), _lib.ptr(out), _lib.ptr(scratch),
                                              B, N, K, int(relu), int(k_split), _lib.cur_stream())
    _lib.check(rc, "tfdl_dense_fwd_splitk")
    return out
