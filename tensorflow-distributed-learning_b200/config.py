"""Process-wide switches: seeds and op determinism (ref: `tf.keras.utils.set_random_seed`,
`tf.config.experimental.enable_op_determinism` - what a user of the reference's stack reaches for when two runs of
/root/reference/tf_dist_example.py have to give the same weights).

The fast path of this framework is not bit-reproducible from run to run: the fused CNN kernels, split-K GEMMs, the
convolution weight gradient and the BatchNorm / LayerNorm column sums accumulate with fp32 atomics, whose order varies.
`enable_op_determinism()` trades that speed for reproducibility, like TensorFlow's switch does:

* models train on the portable engine (autograd over the flat arenas) with the layer ops on PyTorch's library kernels in
  their deterministic modes (`torch.use_deterministic_algorithms`, cuDNN deterministic, no autotuning);
* gradients are still summed by this framework's NVLink kernels, restricted to the ones that add the ranks'
  contributions in a fixed order (push, push2, one-shot, two-shot); the NVLS multicast kernel, whose in-switch reduction
  order is not specified, is not used; the fused optimizer update is element-wise and needs no change;
* data order and initial weights are reproducible once `set_random_seed` has been called.

Call both before the model is built. `TFDL_DETERMINISTIC=1` in the environment enables the switch at import.
"""
from __future__ import annotations

import os
import random
from typing import Optional

import numpy as np

_state = {"deterministic": False, "saved": None}


def set_random_seed(seed: Optional[int]) -> None:
    """Seed everything that draws random numbers: Python, NumPy, PyTorch, the weight initialisers and the dataset
    shuffles that were not given a seed of their own (None restores fresh entropy for the last two)."""
    import torch

    from .data import dataset as _dataset
    from .keras import layers as _layers

    if seed is not None:
        random.seed(seed)
        np.random.seed(seed % (1 << 32))
        torch.manual_seed(seed)
    _layers.set_init_seed(seed)
    _dataset.set_global_seed(seed)


def is_op_determinism_enabled() -> bool:
    return bool(_state["deterministic"])


def enable_op_determinism() -> None:
    import torch

    if _state["deterministic"]:
        return
    _state["saved"] = (torch.are_deterministic_algorithms_enabled(), torch.is_deterministic_algorithms_warn_only_enabled(),
                       torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark)
    # cuBLAS only promises run-to-run reproducibility with a fixed workspace; it reads this when its handle is created
    os.environ.setdefault("CUBLAS_WORKSPACE_CONFIG", ":4096:8")
    torch.use_deterministic_algorithms(True, warn_only=True)
    torch.backends.cudnn.deterministic = True
    torch.backends.cudnn.benchmark = False
    _state["deterministic"] = True


def disable_op_determinism() -> None:
    import torch

    if not _state["deterministic"]:
        return
    det, warn, cd, cb = _state["saved"]
    torch.use_deterministic_algorithms(det, warn_only=warn)
    torch.backends.cudnn.deterministic = cd
    torch.backends.cudnn.benchmark = cb
    _state["deterministic"] = False


if os.environ.get("TFDL_DETERMINISTIC", "0") == "1":
    enable_op_determinism()
