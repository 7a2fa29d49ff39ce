"""dtype names and tf.cast look-alike (ref: /root/reference/tf_dist_example.py:23 `tf.cast(image, tf.float32)`)."""
from __future__ import annotations

import numpy as np

float32 = "float32"
float16 = "float16"
bfloat16 = "bfloat16"
int32 = "int32"
int64 = "int64"
uint8 = "uint8"


def as_numpy_dtype(dt):
    if isinstance(dt, str):
        if dt == "bfloat16":
            raise TypeError("numpy has no bfloat16; cast on the torch side")
        return np.dtype(dt)
    return np.dtype(dt)


def as_torch_dtype(dt):
    import torch

    if isinstance(dt, torch.dtype):
        return dt
    return {"float32": torch.float32, "float16": torch.float16, "bfloat16": torch.bfloat16, "int32": torch.int32,
            "int64": torch.int64, "uint8": torch.uint8, "float64": torch.float64, "bool": torch.bool}[str(np.dtype(dt)) if not isinstance(dt, str) else dt]


def cast(x, dtype):
    """Works on numpy arrays (dataset map functions) and torch tensors."""
    try:
        import torch

        if isinstance(x, torch.Tensor):
            return x.to(as_torch_dtype(dtype))
    except ImportError:  # pragma: no cover
        pass
    return np.asarray(x).astype(as_numpy_dtype(dtype))
