"""Small shared helpers."""
from .dtypes import cast, float32, float16, bfloat16, int32, int64, uint8, as_torch_dtype, as_numpy_dtype
from .trace import StepTimer, nvtx_range
from . import trace

__all__ = ["cast", "float32", "float16", "bfloat16", "int32", "int64", "uint8", "as_torch_dtype", "as_numpy_dtype", "StepTimer", "nvtx_range", "trace"]
