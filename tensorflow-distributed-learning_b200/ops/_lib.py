"""ctypes loader for the in-tree native libraries.

libtfdl_kernels.so (CUDA, sm_100a) and libtfdl_runtime.so (host C++) are built by
tools/build_native.py into <package>/_native/. On a machine with a GPU a missing kernels
library is a hard error: the framework never silently falls back to eager PyTorch there.
"""
from __future__ import annotations

import ctypes
import os
import threading

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_NATIVE = os.path.join(_PKG, "_native")
_lock = threading.Lock()
_kernels = None
_runtime = None


class NativeError(RuntimeError):
    pass


def native_dir() -> str:
    return _NATIVE


def _load(name: str) -> ctypes.CDLL:
    path = os.path.join(_NATIVE, name)
    if not os.path.exists(path):
        raise NativeError(
            f"{path} is missing; run `python tools/build_native.py` (or __graft_entry__.build()) first")
    return ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)


def kernels() -> ctypes.CDLL:
    """The CUDA kernel library. Raises NativeError if it has not been built."""
    global _kernels
    with _lock:
        if _kernels is None:
            lib = _load("libtfdl_kernels.so")
            lib.tfdl_last_error.restype = ctypes.c_char_p
            _kernels = lib
    return _kernels


def runtime() -> ctypes.CDLL:
    """The host runtime library (control plane, TCP ring, data pipeline, checkpoint IO)."""
    global _runtime
    with _lock:
        if _runtime is None:
            lib = _load("libtfdl_runtime.so")
            lib.tfdl_rt_last_error.restype = ctypes.c_char_p
            _runtime = lib
    return _runtime


def kernels_available() -> bool:
    return os.path.exists(os.path.join(_NATIVE, "libtfdl_kernels.so"))


native_calls = 0  # entry points of libtfdl_kernels.so invoked so far (each launches at least one kernel)


def check(rc: int, what: str = "") -> None:
    global native_calls
    native_calls += 1
    if rc != 0:
        msg = kernels().tfdl_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'native call'} failed (rc={rc}): {msg}")


def check_rt(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = runtime().tfdl_rt_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'runtime call'} failed (rc={rc}): {msg}")


def ptr(t) -> ctypes.c_void_p:
    """Device/host pointer of a torch tensor (None -> NULL)."""
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def cur_stream() -> ctypes.c_void_p:
    import torch

    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
