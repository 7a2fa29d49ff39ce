"""Tracing helpers: NVTX ranges around step / forward / backward / all-reduce, and a CUDA-event
step timer. The reference configures no profiler (tf_dist_example.py:59 passes no callbacks; README.md:51 only
says the chief "generates TensorBoard"), so this is the TensorBoard-callback `profile_batch` equivalent.

NVTX ranges cost a Python call each, so they are off unless ``TFDL_NVTX=1`` (or ``enable(True)``); with them on,
an Nsight Systems / ncu timeline shows ``tfdl/step``, ``tfdl/forward``, ``tfdl/backward``, ``tfdl/allreduce+sgd``
and, for the native engine, ``tfdl/native_chunk[n]``."""
from __future__ import annotations

import contextlib
import os
from typing import Iterator, List, Optional

import torch

_ENABLED = os.environ.get("TFDL_NVTX", "0") == "1"


def enable(on: bool = True) -> None:
    global _ENABLED
    _ENABLED = bool(on)


def enabled() -> bool:
    return _ENABLED


@contextlib.contextmanager
def nvtx_range(name: str) -> Iterator[None]:
    """NVTX push/pop when tracing is on and CUDA is present; free otherwise."""
    if _ENABLED and torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


class StepTimer:
    """Device-side step timing with CUDA events (host clock on CPU): ``with timer.step(): ...`` per step, then
    ``timer.summary()`` -> {"steps", "mean_ms", "p50_ms", "max_ms"}. Events are read only in ``summary`` so
    timing adds no synchronisation to the loop."""

    def __init__(self, device: Optional[torch.device] = None) -> None:
        self.device = device
        self.cuda = torch.cuda.is_available() and (device is None or device.type == "cuda")
        self._pairs: List[tuple] = []
        self._host: List[float] = []

    @contextlib.contextmanager
    def step(self) -> Iterator[None]:
        if self.cuda:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            yield
            e1.record()
            self._pairs.append((e0, e1))
        else:
            import time

            t0 = time.perf_counter()
            yield
            self._host.append((time.perf_counter() - t0) * 1e3)

    def summary(self) -> dict:
        if self.cuda:
            torch.cuda.synchronize(self.device)
            ms = [a.elapsed_time(b) for a, b in self._pairs]
        else:
            ms = list(self._host)
        if not ms:
            return {"steps": 0}
        s = sorted(ms)
        return {"steps": len(ms), "mean_ms": sum(ms) / len(ms), "p50_ms": s[len(s) // 2], "max_ms": s[-1]}
