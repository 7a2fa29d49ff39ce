"""Distributed values: the flat parameter arena, mirrored variables, sync-on-read variables.

Reference behaviour (README.md:15,17,21): every variable of the model is *mirrored* - one copy per
replica, kept identical by applying the same all-reduced update everywhere. Metric accumulators
are *sync-on-read*: each replica adds locally and the copies are summed when read.

B200-first layout (not TensorFlow's): all trainable variables of a model live back to back in ONE
flat fp32 buffer per replica (the *arena*), with a second buffer of identical layout for
gradients. The gradient all-reduce therefore needs no pack/unpack copy (TF's `num_packs=1`
concatenation), the fused all-reduce+SGD kernel walks both buffers linearly, and on multi-GPU runs
the gradient arena is allocated in symmetric (peer-mapped) memory so peers can read it directly.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

ALIGN_ELEMS = 64  # every variable starts on a 256-byte boundary (vector loads, TMA, multimem)


@dataclass
class VarSpec:
    name: str
    keras_shape: Tuple[int, ...]        # shape as Keras reports it (e.g. HWIO conv kernels)
    storage_shape: Tuple[int, ...]      # shape as stored in the arena (e.g. OHWI)
    to_storage: Optional[Tuple[int, ...]]  # permutation keras -> storage (None = identity)
    offset: int = 0                     # element offset in the arena
    trainable: bool = True

    @property
    def numel(self) -> int:
        n = 1
        for s in self.storage_shape:
            n *= int(s)
        return n

    @property
    def to_keras(self) -> Optional[Tuple[int, ...]]:
        if self.to_storage is None:
            return None
        inv = [0] * len(self.to_storage)
        for i, p in enumerate(self.to_storage):
            inv[p] = i
        return tuple(inv)


class ParamArena:
    """Flat fp32 parameter + gradient buffers for one replica."""

    def __init__(self, specs: Sequence[VarSpec], device: torch.device,
                 allocator: Optional[Callable[[int, torch.device], torch.Tensor]] = None) -> None:
        off = 0
        for s in specs:
            s.offset = off
            off += (s.numel + ALIGN_ELEMS - 1) // ALIGN_ELEMS * ALIGN_ELEMS
        self.specs = list(specs)
        self.numel = max(off, ALIGN_ELEMS)
        self.device = device
        alloc = allocator or (lambda n, d: torch.zeros(n, dtype=torch.float32, device=d))
        self.params = alloc(self.numel, device)
        self.grads = alloc(self.numel, device)
        self.params.zero_()
        self.grads.zero_()

    def view(self, spec: VarSpec, grads: bool = False) -> torch.Tensor:
        buf = self.grads if grads else self.params
        return buf[spec.offset:spec.offset + spec.numel].view(spec.storage_shape)

    @property
    def nbytes(self) -> int:
        return self.numel * 4


class MirroredVariable:
    """One logical variable, one storage view per local replica (README.md:15)."""

    def __init__(self, spec: VarSpec, init_keras: np.ndarray) -> None:
        self.spec = spec
        self._init = np.asarray(init_keras, dtype=np.float32)
        assert tuple(self._init.shape) == tuple(spec.keras_shape), (self._init.shape, spec.keras_shape)
        self._views: List[torch.Tensor] = []
        self._grad_views: List[torch.Tensor] = []

    # -- binding to arenas (done once by the model when it finalises) ------------------------
    def bind(self, arenas: Sequence[ParamArena]) -> None:
        self._views = [a.view(self.spec) for a in arenas]
        self._grad_views = [a.view(self.spec, grads=True) for a in arenas]
        self.assign(self._init)

    @property
    def bound(self) -> bool:
        return bool(self._views)

    # -- Keras-facing API ---------------------------------------------------------------------
    @property
    def name(self) -> str:
        return self.spec.name

    @property
    def shape(self) -> Tuple[int, ...]:
        return self.spec.keras_shape

    @property
    def trainable(self) -> bool:
        return self.spec.trainable

    @property
    def num_replicas(self) -> int:
        return len(self._views)

    def _to_storage(self, a: np.ndarray) -> np.ndarray:
        return a if self.spec.to_storage is None else np.transpose(a, self.spec.to_storage)

    def assign(self, value) -> "MirroredVariable":
        """Assign a value given in Keras layout to every replica copy."""
        a = np.asarray(value, dtype=np.float32)
        if tuple(a.shape) != tuple(self.spec.keras_shape):
            raise ValueError(f"{self.name}: cannot assign shape {a.shape} to variable of shape {self.spec.keras_shape}")
        if not self._views:
            self._init = a.copy()
            return self
        st = torch.from_numpy(np.ascontiguousarray(self._to_storage(a)))
        for v in self._views:
            v.copy_(st)
        return self

    def numpy(self, replica: int = 0) -> np.ndarray:
        """Value in Keras layout (replica 0 unless asked otherwise; copies are identical)."""
        if not self._views:
            return self._init.copy()
        a = self._views[replica].detach().cpu().numpy()
        if self.spec.to_keras is None:
            return a.copy()  # never hand out a view of the live arena
        return np.ascontiguousarray(np.transpose(a, self.spec.to_keras))

    def value(self, replica: int = 0) -> torch.Tensor:
        """Storage-layout tensor of one replica (a view into its arena)."""
        return self._views[replica]

    def grad(self, replica: int = 0) -> torch.Tensor:
        return self._grad_views[replica]

    @property
    def values(self) -> List[torch.Tensor]:
        return list(self._views)

    def __repr__(self) -> str:
        return f"<MirroredVariable {self.name} shape={self.shape} replicas={len(self._views)}>"


class SyncOnReadVariable:
    """Per-replica accumulator, summed across replicas (and workers) when read - the
    aggregation Keras metrics use (`total` / `count`)."""

    def __init__(self, name: str, devices: Sequence[torch.device], dtype: torch.dtype = torch.float64) -> None:
        self.name = name
        self._vals = [torch.zeros((), dtype=dtype if d.type != "cuda" else torch.float32, device=d) for d in devices]

    def add(self, replica: int, v) -> None:
        self._vals[replica] += v

    def local_sum(self) -> float:
        return float(sum(v.item() for v in self._vals))

    def reset(self) -> None:
        for v in self._vals:
            v.zero_()


class PerReplica:
    """A value per local replica (what `strategy.run` consumes and returns)."""

    def __init__(self, values: Sequence[object]) -> None:
        self.values = tuple(values)

    def __len__(self) -> int:
        return len(self.values)

    def __iter__(self):
        return iter(self.values)

    def __getitem__(self, i: int):
        return self.values[i]

    def __repr__(self) -> str:
        return f"PerReplica({list(self.values)!r})"
