"""Metrics (ref: /root/reference/tf_dist_example.py:52 SparseCategoricalAccuracy).

State is a (total, count) pair of sync-on-read accumulators: every replica adds locally on its own
device, and `result()` sums replicas and workers (one tiny all-reduce) only when somebody reads it -
so the training loop never synchronises with the device for metrics."""
from __future__ import annotations

from typing import List, Sequence

import numpy as np
import torch


class Metric:
    name = "metric"

    def __init__(self, name: str | None = None) -> None:
        if name:
            self.name = name
        self._total: List[torch.Tensor] = []
        self._count: List[torch.Tensor] = []

    def bind(self, devices: Sequence[torch.device]) -> None:
        self._total = [torch.zeros((), dtype=torch.float32, device=d) for d in devices]
        self._count = [torch.zeros((), dtype=torch.float32, device=d) for d in devices]

    def batch_values(self, y_true: torch.Tensor, y_pred: torch.Tensor) -> torch.Tensor:
        """One value per example."""
        raise NotImplementedError

    def update_state(self, y_true: torch.Tensor, y_pred: torch.Tensor, replica: int = 0) -> None:
        v = self.batch_values(y_true, y_pred)
        self._total[replica] += v.sum()
        self._count[replica] += float(v.numel())

    def add(self, replica: int, total, count) -> None:
        self._total[replica] += total
        self._count[replica] += count

    def local_state(self) -> np.ndarray:
        return np.array([sum(float(t.item()) for t in self._total), sum(float(c.item()) for c in self._count)])

    def reset_state(self) -> None:
        for t in self._total + self._count:
            t.zero_()

    reset_states = reset_state


class Mean(Metric):
    name = "mean"

    def update_state(self, values: torch.Tensor, weight: float = 1.0, replica: int = 0) -> None:  # type: ignore[override]
        self._total[replica] += values.detach().sum() * weight
        self._count[replica] += float(values.numel()) * weight


class SparseCategoricalAccuracy(Metric):
    name = "sparse_categorical_accuracy"

    def batch_values(self, y_true: torch.Tensor, y_pred: torch.Tensor) -> torch.Tensor:
        yp = y_pred.reshape(-1, y_pred.shape[-1])
        return (yp.argmax(-1) == y_true.reshape(-1).long()).float()


class CategoricalAccuracy(Metric):
    name = "categorical_accuracy"

    def batch_values(self, y_true: torch.Tensor, y_pred: torch.Tensor) -> torch.Tensor:
        return (y_pred.argmax(-1) == y_true.argmax(-1)).float()


class BinaryAccuracy(Metric):
    name = "binary_accuracy"

    def __init__(self, name: str | None = None, threshold: float = 0.5) -> None:
        super().__init__(name)
        self.threshold = threshold

    def batch_values(self, y_true: torch.Tensor, y_pred: torch.Tensor) -> torch.Tensor:
        hit = ((y_pred.float() > self.threshold) == (y_true.float().reshape(y_pred.shape) > 0.5)).float()
        return hit.reshape(hit.shape[0], -1).mean(-1)


def get(x, loss=None) -> Metric:
    if isinstance(x, Metric):
        return x
    if x in ("accuracy", "acc"):
        from .losses import BinaryCrossentropy, CategoricalCrossentropy

        # Keras picks the accuracy that matches the loss
        m: Metric = (CategoricalAccuracy() if isinstance(loss, CategoricalCrossentropy)
                     else BinaryAccuracy(threshold=0.0 if getattr(loss, "from_logits", False) else 0.5)
                     if isinstance(loss, BinaryCrossentropy) else SparseCategoricalAccuracy())
        m.name = "accuracy"
        return m
    if x == "binary_accuracy":
        return BinaryAccuracy()
    if x == "sparse_categorical_accuracy":
        return SparseCategoricalAccuracy()
    if x == "categorical_accuracy":
        return CategoricalAccuracy()
    raise ValueError(f"unknown metric {x!r}")
