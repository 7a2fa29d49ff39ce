"""Losses (ref: /root/reference/tf_dist_example.py:50 SparseCategoricalCrossentropy(from_logits=True)).

`per_example` returns one value per sample; the training engines turn it into the Keras distributed
loss: mean over the replica batch, divided by num_replicas_in_sync, so that the SUM all-reduce of
gradients yields the gradient of the mean over the GLOBAL batch."""
from __future__ import annotations

import torch
import torch.nn.functional as F


class Loss:
    name = "loss"

    def per_example(self, y_true: torch.Tensor, y_pred: torch.Tensor) -> torch.Tensor:
        raise NotImplementedError

    def __call__(self, y_true: torch.Tensor, y_pred: torch.Tensor) -> torch.Tensor:
        return self.per_example(y_true, y_pred).mean()

    def get_config(self) -> dict:
        return {k: v for k, v in vars(self).items() if not k.startswith("_")}

    @classmethod
    def from_config(cls, config: dict) -> "Loss":
        return cls(**config)


class SparseCategoricalCrossentropy(Loss):
    name = "sparse_categorical_crossentropy"

    def __init__(self, from_logits: bool = False, ignore_class: int | None = None) -> None:
        self.from_logits = from_logits
        self.ignore_class = ignore_class

    def per_example(self, y_true: torch.Tensor, y_pred: torch.Tensor) -> torch.Tensor:
        if self.from_logits and y_pred.is_cuda and y_pred.dtype == torch.bfloat16:
            # bf16 logits on a GPU (wide vocabularies): one fused forward and one in-place backward kernel
            from ..ops import xent as _xent

            if _xent.supported(y_pred):
                ig = self.ignore_class if self.ignore_class is not None else -(1 << 62)
                return _xent.sparse_softmax_xent(y_pred, y_true, ig)
        yp = y_pred.float()
        if yp.dim() > 2:
            yp = yp.reshape(-1, yp.shape[-1])
        y = y_true.reshape(-1).long()
        logp = F.log_softmax(yp, dim=-1) if self.from_logits else torch.log(yp.clamp_min(1e-7))
        if self.ignore_class is not None:
            return F.nll_loss(logp, y, reduction="none", ignore_index=self.ignore_class)
        return F.nll_loss(logp, y, reduction="none")


class CategoricalCrossentropy(Loss):
    name = "categorical_crossentropy"

    def __init__(self, from_logits: bool = False) -> None:
        self.from_logits = from_logits

    def per_example(self, y_true: torch.Tensor, y_pred: torch.Tensor) -> torch.Tensor:
        yp = y_pred.float()
        logp = F.log_softmax(yp, dim=-1) if self.from_logits else torch.log(yp.clamp_min(1e-7))
        return -(y_true.float() * logp).sum(-1)


class MeanSquaredError(Loss):
    name = "mean_squared_error"

    def per_example(self, y_true: torch.Tensor, y_pred: torch.Tensor) -> torch.Tensor:
        d = (y_pred.float() - y_true.float().reshape(y_pred.shape)) ** 2
        return d.reshape(d.shape[0], -1).mean(-1)


class MeanAbsoluteError(Loss):
    name = "mean_absolute_error"

    def per_example(self, y_true: torch.Tensor, y_pred: torch.Tensor) -> torch.Tensor:
        d = (y_pred.float() - y_true.float().reshape(y_pred.shape)).abs()
        return d.reshape(d.shape[0], -1).mean(-1)


class BinaryCrossentropy(Loss):
    name = "binary_crossentropy"

    def __init__(self, from_logits: bool = False) -> None:
        self.from_logits = from_logits

    def per_example(self, y_true: torch.Tensor, y_pred: torch.Tensor) -> torch.Tensor:
        yp = y_pred.float()
        yt = y_true.float().reshape(yp.shape)
        if self.from_logits:
            per = F.binary_cross_entropy_with_logits(yp, yt, reduction="none")
        else:
            per = F.binary_cross_entropy(yp.clamp(1e-7, 1 - 1e-7), yt, reduction="none")
        return per.reshape(per.shape[0], -1).mean(-1)


_BY_NAME = {
    "sparse_categorical_crossentropy": lambda: SparseCategoricalCrossentropy(),
    "categorical_crossentropy": lambda: CategoricalCrossentropy(),
    "mse": lambda: MeanSquaredError(),
    "mean_squared_error": lambda: MeanSquaredError(),
    "mae": lambda: MeanAbsoluteError(),
    "mean_absolute_error": lambda: MeanAbsoluteError(),
    "binary_crossentropy": lambda: BinaryCrossentropy(),
}


def get(x) -> Loss:
    if isinstance(x, Loss):
        return x
    if isinstance(x, str) and x in _BY_NAME:
        return _BY_NAME[x]()
    raise ValueError(f"unknown loss {x!r}")
