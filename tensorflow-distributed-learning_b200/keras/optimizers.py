"""Optimizers (ref: /root/reference/tf_dist_example.py:51 SGD(learning_rate=0.001)).

They act on the flat arenas (parallel/values.py): one fused update over the whole parameter
buffer instead of one `ResourceApplyGradientDescent` launch per variable. On the native GPU path
the SGD / momentum update is not even a separate kernel - it is fused into the gradient all-reduce
(ops/allreduce.py); `apply_flat` below is the portable torch version with identical arithmetic."""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Union

import torch

Schedule = Callable[[int], float]


class Optimizer:
    name = "optimizer"

    def __init__(self, learning_rate: Union[float, Schedule] = 0.01) -> None:
        self._lr = learning_rate
        self.iterations = 0
        self._slots: Dict[str, List[torch.Tensor]] = {}

    @property
    def learning_rate(self) -> float:
        return float(self._lr(self.iterations)) if callable(self._lr) else float(self._lr)

    @learning_rate.setter
    def learning_rate(self, v: Union[float, Schedule]) -> None:
        self._lr = v

    lr = learning_rate

    def slot_names(self) -> List[str]:
        return []

    def bind(self, params: List[torch.Tensor]) -> None:
        """Allocate slot arenas (one per local replica) shaped like the parameter arenas."""
        self._slots = {n: [torch.zeros_like(p) for p in params] for n in self.slot_names()}

    def slots(self, name: str) -> List[torch.Tensor]:
        return self._slots[name]

    def apply_flat(self, params: torch.Tensor, grads: torch.Tensor, replica: int, mask: Optional[torch.Tensor]) -> None:
        raise NotImplementedError

    def step_done(self) -> None:
        self.iterations += 1

    def get_config(self) -> Dict[str, float]:
        return {"name": self.name, "learning_rate": self.learning_rate}


class SGD(Optimizer):
    """w -= lr * g, or with momentum m: v = m*v - lr*g; w += v (Keras convention; nesterov:
    w += m*v - lr*g)."""

    name = "SGD"

    def __init__(self, learning_rate: Union[float, Schedule] = 0.01, momentum: float = 0.0,
                 nesterov: bool = False) -> None:
        super().__init__(learning_rate)
        self.momentum, self.nesterov = float(momentum), bool(nesterov)

    def slot_names(self) -> List[str]:
        return ["momentum"] if self.momentum else []

    def apply_flat(self, params, grads, replica, mask):
        lr = self.learning_rate
        g = grads if mask is None else grads * mask
        if not self.momentum:
            params.add_(g, alpha=-lr)
            return
        v = self._slots["momentum"][replica]
        v.mul_(self.momentum).add_(g, alpha=-lr)
        if self.nesterov:
            params.add_(v, alpha=self.momentum).add_(g, alpha=-lr)
        else:
            params.add_(v)

    def get_config(self):
        return {**super().get_config(), "momentum": self.momentum, "nesterov": self.nesterov}


class Adam(Optimizer):
    name = "Adam"

    def __init__(self, learning_rate: Union[float, Schedule] = 0.001, beta_1: float = 0.9, beta_2: float = 0.999,
                 epsilon: float = 1e-7, weight_decay: float = 0.0) -> None:
        super().__init__(learning_rate)
        self.beta_1, self.beta_2, self.epsilon, self.weight_decay = beta_1, beta_2, epsilon, weight_decay

    def get_config(self):
        return {**super().get_config(), "beta_1": self.beta_1, "beta_2": self.beta_2, "epsilon": self.epsilon,
                "weight_decay": self.weight_decay}

    def slot_names(self) -> List[str]:
        return ["m", "v"]

    def apply_flat(self, params, grads, replica, mask):
        t = self.iterations + 1
        g = grads if mask is None else grads * mask
        m, v = self._slots["m"][replica], self._slots["v"][replica]
        m.mul_(self.beta_1).add_(g, alpha=1 - self.beta_1)
        v.mul_(self.beta_2).addcmul_(g, g, value=1 - self.beta_2)
        lr_t = self.learning_rate * (1 - self.beta_2 ** t) ** 0.5 / (1 - self.beta_1 ** t)
        upd = m / (v.sqrt() + self.epsilon)
        if self.weight_decay:
            upd = upd + self.weight_decay * (params if mask is None else params * mask)
        params.add_(upd, alpha=-lr_t)


class RMSprop(Optimizer):
    """tf.keras.optimizers.RMSprop: v = rho*v + (1-rho)*g^2; w -= lr * g / (sqrt(v) + eps), optionally through a momentum
    buffer. Runs as tensor ops on the flat arenas behind the plain-sum all-reduce (no fused kernel)."""

    name = "RMSprop"

    def __init__(self, learning_rate: Union[float, Schedule] = 0.001, rho: float = 0.9, momentum: float = 0.0,
                 epsilon: float = 1e-7) -> None:
        super().__init__(learning_rate)
        self.rho, self.momentum, self.epsilon = float(rho), float(momentum), float(epsilon)

    def slot_names(self) -> List[str]:
        return ["rms", "momentum"] if self.momentum else ["rms"]

    def apply_flat(self, params, grads, replica, mask):
        g = grads if mask is None else grads * mask
        v = self._slots["rms"][replica]
        v.mul_(self.rho).addcmul_(g, g, value=1 - self.rho)
        upd = g / (v.sqrt() + self.epsilon)
        if self.momentum:
            mom = self._slots["momentum"][replica]
            mom.mul_(self.momentum).add_(upd, alpha=self.learning_rate)
            params.sub_(mom)
        else:
            params.add_(upd, alpha=-self.learning_rate)

    def get_config(self):
        return {**super().get_config(), "rho": self.rho, "momentum": self.momentum, "epsilon": self.epsilon}


_BY_NAME = {"sgd": SGD, "adam": Adam, "rmsprop": RMSprop}


def from_config(config: Dict[str, float]) -> Optimizer:
    """Rebuild an optimizer from get_config() (a schedule given as a callable is saved as its current value)."""
    cfg = dict(config)
    cls = _BY_NAME[str(cfg.pop("name")).lower()]
    return cls(**cfg)


def get(x) -> Optimizer:
    if isinstance(x, Optimizer):
        return x
    if isinstance(x, str):
        if x.lower() not in _BY_NAME:
            raise ValueError(f"unknown optimizer {x!r}; known: {sorted(_BY_NAME)}")
        return _BY_NAME[x.lower()]()
    raise ValueError(f"unknown optimizer {x!r}")
