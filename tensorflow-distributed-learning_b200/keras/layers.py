"""Keras-style layers (ref: /root/reference/tf_dist_example.py:40-48).

Shapes follow Keras: activations are NHWC, Conv2D kernels report HWIO, Dense kernels (in, out),
Flatten orders features (h, w, c). *Storage* differs (see parallel/values.py): conv kernels are held
OHWI and dense kernels [out, in] so that the forward GEMM operands are K-major for tcgen05 and so
that torch's channels_last conv can use them without a copy; `get_weights()` / checkpoints convert
back to the Keras layout.

`call()` dispatches per layer: on CUDA replicas (CallContext.native) Conv2D, MaxPooling2D, GlobalAveragePooling2D, Dense,
BatchNormalization and LayerNormalization run on this library's sm_100a kernels (ops/conv.py, ops/pool.py, ops/gemm.py,
ops/batchnorm.py, ops/layernorm.py); the torch implementation below each dispatch is the portable path (CPU tests, the
independent reference of the GPU tests, shapes a kernel does not take). The reference CNN's exact stack is executed by the
native step program instead (keras/native_engine.py), which never calls these functions.
"""
from __future__ import annotations

import functools
import inspect
import math
import os
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple, Union

import numpy as np
import torch
import torch.nn.functional as F

from ..parallel.values import MirroredVariable, VarSpec

_uids: Dict[str, int] = {}


def _unique_name(prefix: str) -> str:
    n = _uids.get(prefix, 0)
    _uids[prefix] = n + 1
    return prefix if n == 0 else f"{prefix}_{n}"


def clear_session() -> None:
    """keras.backend.clear_session(): restart the auto-naming counters."""
    _uids.clear()


def _pair(v: Union[int, Sequence[int]]) -> Tuple[int, int]:
    return (int(v), int(v)) if isinstance(v, int) else (int(v[0]), int(v[1]))


# ---- initialisers -------------------------------------------------------------------------------

_init_rng = np.random.default_rng()


def set_init_seed(seed: Optional[int]) -> None:
    global _init_rng
    _init_rng = np.random.default_rng(seed)


def glorot_uniform(shape: Tuple[int, ...]) -> np.ndarray:
    """Keras default kernel initialiser. fans for a conv kernel HWIO: in = h*w*i, out = h*w*o."""
    if len(shape) == 2:
        fan_in, fan_out = shape
    else:
        rf = int(np.prod(shape[:-2]))
        fan_in, fan_out = shape[-2] * rf, shape[-1] * rf
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    return _init_rng.uniform(-lim, lim, size=shape).astype(np.float32)


def he_normal(shape: Tuple[int, ...]) -> np.ndarray:
    rf = int(np.prod(shape[:-2])) if len(shape) > 2 else 1
    fan_in = shape[-2] * rf if len(shape) >= 2 else shape[0]
    return (_init_rng.standard_normal(size=shape) * math.sqrt(2.0 / fan_in)).astype(np.float32)


def truncated_normal(stddev: float = 0.02) -> Callable[[Tuple[int, ...]], np.ndarray]:
    def init(shape: Tuple[int, ...]) -> np.ndarray:
        a = _init_rng.standard_normal(size=shape)
        bad = np.abs(a) > 2
        while bad.any():
            a[bad] = _init_rng.standard_normal(size=int(bad.sum()))
            bad = np.abs(a) > 2
        return (a * stddev).astype(np.float32)

    return init


def zeros(shape: Tuple[int, ...]) -> np.ndarray:
    return np.zeros(shape, np.float32)


def ones(shape: Tuple[int, ...]) -> np.ndarray:
    return np.ones(shape, np.float32)


_INITS = {"glorot_uniform": glorot_uniform, "he_normal": he_normal, "zeros": zeros, "ones": ones}


def get_initializer(x: Union[str, Callable]) -> Callable[[Tuple[int, ...]], np.ndarray]:
    return _INITS[x] if isinstance(x, str) else x


# ---- activations --------------------------------------------------------------------------------

def _gelu_tanh(x: torch.Tensor) -> torch.Tensor:
    return F.gelu(x, approximate="tanh")


_ACTS: Dict[Optional[str], Optional[Callable[[torch.Tensor], torch.Tensor]]] = {
    None: None, "linear": None, "relu": torch.relu, "gelu": _gelu_tanh, "tanh": torch.tanh,
    "sigmoid": torch.sigmoid, "softmax": lambda x: torch.softmax(x, dim=-1),
}


class CallContext:
    """Per-call state handed to layers: which replica, training or not, and how to fetch the
    autograd leaf that aliases a variable in that replica's arena."""

    def __init__(self, replica: int, training: bool, param: Callable[[MirroredVariable], torch.Tensor],
                 compute_dtype: torch.dtype = torch.float32, native: Optional[bool] = None) -> None:
        self.replica, self.training, self.param, self.compute_dtype = replica, training, param, compute_dtype
        # native: run Conv2D / Dense / pooling on the sm_100a kernels (bf16 tensor-core operands, fp32 accumulation,
        # fp32 master weights - the precision contract of the native step program) whenever the shapes allow it.
        # Default: under the mixed_bfloat16 policy only; the training engine turns it on for CUDA replicas.
        self.native = (compute_dtype == torch.bfloat16) if native is None else bool(native)
        # Optional services of the training engine (None elsewhere): param_lp(var) -> a current bf16 copy of the variable
        # (the shadow arena the fused all-reduce refreshes), grad_sink(var) -> (fp32 gradient view in the arena, callback
        # to run once the gradient is complete) so a kernel can write the weight gradient in place.
        self.param_lp: Optional[Callable[[MirroredVariable], Optional[torch.Tensor]]] = None
        self.grad_sink: Optional[Callable[[MirroredVariable], Any]] = None
        # in-place updates of non-trainable variables (BatchNorm statistics) that must wait until autograd
        # has finished: every variable is a view of one arena, so they share a version counter
        self.deferred: List[Callable[[], None]] = []

    def lp(self, var: MirroredVariable) -> Optional[torch.Tensor]:
        return self.param_lp(var) if self.param_lp is not None else None

    def sink(self, var: MirroredVariable) -> Any:
        return self.grad_sink(var) if (self.grad_sink is not None and self.training) else None


def capture_init_config(cls: type) -> None:
    """Wrap cls.__init__ so that the arguments of the OUTERMOST constructor call are remembered on the instance
    (`_init_config`): that is what get_config() returns and from_config() feeds back (tf.keras layer / model configs)."""
    init = cls.__dict__.get("__init__")
    if init is None or getattr(init, "_captures_config", False):
        return
    sig = inspect.signature(init)

    @functools.wraps(init)
    def wrapped(self, *a: Any, **k: Any) -> None:
        if not hasattr(self, "_init_config"):
            cfg: Dict[str, Any] = {}
            try:
                bound = sig.bind(self, *a, **k)
                for name, val in list(bound.arguments.items())[1:]:
                    kind = sig.parameters[name].kind
                    if kind is inspect.Parameter.VAR_KEYWORD:
                        cfg.update(val)
                    elif kind is not inspect.Parameter.VAR_POSITIONAL:
                        cfg[name] = val
            except TypeError:
                pass  # the real call below raises the proper error
            self._init_config = cfg
        init(self, *a, **k)

    wrapped._captures_config = True  # type: ignore[attr-defined]
    cls.__init__ = wrapped  # type: ignore[assignment]


def jsonable_config(cfg: Dict[str, Any]) -> Dict[str, Any]:
    """Constructor arguments as JSON values: tuples become lists, callables (custom initialisers) are left out - a
    restored model takes its values from the saved variables, not from an initialiser."""
    out: Dict[str, Any] = {}
    for k, v in cfg.items():
        if callable(v):
            continue
        if isinstance(v, (tuple, list)):
            v = [int(e) if isinstance(e, (int, np.integer)) else e for e in v]
        elif isinstance(v, np.integer):
            v = int(v)
        elif isinstance(v, np.floating):
            v = float(v)
        out[k] = v
    return out


class Layer:
    """Base layer: lazily built from the input shape (excluding the batch dimension)."""

    def __init_subclass__(cls, **kw: Any) -> None:
        super().__init_subclass__(**kw)
        capture_init_config(cls)

    def __init__(self, name: Optional[str] = None, input_shape: Optional[Sequence[int]] = None,
                 trainable: bool = True, **kw: Any) -> None:
        cls = type(self).__name__
        snake = "".join("_" + c.lower() if c.isupper() and i and not cls[i - 1].isdigit() and not cls[i - 1].isupper()
                        else c.lower() for i, c in enumerate(cls))
        self.name = name or _unique_name(snake)
        self.trainable = trainable
        self.built = False
        self.input_shape_arg = tuple(input_shape) if input_shape is not None else None
        self.input_shape: Optional[Tuple[int, ...]] = None
        self.output_shape: Optional[Tuple[int, ...]] = None
        self._weights: List[MirroredVariable] = []

    # -- weights ------------------------------------------------------------------------------
    def add_weight(self, name: str, shape: Sequence[int], initializer: Union[str, Callable] = "glorot_uniform",
                   trainable: bool = True, to_storage: Optional[Tuple[int, ...]] = None) -> MirroredVariable:
        shape = tuple(int(s) for s in shape)
        st_shape = shape if to_storage is None else tuple(shape[p] for p in to_storage)
        spec = VarSpec(f"{self.name}/{name}:0", shape, st_shape, to_storage, trainable=trainable and self.trainable)
        v = MirroredVariable(spec, get_initializer(initializer)(shape))
        self._weights.append(v)
        return v

    @property
    def weights(self) -> List[MirroredVariable]:
        return list(self._weights)

    @property
    def trainable_weights(self) -> List[MirroredVariable]:
        return [w for w in self._weights if w.trainable]

    def count_params(self) -> int:
        return sum(int(np.prod(w.shape)) for w in self._weights)

    def get_weights(self) -> List[np.ndarray]:
        return [w.numpy() for w in self._weights]

    def set_weights(self, ws: Sequence[np.ndarray]) -> None:
        if len(ws) != len(self._weights):
            raise ValueError(f"{self.name}: expected {len(self._weights)} arrays, got {len(ws)}")
        for w, a in zip(self._weights, ws):
            w.assign(a)

    # -- build / call -------------------------------------------------------------------------
    def build(self, input_shape: Tuple[int, ...]) -> None:
        pass

    def compute_output_shape(self, input_shape: Tuple[int, ...]) -> Tuple[int, ...]:
        return input_shape

    def _maybe_build(self, input_shape: Tuple[int, ...]) -> Tuple[int, ...]:
        input_shape = tuple(int(s) for s in input_shape)
        if not self.built:
            self.input_shape = input_shape
            self.build(input_shape)
            self.output_shape = self.compute_output_shape(input_shape)
            self.built = True
        elif self.input_shape != input_shape:
            raise ValueError(f"{self.name} was built for input shape {self.input_shape}, got {input_shape}")
        assert self.output_shape is not None
        return self.output_shape

    def call(self, x: torch.Tensor, ctx: CallContext) -> torch.Tensor:
        return x

    def get_config(self) -> Dict[str, Any]:
        """The constructor arguments (tf.keras.layers.Layer.get_config); `from_config` rebuilds an unbuilt layer."""
        cfg = jsonable_config(getattr(self, "_init_config", {}))
        cfg["name"] = self.name
        return cfg

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "Layer":
        return cls(**config)

    def __repr__(self) -> str:
        return f"<{type(self).__name__} {self.name}>"


class InputLayer(Layer):
    def __init__(self, input_shape: Sequence[int], **kw: Any) -> None:
        super().__init__(input_shape=input_shape, **kw)


class Conv2D(Layer):
    """tf.keras.layers.Conv2D (tf_dist_example.py:41,43): NHWC, kernel HWIO, stride 1 / 'valid' by default."""

    def __init__(self, filters: int, kernel_size: Union[int, Sequence[int]], strides: Union[int, Sequence[int]] = 1,
                 padding: str = "valid", activation: Optional[str] = None, use_bias: bool = True,
                 kernel_initializer: Union[str, Callable] = "glorot_uniform",
                 bias_initializer: Union[str, Callable] = "zeros", **kw: Any) -> None:
        super().__init__(**kw)
        self.filters, self.kernel_size, self.strides = int(filters), _pair(kernel_size), _pair(strides)
        self.padding = padding.lower()
        if self.padding not in ("valid", "same"):
            raise ValueError("padding must be 'valid' or 'same'")
        if activation not in _ACTS:
            raise ValueError(f"unknown activation {activation!r}")
        self.activation, self.use_bias = activation, use_bias
        self._kinit, self._binit = kernel_initializer, bias_initializer
        self.kernel: Optional[MirroredVariable] = None
        self.bias: Optional[MirroredVariable] = None

    def build(self, input_shape: Tuple[int, ...]) -> None:
        if len(input_shape) != 3:
            raise ValueError(f"{self.name}: Conv2D expects (H, W, C) inputs, got {input_shape}")
        kh, kw = self.kernel_size
        # Keras HWIO -> storage OHWI
        self.kernel = self.add_weight("kernel", (kh, kw, input_shape[2], self.filters), self._kinit,
                                      to_storage=(3, 0, 1, 2))
        if self.use_bias:
            self.bias = self.add_weight("bias", (self.filters,), self._binit)

    def _pads(self, size: int, k: int, s: int) -> Tuple[int, int]:
        if self.padding == "valid":
            return 0, 0
        out = -(-size // s)
        tot = max((out - 1) * s + k - size, 0)
        return tot // 2, tot - tot // 2

    def compute_output_shape(self, input_shape: Tuple[int, ...]) -> Tuple[int, ...]:
        h, w, _ = input_shape
        (kh, kw), (sh, sw) = self.kernel_size, self.strides
        if self.padding == "valid":
            return ((h - kh) // sh + 1, (w - kw) // sw + 1, self.filters)
        return (-(-h // sh), -(-w // sw), self.filters)

    def call(self, x: torch.Tensor, ctx: CallContext, add_bias: bool = True, defer_relu: bool = False) -> torch.Tensor:
        """``add_bias=False``: leave the bias to a following BatchNormalization(in_bias=...), which absorbs it for
        free (no extra pass over the activations, no bias-gradient reduction). ``defer_relu=True`` (library-kernel path
        only): return the pre-activation; the caller applies the ReLU after the max-pool that follows (see
        Sequential._forward)."""
        assert self.kernel is not None
        b = ctx.param(self.bias) if (self.bias is not None and add_bias) else None
        fused_act = self.activation if self.activation in (None, "linear", "relu") else None
        post = None if fused_act == self.activation or self.activation == "linear" else _ACTS[self.activation]
        if (self.kernel_size == (1, 1) and self.strides == (1, 1) and x.is_cuda and ctx.native
                and x.shape[-1] % 8 == 0 and self.filters % 8 == 0 and os.environ.get("TFDL_CONV1X1", "gemm") == "gemm"):
            # a 1x1 stride-1 convolution over NHWC activations IS a GEMM over [N*H*W, Cin]: forward, dgrad and wgrad
            # run on the TMA + tcgen05 kernels (OHWI storage of a 1x1 kernel is the [out][in] matrix itself)
            from ..ops.gemm import linear

            w2 = ctx.param(self.kernel).reshape(self.filters, x.shape[-1])
            lp = ctx.lp(self.kernel)
            y = linear(x, w2, b, self.activation if self.activation in (None, "relu", "gelu") else None,
                       None if lp is None else lp.reshape(self.filters, x.shape[-1]), ctx.sink(self.kernel))
            if self.activation not in (None, "relu", "gelu"):
                y = _ACTS[self.activation](y)
            return y
        if x.is_cuda and ctx.native and self.filters % 8 == 0 and os.environ.get("TFDL_CONV", "native") == "native":
            # general Conv2D on the implicit-GEMM kernels (TMA im2col + tcgen05): any filter size / stride / padding
            from ..ops import conv as _conv

            pt, pb = self._pads(x.shape[1], self.kernel_size[0], self.strides[0])
            pl, pr = self._pads(x.shape[2], self.kernel_size[1], self.strides[1])
            pads = (pt, pl, pb, pr)
            y = None
            if x.shape[-1] % 8 == 0:
                y = _conv.conv2d(x, ctx.param(self.kernel), b, self.strides, pads, None if fused_act == "linear" else fused_act,
                                 ctx.lp(self.kernel), ctx.sink(self.kernel))
            elif not x.requires_grad and self.kernel_size[0] * self.kernel_size[1] * x.shape[-1] <= 512:
                y = _conv.conv2d_small_cin(x, ctx.param(self.kernel), b, self.strides, pads,
                                           None if fused_act == "linear" else fused_act)
            if y is not None:
                return post(y) if post else y
        w = ctx.param(self.kernel).permute(0, 3, 1, 2)  # OHWI storage viewed as OIHW (channels_last strides)
        xin = x.permute(0, 3, 1, 2)
        pad = (0, 0)
        if self.padding == "same":
            pt, pb = self._pads(x.shape[1], self.kernel_size[0], self.strides[0])
            pl, pr = self._pads(x.shape[2], self.kernel_size[1], self.strides[1])
            pad = (min(pt, pb), min(pl, pr))  # the symmetric part is the convolution's own padding: no copy
            pt, pb, pl, pr = pt - pad[0], pb - pad[0], pl - pad[1], pr - pad[1]
            if pt or pb or pl or pr:
                xin = F.pad(xin, (pl, pr, pt, pb))
        if ctx.compute_dtype != torch.float32:
            xin, w = xin.to(ctx.compute_dtype), w.to(ctx.compute_dtype)
            b = b.to(ctx.compute_dtype) if b is not None else None
        elif xin.dtype != w.dtype:  # bf16 activations from a native layer upstream
            xin = xin.to(w.dtype)
        y = F.conv2d(xin, w, b, stride=self.strides, padding=pad).permute(0, 2, 3, 1)
        act = None if (defer_relu and self.activation == "relu") else _ACTS[self.activation]
        return act(y) if act else y


class MaxPooling2D(Layer):
    """tf.keras.layers.MaxPooling2D (tf_dist_example.py:42,44): 2x2, stride 2, 'valid' by default."""

    def __init__(self, pool_size: Union[int, Sequence[int]] = 2, strides: Union[None, int, Sequence[int]] = None,
                 padding: str = "valid", **kw: Any) -> None:
        super().__init__(**kw)
        self.pool_size = _pair(pool_size)
        self.strides = _pair(strides) if strides is not None else self.pool_size
        self.padding = padding.lower()

    def compute_output_shape(self, input_shape: Tuple[int, ...]) -> Tuple[int, ...]:
        h, w, c = input_shape
        if self.padding == "same":
            return (-(-h // self.strides[0]), -(-w // self.strides[1]), c)
        return ((h - self.pool_size[0]) // self.strides[0] + 1, (w - self.pool_size[1]) // self.strides[1] + 1, c)

    def call(self, x: torch.Tensor, ctx: CallContext) -> torch.Tensor:
        if type(self) is MaxPooling2D and ctx.native:
            from ..ops import pool as _pool

            if _pool.supported(x) and self.pool_size[0] * self.pool_size[1] <= 255:
                oh, ow, _ = self.compute_output_shape(tuple(x.shape[1:]))
                pt = pl = 0
                if self.padding == "same":
                    pt = max((oh - 1) * self.strides[0] + self.pool_size[0] - x.shape[1], 0) // 2
                    pl = max((ow - 1) * self.strides[1] + self.pool_size[1] - x.shape[2], 0) // 2
                return _pool.max_pool(x, self.pool_size, self.strides, (pt, pl), (oh, ow))
        xin = x.permute(0, 3, 1, 2)
        if self.padding == "same":
            oh, ow, _ = self.compute_output_shape(tuple(x.shape[1:]))
            ph = max((oh - 1) * self.strides[0] + self.pool_size[0] - x.shape[1], 0)
            pw = max((ow - 1) * self.strides[1] + self.pool_size[1] - x.shape[2], 0)
            xin = F.pad(xin, (pw // 2, pw - pw // 2, ph // 2, ph - ph // 2), value=float("-inf"))
        return F.max_pool2d(xin, self.pool_size, self.strides).permute(0, 2, 3, 1)


class AveragePooling2D(MaxPooling2D):
    def call(self, x: torch.Tensor, ctx: CallContext) -> torch.Tensor:
        return F.avg_pool2d(x.permute(0, 3, 1, 2), self.pool_size, self.strides).permute(0, 2, 3, 1)


class GlobalAveragePooling2D(Layer):
    def compute_output_shape(self, input_shape: Tuple[int, ...]) -> Tuple[int, ...]:
        return (input_shape[2],)

    def call(self, x: torch.Tensor, ctx: CallContext) -> torch.Tensor:
        if ctx.native:
            from ..ops import pool as _pool

            if _pool.supported(x) and (x.shape[-1] <= 2048 or x.shape[-1] % 2048 == 0):
                return _pool.global_avg_pool(x)
        return x.mean(dim=(1, 2))


class Flatten(Layer):
    """tf.keras.layers.Flatten (tf_dist_example.py:45): NHWC order (h, w, c)."""

    def compute_output_shape(self, input_shape: Tuple[int, ...]) -> Tuple[int, ...]:
        return (int(np.prod(input_shape)),)

    def call(self, x: torch.Tensor, ctx: CallContext) -> torch.Tensor:
        return x.reshape(x.shape[0], -1)


class Dense(Layer):
    """tf.keras.layers.Dense (tf_dist_example.py:46-47): kernel (in, out) in Keras layout."""

    def __init__(self, units: int, activation: Optional[str] = None, use_bias: bool = True,
                 kernel_initializer: Union[str, Callable] = "glorot_uniform",
                 bias_initializer: Union[str, Callable] = "zeros", **kw: Any) -> None:
        super().__init__(**kw)
        if activation not in _ACTS:
            raise ValueError(f"unknown activation {activation!r}")
        self.units, self.activation, self.use_bias = int(units), activation, use_bias
        self._kinit, self._binit = kernel_initializer, bias_initializer
        self.kernel: Optional[MirroredVariable] = None
        self.bias: Optional[MirroredVariable] = None

    def build(self, input_shape: Tuple[int, ...]) -> None:
        self.kernel = self.add_weight("kernel", (input_shape[-1], self.units), self._kinit, to_storage=(1, 0))
        if self.use_bias:
            self.bias = self.add_weight("bias", (self.units,), self._binit)

    def compute_output_shape(self, input_shape: Tuple[int, ...]) -> Tuple[int, ...]:
        return tuple(input_shape[:-1]) + (self.units,)

    def call(self, x: torch.Tensor, ctx: CallContext) -> torch.Tensor:
        assert self.kernel is not None
        w = ctx.param(self.kernel)
        b = ctx.param(self.bias) if self.bias is not None else None
        if (x.is_cuda and ctx.native and self.activation in (None, "linear", "relu", "gelu")
                and x.shape[-1] % 8 == 0 and self.units % 8 == 0 and (x.numel() // x.shape[-1]) % 8 == 0):
            # mixed_bfloat16 on a B200: forward, input-gradient and weight-gradient GEMMs on the persistent
            # TMA + tcgen05 kernel (ops/gemm.py), fp32 master weights and fp32 weight gradients
            from ..ops.gemm import linear as _tc_linear

            return _tc_linear(x, w, b, None if self.activation == "linear" else self.activation, ctx.lp(self.kernel),
                              ctx.sink(self.kernel))
        if ctx.compute_dtype != torch.float32:
            x, w = x.to(ctx.compute_dtype), w.to(ctx.compute_dtype)
            b = b.to(ctx.compute_dtype) if b is not None else None
        elif x.dtype != w.dtype:  # bf16 activations from a native layer upstream
            x = x.to(w.dtype)
        y = F.linear(x, w, b)
        act = _ACTS[self.activation]
        return act(y) if act else y


class Activation(Layer):
    def __init__(self, activation: str, **kw: Any) -> None:
        super().__init__(**kw)
        if activation not in _ACTS:
            raise ValueError(f"unknown activation {activation!r}")
        self.activation = activation

    def call(self, x: torch.Tensor, ctx: CallContext) -> torch.Tensor:
        act = _ACTS[self.activation]
        return act(x) if act else x


class ReLU(Activation):
    def __init__(self, **kw: Any) -> None:
        super().__init__("relu", **kw)


class Dropout(Layer):
    def __init__(self, rate: float, **kw: Any) -> None:
        super().__init__(**kw)
        self.rate = float(rate)

    def call(self, x: torch.Tensor, ctx: CallContext) -> torch.Tensor:
        return F.dropout(x, self.rate, training=ctx.training)


class BatchNormalization(Layer):
    """Per-replica batch statistics (Keras default under MirroredStrategy: not synchronised).
    gamma/beta are mirrored trainable variables; moving statistics are non-trainable."""

    def __init__(self, momentum: float = 0.99, epsilon: float = 1e-3, **kw: Any) -> None:
        super().__init__(**kw)
        self.momentum, self.epsilon = momentum, epsilon

    def build(self, input_shape: Tuple[int, ...]) -> None:
        c = input_shape[-1]
        self.gamma = self.add_weight("gamma", (c,), "ones")
        self.beta = self.add_weight("beta", (c,), "zeros")
        self.moving_mean = self.add_weight("moving_mean", (c,), "zeros", trainable=False)
        self.moving_variance = self.add_weight("moving_variance", (c,), "ones", trainable=False)

    def call(self, x: torch.Tensor, ctx: CallContext, relu: bool = False, residual: Optional[torch.Tensor] = None,
             in_bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        """y = BN(x + in_bias); with ``relu`` / ``residual`` the bottleneck tail relu(BN(x) + residual) in the same
        pass. ``in_bias``: the bias of the producing Conv2D when it was called with add_bias=False.
        NHWC bf16 CUDA activations run on the native streaming kernels (ops/batchnorm.py)."""
        from ..ops import batchnorm as _bn

        g, b = ctx.param(self.gamma), ctx.param(self.beta)
        mm, mv = ctx.param(self.moving_mean), ctx.param(self.moving_variance)
        if _bn.supported(x) and (residual is None or (residual.dtype == x.dtype and residual.shape == x.shape)):
            if ctx.training:
                # the kernel writes the running statistics through their arena pointers (no autograd version bump)
                return _bn.batch_norm_train(x, g, b, mm.detach(), mv.detach(), 1.0 - self.momentum, self.epsilon, relu,
                                            residual, in_bias)
            return _bn.batch_norm_infer(x, g, b, mm, mv, self.epsilon, relu, residual, in_bias)
        if in_bias is not None:
            x = x + in_bias.to(x.dtype)
        y = self._call_torch(x, ctx, g, b, mm, mv)
        if residual is not None:
            y = y + residual
        return torch.relu(y) if relu else y

    def _call_torch(self, x: torch.Tensor, ctx: CallContext, g: torch.Tensor, b: torch.Tensor, mm: torch.Tensor,
                    mv: torch.Tensor) -> torch.Tensor:
        if x.dim() == 4:
            # NHWC activations viewed as channels_last NCHW: one fused (cuDNN) batch-norm kernel. The running
            # statistics are updated on private copies and written back after autograd has finished, because
            # every variable is a view of one arena and in-place writes would invalidate saved tensors.
            xin = x.permute(0, 3, 1, 2)
            if ctx.training:
                rm, rv = mm.detach().clone(), mv.detach().clone()
                y = F.batch_norm(xin, rm, rv, g, b, True, 1.0 - self.momentum, self.epsilon)

                def _update(mm=mm, mv=mv, rm=rm, rv=rv) -> None:
                    with torch.no_grad():
                        mm.copy_(rm)
                        mv.copy_(rv)

                ctx.deferred.append(_update)
            else:
                y = F.batch_norm(xin, mm, mv, g, b, False, 0.0, self.epsilon)
            return y.permute(0, 2, 3, 1)
        dims = tuple(range(x.dim() - 1))
        xf = x.float()
        if ctx.training:
            mean, var = xf.mean(dims), xf.var(dims, unbiased=False)
            m_d, v_d, mom = mean.detach(), var.detach(), self.momentum

            def _update2(mm=mm, mv=mv, m_d=m_d, v_d=v_d, mom=mom) -> None:
                with torch.no_grad():
                    mm.mul_(mom).add_(m_d, alpha=1 - mom)
                    mv.mul_(mom).add_(v_d, alpha=1 - mom)

            ctx.deferred.append(_update2)
        else:
            mean, var = mm, mv
        y = (xf - mean) * torch.rsqrt(var + self.epsilon) * g + b
        return y.to(x.dtype)


class LayerNormalization(Layer):
    def __init__(self, epsilon: float = 1e-12, **kw: Any) -> None:
        super().__init__(**kw)
        self.epsilon = epsilon

    def build(self, input_shape: Tuple[int, ...]) -> None:
        self.gamma = self.add_weight("gamma", (input_shape[-1],), "ones")
        self.beta = self.add_weight("beta", (input_shape[-1],), "zeros")

    def call(self, x: torch.Tensor, ctx: CallContext, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        """LayerNorm(x + residual) over the last axis. bf16 CUDA activations run on the native streaming kernel
        (ops/layernorm.py: residual add fused, fp32 statistics); everything else goes through torch in fp32."""
        from ..ops import layernorm as _ln

        if _ln.supported(x) and (residual is None or (residual.dtype == x.dtype and residual.shape == x.shape)):
            return _ln.layer_norm(x, ctx.param(self.gamma), ctx.param(self.beta), self.epsilon, residual)
        if residual is not None:
            x = x + residual
        return F.layer_norm(x.float(), (x.shape[-1],), ctx.param(self.gamma), ctx.param(self.beta),
                            self.epsilon).to(x.dtype)


class Embedding(Layer):
    def __init__(self, input_dim: int, output_dim: int, embeddings_initializer: Union[str, Callable] = None,
                 **kw: Any) -> None:
        super().__init__(**kw)
        self.input_dim, self.output_dim = int(input_dim), int(output_dim)
        self._init = embeddings_initializer or truncated_normal(0.02)

    def build(self, input_shape: Tuple[int, ...]) -> None:
        self.embeddings = self.add_weight("embeddings", (self.input_dim, self.output_dim), self._init)

    def compute_output_shape(self, input_shape: Tuple[int, ...]) -> Tuple[int, ...]:
        return tuple(input_shape) + (self.output_dim,)

    def call(self, x: torch.Tensor, ctx: CallContext) -> torch.Tensor:
        return F.embedding(x.long(), ctx.param(self.embeddings))
