"""Model / Sequential with compile() and fit() (ref: /root/reference/tf_dist_example.py:39-59).

Call path of `fit` (mirrors SURVEY 3.3, re-designed):
    fit -> strategy.experimental_distribute_dataset(x)      global batch -> per-replica batches
        -> for epoch / for step: engine.train_step(PerReplica batch)
              forward + loss + backward on every local replica (gradients land in the flat arena)
              cross_device_ops.allreduce_grads(...)         ONE collective over the whole arena
              optimizer update on the arena                 (fused into the collective on B200)
              metric accumulators updated on device         (no host sync)
        -> logs are read (device->host, plus a scalar all-reduce across workers) at epoch end only

Two engines implement `train_step`: the portable TorchEngine below (CPU, stand-in baseline, anything
without a native kernel) and keras/native_engine.NativeCnnEngine (sm_100a kernels + CUDA graph).
"""
from __future__ import annotations

import os
import sys
import time
from typing import Any, Dict, Iterator, List, Optional, Sequence, Tuple, Union

import numpy as np
import torch

from .. import config as _config
from ..data import Dataset
from ..parallel import strategy as _strat
from ..parallel.strategy import DistributedDataset, Strategy, get_strategy
from ..parallel.values import MirroredVariable, ParamArena, PerReplica
from ..utils.trace import nvtx_range
from . import callbacks as _cb
from . import layers as _klayers
from . import losses as _losses
from . import metrics as _metrics
from . import optimizers as _optimizers
from .layers import CallContext, InputLayer, Layer

_policy = {"compute_dtype": torch.float32}


class mixed_precision:  # tf.keras.mixed_precision look-alike
    @staticmethod
    def set_global_policy(name: str) -> None:
        _policy["compute_dtype"] = {"float32": torch.float32, "mixed_bfloat16": torch.bfloat16,
                                    "mixed_float16": torch.float16}[name]

    @staticmethod
    def global_policy() -> str:
        return {torch.float32: "float32", torch.bfloat16: "mixed_bfloat16", torch.float16: "mixed_float16"}[
            _policy["compute_dtype"]]


class History(_cb.Callback):
    lag_ok = True

    def __init__(self) -> None:
        super().__init__()
        self.history: Dict[str, List[float]] = {}
        self.epoch: List[int] = []

    def on_epoch_end(self, epoch: int, logs: Optional[Dict[str, float]] = None) -> None:
        self.epoch.append(epoch)
        for k, v in (logs or {}).items():
            self.history.setdefault(k, []).append(v)


def _as_dataset(x: Any, y: Any, batch_size: Optional[int], shuffle: bool = False) -> Dataset:
    if isinstance(x, Dataset):
        if y is not None:
            raise ValueError("when x is a Dataset, y must not be given (the dataset yields (x, y))")
        return x
    xs = np.asarray(x)
    tensors = (xs, np.asarray(y)) if y is not None else (xs,)
    ds = Dataset.from_tensor_slices(tensors)
    if shuffle:
        ds = ds.shuffle(len(xs))
    return ds.batch(batch_size or 32)


class Model:
    """Base model: a stack (or custom graph) of layers with Keras-style training methods."""

    def __init_subclass__(cls, **kw: Any) -> None:
        super().__init_subclass__(**kw)
        _klayers.capture_init_config(cls)  # constructor arguments -> get_config(), like the layers

    def __init__(self, name: Optional[str] = None) -> None:
        self.name = name or "model"
        self.strategy: Strategy = get_strategy()  # captured at construction: build under strategy.scope()
        self.built = False
        self.optimizer: Optional[_optimizers.Optimizer] = None
        self.loss: Optional[_losses.Loss] = None
        self.compiled_metrics: List[_metrics.Metric] = []
        self.loss_tracker = _metrics.Mean("loss")
        self.stop_training = False
        self.compute_dtype = _policy["compute_dtype"]
        self._arenas: List[ParamArena] = []
        self._engine: Any = None
        self._engine_pref = os.environ.get("TFDL_ENGINE", "auto")
        self.history: Optional[History] = None
        self._train_iter: Optional[Iterator[PerReplica]] = None
        self._data_state: Dict[str, Any] = {"shuffle": [], "batches": 0}  # position of fit()'s data iterator
        self._resume_data: Optional[Dict[str, Any]] = None                # set by load_training_state
        self._steps_done = 0
        self._input_shape: Optional[Tuple[int, ...]] = None

    # ---- to be provided by subclasses -------------------------------------------------------
    def _layers(self) -> List[Layer]:
        raise NotImplementedError

    def _forward(self, x: torch.Tensor, ctx: CallContext) -> torch.Tensor:
        raise NotImplementedError

    def build(self, input_shape: Sequence[int]) -> None:
        raise NotImplementedError

    # ---- variables --------------------------------------------------------------------------
    @property
    def layers(self) -> List[Layer]:
        return [l for l in self._layers() if not isinstance(l, InputLayer)]

    @property
    def weights(self) -> List[MirroredVariable]:
        return [w for l in self._layers() for w in l.weights]

    variables = weights

    @property
    def trainable_weights(self) -> List[MirroredVariable]:
        return [w for w in self.weights if w.trainable]

    trainable_variables = trainable_weights

    def count_params(self) -> int:
        return sum(int(np.prod(w.shape)) for w in self.weights)

    def get_weights(self) -> List[np.ndarray]:
        return [w.numpy() for w in self.weights]

    def set_weights(self, ws: Sequence[np.ndarray]) -> None:
        vs = self.weights
        if len(ws) != len(vs):
            raise ValueError(f"expected {len(vs)} weight arrays, got {len(ws)}")
        for v, a in zip(vs, ws):
            v.assign(a)
        self._weights_changed()

    def _weights_changed(self) -> None:
        if self._engine is not None and hasattr(self._engine, "weights_changed"):
            self._engine.weights_changed()

    def summary(self, print_fn=print) -> None:
        print_fn(f'Model: "{self.name}"')
        print_fn(f"{'Layer (type)':<34}{'Output Shape':<26}{'Param #':>10}")
        print_fn("=" * 70)
        for l in self.layers:
            shp = (None,) + tuple(l.output_shape) if l.output_shape is not None else "?"
            print_fn(f"{l.name + ' (' + type(l).__name__ + ')':<34}{str(shp):<26}{l.count_params():>10}")
        print_fn("=" * 70)
        print_fn(f"Total params: {self.count_params():,}")

    # ---- finalisation: arenas, mirrored copies, chief broadcast ------------------------------
    def _finalize(self) -> None:
        """Allocate one flat (params, grads) arena per local replica, bind every variable to its
        slice, and make all replicas of all workers start from the chief's values (K15)."""
        if self._arenas:
            return
        if not self.built:
            raise RuntimeError("the model has no input shape yet: pass input_shape= to the first layer, call "
                               "model.build(input_shape), or call fit() which builds from the data")
        s = self.strategy
        specs = [w.spec for w in self.weights]
        ops = s.cross_device_ops
        alloc = getattr(ops, "alloc_arena", None)
        self._arenas = [ParamArena(specs, d, allocator=alloc) for d in s.local_devices]
        for w in self.weights:
            w.bind(self._arenas)
        mask = torch.zeros(self._arenas[0].numel)
        for sp in specs:
            if sp.trainable:
                mask[sp.offset:sp.offset + sp.numel] = 1
        self._all_trainable = bool(all(sp.trainable for sp in specs))
        self._masks = [mask.to(a.device) for a in self._arenas]
        ops.broadcast_params([a.params for a in self._arenas])

    @property
    def arenas(self) -> List[ParamArena]:
        self._finalize()
        return self._arenas

    # ---- compile ----------------------------------------------------------------------------
    def compile(self, optimizer: Any = "sgd", loss: Any = None, metrics: Optional[Sequence[Any]] = None,
                **_: Any) -> None:
        """tf.keras.Model.compile (tf_dist_example.py:49-52). Creates optimizer slots and metric
        accumulators; call it inside strategy.scope() like the reference does."""
        scope = get_strategy()
        if _strat.has_strategy() and scope is not self.strategy:
            raise RuntimeError("compile() is running under a different strategy scope than the one the model "
                               "was created in")
        self.optimizer = _optimizers.get(optimizer)
        self.loss = _losses.get(loss) if loss is not None else None
        self.compiled_metrics = [_metrics.get(m, self.loss) for m in (metrics or [])]
        devs = self.strategy.local_devices
        self.loss_tracker.bind(devs)
        for m in self.compiled_metrics:
            m.bind(devs)
        self._engine = None
        if self.built:
            self._finalize()
            self.optimizer.bind([a.params for a in self._arenas])

    @property
    def metrics_names(self) -> List[str]:
        return ["loss"] + [m.name for m in self.compiled_metrics]

    # ---- engine -----------------------------------------------------------------------------
    def _make_engine(self) -> Any:
        if self._engine is not None:
            return self._engine
        self._finalize()
        assert self.optimizer is not None, "call compile() before fit()/evaluate()"
        if not self.optimizer._slots and self.optimizer.slot_names():
            self.optimizer.bind([a.params for a in self._arenas])
        elif not self.optimizer.slot_names():
            self.optimizer._slots = {}
        pref = self._engine_pref
        if _config.is_op_determinism_enabled():
            pref = "torch"  # the native step programs accumulate with atomics (config.py)
        on_cuda = all(d.type == "cuda" for d in self.strategy.local_devices)
        engine = None
        if pref in ("auto", "native") and on_cuda:
            from . import native_engine

            engine = native_engine.try_build(self)  # raises if the .so is missing on a GPU box
            if engine is None and pref == "native":
                raise RuntimeError("TFDL_ENGINE=native but this model/strategy has no native step program")
        elif pref == "native":
            raise RuntimeError("TFDL_ENGINE=native needs CUDA replicas")
        self._engine = engine or TorchEngine(self)
        return self._engine

    @property
    def engine_name(self) -> str:
        return type(self._make_engine()).__name__

    # ---- logs -------------------------------------------------------------------------------
    def _read_logs(self, prefix: str = "") -> Dict[str, float]:
        eng = self._engine
        if eng is not None and hasattr(eng, "flush_metrics"):
            eng.flush_metrics()
        trackers = [self.loss_tracker] + self.compiled_metrics
        local = np.concatenate([t.local_state() for t in trackers])
        tot = self.strategy.cross_device_ops.allreduce_scalars(local)
        out = {}
        for i, t in enumerate(trackers):
            total, count = tot[2 * i], tot[2 * i + 1]
            out[prefix + t.name] = float(total / count) if count else 0.0
        return out

    def reset_metrics(self) -> None:
        if self._engine is not None and hasattr(self._engine, "flush_metrics"):
            self._engine.flush_metrics()
        for t in [self.loss_tracker] + self.compiled_metrics:
            t.reset_state()

    # ---- fit --------------------------------------------------------------------------------
    def fit(self, x: Any = None, y: Any = None, batch_size: Optional[int] = None, epochs: int = 1,
            verbose: Union[int, str] = 1, callbacks: Optional[Sequence[_cb.Callback]] = None,
            validation_data: Any = None, shuffle: bool = True, initial_epoch: int = 0,
            steps_per_epoch: Optional[int] = None, validation_steps: Optional[int] = None,
            validation_split: float = 0.0, **unused: Any) -> History:
        """tf.keras.Model.fit (tf_dist_example.py:59: fit(x=dataset, epochs=10, steps_per_epoch=20)).

        With `steps_per_epoch` the dataset iterator is carried across epochs (the reference has no
        .repeat(): 200 steps x 64 samples stay below one pass of the data); without it every epoch
        is one full pass."""
        if self.optimizer is None or self.loss is None:
            raise RuntimeError("compile(optimizer=..., loss=...) must be called before fit()")
        t_fit = time.time()
        for k, v in unused.items():
            # arguments that change the result must not be dropped silently; the ones that only tune Keras' own input
            # machinery (workers, use_multiprocessing, max_queue_size, validation_freq, ...) have no meaning here
            if k in ("class_weight", "sample_weight") and v is not None:
                raise NotImplementedError(f"fit({k}=...) is not supported")
        if validation_split:
            # Keras: the LAST fraction of the arrays, taken before any shuffling, is held out for validation
            if isinstance(x, (Dataset, DistributedDataset)) or y is None:
                raise ValueError("validation_split needs array inputs x and y (a dataset cannot be split this way)")
            if not 0.0 < validation_split < 1.0:
                raise ValueError("validation_split must be in (0, 1)")
            xs, ys = np.asarray(x), np.asarray(y)
            cut = int(len(xs) * (1.0 - validation_split))
            if validation_data is None:
                validation_data = (xs[cut:], ys[cut:])
            x, y = xs[:cut], ys[:cut]
        ds = _as_dataset(x, y, batch_size, shuffle=shuffle and not isinstance(x, Dataset))
        dist = ds if isinstance(x, DistributedDataset) else self.strategy.experimental_distribute_dataset(ds)
        if not self.built:
            first = next(iter(dist))
            xb = first[0][0] if isinstance(first[0], (tuple, list)) else first[0]
            self.build(tuple(np.asarray(xb).shape[1:]))
        engine = self._make_engine()
        history = History()
        callbacks = list(callbacks or [])
        if os.environ.get("TFDL_FAULT") and not any(isinstance(c, _cb.FaultInjection) for c in callbacks):
            callbacks.append(_cb.FaultInjection())  # fault injection by environment alone
        # History after the user's callbacks, as in Keras: what they add to the logs (ReduceLROnPlateau's "lr") is recorded
        cbs = _cb.CallbackList(callbacks + [history], model=self, verbose=verbose, epochs=epochs,
                               steps=steps_per_epoch)
        self.stop_training = False
        logs: Dict[str, float] = {}
        cbs.on_train_begin()
        initial_epoch = cbs.initial_epoch(initial_epoch)
        chunked = hasattr(engine, "train_chunk") and not cbs.has_batch_hooks
        # Index path: the engine gathers its batches itself, straight from the dataset arrays into its pinned ring.
        idx_mode = None
        if chunked and hasattr(engine, "train_index_chunk") and os.environ.get("TFDL_FIT_INDEX", "1") == "1":
            st = dist.index_stream() if hasattr(dist, "index_stream") else None
            if st is not None:
                arrs = engine.index_arrays(st[0])
                if arrs is not None:
                    idx_mode = arrs
        # Lagged epoch ends: when every callback only consumes logs, epoch k's totals are read (and its callbacks run)
        # after epoch k+1 has been queued, so the device never idles at an epoch boundary. The totals themselves are
        # exact - they are closed on the device, stream-ordered after the epoch's last step.
        lag = (chunked and hasattr(engine, "epoch_mark") and validation_data is None and cbs.lag_ok
               and os.environ.get("TFDL_FIT_LAG", "1") == "1")
        pending: Optional[Tuple[int, Any, float, int]] = None

        def deliver(p: Tuple[int, Any, float, int]) -> Dict[str, float]:
            ep, handle, t_start, n_steps = p
            loss_sum, correct, count = engine.epoch_read(handle)
            out = {"loss": loss_sum / count if count else 0.0}
            for met in self.compiled_metrics:
                out[met.name] = correct / count if count else 0.0
            out["_epoch_time"] = time.time() - t_start
            out["_steps"] = n_steps
            cbs.on_epoch_end(ep, out)
            return out

        it: Any = None
        if lag:
            # epoch totals come from the engine's device counters (epoch_mark), never from the Keras accumulators: only
            # those counters have to be clean - zeroing the accumulators would be four kernel launches in front of the
            # first step of every fit()
            engine.flush_metrics()
        else:
            self.reset_metrics()
        t_loop = time.time()
        for epoch in range(initial_epoch, epochs):
            if not lag and epoch > initial_epoch:
                self.reset_metrics()
            cbs.on_epoch_begin(epoch)
            t0 = time.time()
            if it is None or steps_per_epoch is None:
                from ..data import dataset as _dsmod

                resume = self._resume_data
                self._resume_data = None
                if resume is not None and steps_per_epoch is not None:
                    _dsmod.restore_iterator_state(dist.dataset, resume["shuffle"])  # same order as the interrupted run
                # remembered for backups: the seeds this pass starts from and how many steps it has handed out
                self._data_state = {"shuffle": _dsmod.iterator_state(dist.dataset), "batches": 0}
                if idx_mode is not None:
                    if it is not None or resume is not None:
                        st = dist.index_stream()  # a new pass over the data: fresh shuffle order
                    it = iter(st[1])
                else:
                    it = iter(dist)
                if resume is not None and steps_per_epoch is not None:
                    for _ in range(int(resume["batches"])):  # skip what the interrupted run already consumed
                        if next(it, None) is None:
                            break
                    self._data_state["batches"] = int(resume["batches"])
            step = 0
            while steps_per_epoch is None or step < steps_per_epoch:
                want = 1
                if chunked:
                    if steps_per_epoch is not None:
                        want = min(engine.CHUNK, steps_per_epoch - step)  # the epoch length is known: no ramp
                    else:
                        # chunk sizes ramp 8, 16, 32, ... so the device starts after 8 batches are staged, not after CHUNK
                        want = min(engine.CHUNK, 8 << min(step // 8, 16).bit_length()) if step < engine.CHUNK else engine.CHUNK
                group, exhausted = [], False
                for _ in range(want):
                    try:
                        group.append(next(it))
                    except StopIteration:
                        exhausted = True
                        break
                if group:
                    if idx_mode is not None:
                        engine.train_index_chunk(idx_mode[0], idx_mode[1], group)
                    elif chunked:
                        engine.train_chunk(group)  # the native executor runs the whole chunk without Python
                    else:
                        cbs.on_train_batch_begin(step)
                        engine.train_step(group[0])
                    self._steps_done += len(group)
                    self._data_state["batches"] += len(group)
                    if not chunked:
                        cbs.on_train_batch_end(step, None)
                    step += len(group)
                if exhausted:
                    if steps_per_epoch is not None:
                        if verbose:
                            print(f"WARNING: input ran out of data after {step} of {steps_per_epoch} steps "
                                  f"in epoch {epoch + 1}; use .repeat() on the dataset", file=sys.stderr)
                        self.stop_training = True
                    break
                if self.stop_training:
                    break
            if lag:
                handle = engine.epoch_mark()
                if pending is not None:
                    logs = deliver(pending)
                pending = (epoch, handle, t0, step)
            else:
                logs = self._read_logs()
                if validation_data is not None:
                    if isinstance(validation_data, (tuple, list)) and len(validation_data) == 2 and \
                            not isinstance(validation_data[0], (Dataset, DistributedDataset)):
                        vx, vy = validation_data  # the (x_val, y_val) arrays form
                    else:
                        vx, vy = validation_data, None
                    val = self.evaluate(vx, vy, batch_size=batch_size, steps=validation_steps, verbose=0,
                                        return_dict=True, _prefix="val_")
                    logs.update(val)
                logs["_epoch_time"] = time.time() - t0
                logs["_steps"] = step
                cbs.on_epoch_end(epoch, logs)
            if self.stop_training:
                break
        if pending is not None:
            logs = deliver(pending)
        if os.environ.get("TFDL_FIT_PROFILE") and hasattr(engine, "prof"):
            print("[tfdl fit profile, host ms] " + ", ".join(f"{k} {1e3 * v:.2f}" for k, v in engine.prof.items())
                  + f", setup {1e3 * (t_loop - t_fit):.2f}, fit total {1e3 * (time.time() - t_fit):.2f}", file=sys.stderr)
            for k in engine.prof:
                engine.prof[k] = 0.0
        cbs.on_train_end(logs)
        for k in ("_epoch_time", "_steps"):
            history.history.pop(k, None)
        self.history = history
        return history

    def train_on_batch(self, x: Any, y: Any = None, return_dict: bool = False) -> Any:
        """One step on one GLOBAL batch given as arrays; the batch is split over local replicas."""
        engine = self._prepare_for_arrays(x)
        self.reset_metrics()
        engine.train_step(self._split_arrays(x, y))
        self._steps_done += 1
        logs = self._read_logs()
        return logs if return_dict else [logs[n] for n in self.metrics_names]

    def test_on_batch(self, x: Any, y: Any = None, return_dict: bool = False) -> Any:
        """Loss and metrics of one GLOBAL batch given as arrays, without a training step."""
        engine = self._prepare_for_arrays(x)
        self.reset_metrics()
        engine.test_step(self._split_arrays(x, y))
        logs = self._read_logs()
        return logs if return_dict else [logs[n] for n in self.metrics_names]

    def predict_on_batch(self, x: Any) -> np.ndarray:
        return self.predict(x, batch_size=max(1, len(np.asarray(x))))

    @property
    def metrics(self) -> List[_metrics.Metric]:
        """The loss tracker followed by the compiled metrics (tf.keras.Model.metrics)."""
        return [self.loss_tracker] + list(self.compiled_metrics)

    def _prepare_for_arrays(self, x: Any) -> Any:
        if not self.built:
            self.build(tuple(np.asarray(x).shape[1:]))
        return self._make_engine()

    def _split_arrays(self, x: Any, y: Any) -> PerReplica:
        n = self.strategy.num_local_replicas
        xs = np.array_split(np.asarray(x), n)
        if y is None:
            return PerReplica([(a,) for a in xs])
        ys = np.array_split(np.asarray(y), n)
        return PerReplica(list(zip(xs, ys)))

    # ---- evaluate / predict --------------------------------------------------------------------
    def evaluate(self, x: Any = None, y: Any = None, batch_size: Optional[int] = None, verbose: Union[int, str] = 0,
                 steps: Optional[int] = None, return_dict: bool = False, _prefix: str = "", **_: Any) -> Any:
        ds = _as_dataset(x, y, batch_size)
        dist = ds if isinstance(x, DistributedDataset) else self.strategy.experimental_distribute_dataset(ds)
        if not self.built:
            first = next(iter(dist))
            self.build(tuple(np.asarray(first[0][0]).shape[1:]))
        engine = self._make_engine()
        saved = [(t, [v.clone() for v in t._total], [v.clone() for v in t._count])
                 for t in [self.loss_tracker] + self.compiled_metrics] if _prefix else []
        self.reset_metrics()
        for i, batch in enumerate(dist):
            if steps is not None and i >= steps:
                break
            engine.test_step(batch)
        logs = self._read_logs(_prefix)
        for t, tot, cnt in saved:  # restore the training accumulators
            for dst, src in zip(t._total, tot):
                dst.copy_(src)
            for dst, src in zip(t._count, cnt):
                dst.copy_(src)
        if verbose:
            print(" - ".join(f"{k}: {v:.4f}" for k, v in logs.items()))
        return logs if return_dict else [logs[_prefix + n] for n in self.metrics_names]

    def predict(self, x: Any, batch_size: Optional[int] = None, steps: Optional[int] = None, **_: Any) -> np.ndarray:
        ds = _as_dataset(x, None, batch_size)
        if not self.built:
            first = next(iter(ds))
            xb = first[0] if isinstance(first, (tuple, list)) else first
            self.build(tuple(np.asarray(xb).shape[1:]))
        # a model that was never compiled can still predict (as in Keras): eager forward on replica 0
        engine = self._make_engine() if self.optimizer is not None else None
        outs = []
        for i, b in enumerate(ds):
            if steps is not None and i >= steps:
                break
            xb = b[0] if isinstance(b, (tuple, list)) else b
            outs.append(engine.predict_step(np.asarray(xb)) if engine is not None
                        else self(np.asarray(xb)).float().cpu().numpy())
        return np.concatenate(outs) if outs else np.zeros((0,))

    def __call__(self, x: Any, training: bool = False) -> torch.Tensor:
        """Eager forward on replica 0 (torch ops)."""
        if not self.built:
            self.build(tuple(np.asarray(x).shape[1:]))
        self._finalize()
        dev = self.strategy.local_devices[0]
        xt = torch.as_tensor(np.asarray(x)).to(dev)
        ctx = CallContext(0, training, lambda v: v.value(0), self.compute_dtype)
        with torch.no_grad():
            return self._forward(xt, ctx)

    # ---- weights IO -----------------------------------------------------------------------------
    def save_weights(self, filepath: str, overwrite: bool = True, save_format: Optional[str] = None) -> None:
        """Keras TF-format weights (TensorBundle with object-graph keys) or `.npz`. Under a
        multi-worker strategy every worker must call this; only the chief writes (README.md:51)."""
        from ..checkpoint import save_model_weights

        save_model_weights(self, filepath, save_format)

    def load_weights(self, filepath: str) -> None:
        from ..checkpoint import load_model_weights

        load_model_weights(self, filepath)
        self._weights_changed()

    # ---- whole-model save (architecture + compile arguments + variables) ----------------------------
    def get_config(self) -> Dict[str, Any]:
        """Constructor arguments of a subclassed model (Sequential overrides this with its layer list)."""
        return _klayers.jsonable_config(getattr(self, "_init_config", {}))

    @classmethod
    def from_config(cls, config: Dict[str, Any], custom_objects: Optional[Dict[str, Any]] = None) -> "Model":
        return cls(**config)

    def to_json(self) -> str:
        from . import models as _models

        return _models.model_to_json(self)

    def save(self, filepath: str, overwrite: bool = True, include_optimizer: bool = True) -> None:
        """tf.keras.Model.save: a directory with the architecture and compile arguments (`model.json`) and the
        variables as a TensorFlow-format checkpoint (`variables/variables.*`, optimizer slots included). Under a
        multi-worker strategy every worker calls this and only the chief writes (README.md:51 of the reference).
        `tfdl.keras.models.load_model` restores it."""
        from . import models as _models

        _models.save_model(self, filepath, overwrite=overwrite, include_optimizer=include_optimizer)


class Sequential(Model):
    """tf.keras.Sequential (tf_dist_example.py:40-48)."""

    def __init__(self, layers: Optional[Sequence[Layer]] = None, name: Optional[str] = None) -> None:
        super().__init__(name or "sequential")
        self._stack: List[Layer] = []
        for l in layers or []:
            self.add(l)

    def add(self, layer: Layer) -> None:
        if self._arenas:
            raise RuntimeError("cannot add layers after the model's variables have been created")
        self._stack.append(layer)
        if self.built:
            assert self._stack[-2].output_shape is not None
            layer._maybe_build(self._stack[-2].output_shape)
        elif self._stack[0].input_shape_arg is not None:
            self.build(self._stack[0].input_shape_arg)

    def _layers(self) -> List[Layer]:
        return self._stack

    def get_config(self) -> Dict[str, Any]:
        return {"name": self.name,
                "layers": [{"class_name": type(l).__name__, "module": type(l).__module__, "config": l.get_config()}
                           for l in self._stack],
                "build_input_shape": list(self._input_shape) if self._input_shape else None}

    @classmethod
    def from_config(cls, config: Dict[str, Any], custom_objects: Optional[Dict[str, Any]] = None) -> "Sequential":
        from . import models as _models

        model = cls(name=config.get("name"))
        for spec in config["layers"]:
            layer_cls = _models.resolve_class(spec["class_name"], spec.get("module"), custom_objects, Layer)
            model.add(layer_cls.from_config(dict(spec["config"])))
        if not model.built and config.get("build_input_shape"):
            model.build(tuple(config["build_input_shape"]))
        return model

    def build(self, input_shape: Sequence[int]) -> None:
        shape = tuple(int(s) for s in input_shape)
        self._input_shape = shape
        for l in self._stack:
            shape = l._maybe_build(shape)
        self.built = True

    @property
    def input_shape(self) -> Optional[Tuple[Optional[int], ...]]:
        return (None,) + self._input_shape if self._input_shape else None

    @property
    def output_shape(self) -> Optional[Tuple[Optional[int], ...]]:
        return (None,) + tuple(self._stack[-1].output_shape) if self.built else None

    def _forward(self, x: torch.Tensor, ctx: CallContext) -> torch.Tensor:
        stack, i = self._stack, 0
        while i < len(stack):
            l = stack[i]
            if (not ctx.native and not x.is_cuda and type(l) is _klayers.Conv2D and l.activation == "relu"
                    and i + 1 < len(stack) and type(stack[i + 1]) is _klayers.MaxPooling2D):
                # CPU workers (library kernels, no fusion): ReLU and max-pool commute exactly - relu(max(w)) ==
                # max(relu(w)) for every window, gradients included - so the ReLU runs after the pool, on a quarter of
                # the elements (tf_dist_example.py:41-44: both convolutions of the reference model)
                x = stack[i + 1].call(l.call(x, ctx, defer_relu=True), ctx)
                x = torch.relu(x)
                i += 2
                continue
            x = l.call(x, ctx)
            i += 1
        return x


class TorchEngine:
    """Portable training engine on torch ops + autograd. Gradients are accumulated straight into the
    flat gradient arena (each parameter leaf's .grad is a view of it), the strategy's cross-device
    op reduces the arena in place, and the optimizer updates the parameter arena in one pass."""

    def __init__(self, model: Model) -> None:
        self.model = model
        self.strategy = model.strategy
        self.devices = self.strategy.local_devices
        self.leaves: List[Dict[int, torch.Tensor]] = []
        self._fused = None
        self._fused_lr = None
        self._graphs: Dict[Any, Any] = {}
        self._graph_seen: Dict[Any, int] = {}
        self._graph_off = False
        self._buckets: List[Any] = []        # (FusedState view, algo) in reverse layer order
        self._bucket_of: Dict[int, int] = {}  # id(variable) -> bucket number
        self._bucket_left: List[int] = []
        self._bucket_need: List[int] = []
        self._comm_stream: Any = None
        self.launches_per_step = 0
        ops = self.strategy.cross_device_ops
        opt0 = model.optimizer
        fusable = isinstance(opt0, _optimizers.SGD) or (isinstance(opt0, _optimizers.Adam) and not opt0.weight_decay)
        if (type(ops).__name__ == "P2PFusedAllReduce" and fusable
                and len(self.devices) == 1 and os.environ.get("TFDL_TORCH_FUSED_AR", "1") == "1"):
            # GPU process group + SGD: the gradient all-reduce, 1/N scale and the (momentum) update are ONE
            # NVLink kernel over the arena (ops/allreduce.py) even though forward/backward run on torch ops.
            # A single gradient buffer is used here, so only the variants with an end barrier are legal.
            arena = model.arenas[0]
            is_adam = isinstance(opt0, _optimizers.Adam)
            st = ops.fused_state(arena.params, arena.grads, momentum=bool(getattr(opt0, "momentum", 0.0)), adam=is_adam)
            if is_adam:
                # the optimizer's slots ARE the kernel's arenas (values restored from a checkpoint before this point move in)
                old = dict(opt0._slots) if opt0._slots else {}
                for name, buf in (("m", st.mom), ("v", st.mom2)):
                    if old.get(name):
                        buf.copy_(old[name][0])
                opt0._slots = {"m": [st.mom], "v": [st.mom2]}
                self._adam_steps0 = int(opt0.iterations)  # Adam's t before this state's first kernel
            elif opt0.momentum:
                opt0._slots = {"momentum": [st.mom]}
            from ..ops.structs import AR_MULTIMEM, AR_ONESHOT, AR_TWOSHOT

            from ..ops.allreduce import multimem_ok

            st.torch_algo = AR_ONESHOT if ops.world == 1 else (
                AR_MULTIMEM if multimem_ok(ops.world, st.has_multicast) else AR_TWOSHOT)
            self._fused = st
            st.refresh_shadow()
            self._make_buckets(ops)
        for r, arena in enumerate(model.arenas):
            d: Dict[int, torch.Tensor] = {}
            for w in model.weights:
                leaf = w.value(r).detach()
                if w.trainable:
                    leaf.requires_grad_(True)
                    leaf.grad = w.grad(r)
                    if r == 0 and id(w) in self._bucket_of:
                        leaf.register_post_accumulate_grad_hook(lambda _t, _i=id(w): self._grad_ready(_i))
                d[id(w)] = leaf
            self.leaves.append(d)

    # ---- gradient buckets reduced while backward is still running ----------------------------------------------
    # The reference reduces ONE packed buffer after the whole backward pass (num_packs=1, no overlap). Here
    # the arena is cut into buckets in reverse layer order; when the last gradient of a bucket has been accumulated
    # (post-accumulate-grad hooks count them down) the bucket's fused all-reduce + update kernel is queued on a
    # communication stream and runs next to the rest of backward. Updating a bucket's weights early is safe: every
    # backward op that reads them has been queued before the hook fires. The whole arrangement is captured in the step's
    # CUDA graph, so the hooks only run while capturing and in the eager warm-up steps.
    def _make_buckets(self, ops: Any) -> None:
        if ops.world == 1 or os.environ.get("TFDL_TORCH_BUCKETS", "1") != "1":
            return
        target = int(float(os.environ.get("TFDL_BUCKET_MB", "32")) * (1 << 20)) // 4  # elements
        ws = [w for w in self.model.weights]
        n = self._fused.n
        cuts, hi = [], n
        for w in reversed(ws):
            if hi - w.spec.offset >= target and w.spec.offset > 0:
                cuts.append((w.spec.offset, hi))
                hi = w.spec.offset
        cuts.append((0, hi))
        if len(cuts) < 2:
            return
        from ..ops.allreduce import pick_algo
        from ..ops.structs import AR_MULTIMEM, AR_PUSH, AR_PUSH2, AR_TWOSHOT

        for k, (lo, hi) in enumerate(cuts):
            b = self._fused.bucket(lo, hi - lo)
            algo = pick_algo((hi - lo) * 4, ops.world, b.has_multicast)
            if algo not in (AR_PUSH, AR_PUSH2, AR_MULTIMEM, AR_TWOSHOT):  # one gradient buffer: no barrier-less pull
                algo = self._fused.torch_algo
            b.ensure_inbox(algo)
            self._buckets.append((b, algo))
            need = 0
            for w in ws:
                if lo <= w.spec.offset < hi and w.trainable:
                    self._bucket_of[id(w)] = k
                    need += 1
            self._bucket_need.append(need)
        # ONE communication stream on purpose: two bucket kernels in flight at once could each hold the SM slots the
        # other one's not-yet-resident blocks need on a different GPU (every block waits for its peers' matching block),
        # a cross-GPU cycle; a single kernel at a time only ever waits for compute kernels, which always finish.
        self._comm_streams = [torch.cuda.Stream(self.devices[0])]
        self._comm_stream = self._comm_streams[0]
        self.bucket_info = [{"elements": hi - lo, "algo": int(a)} for (lo, hi), (_, a) in zip(cuts, self._buckets)]

    def _fire_bucket(self, k: int) -> None:
        b, algo = self._buckets[k]
        cur = torch.cuda.current_stream(self.devices[0])
        ev = torch.cuda.Event()
        ev.record(cur)
        cs = self._comm_streams[k % len(self._comm_streams)]
        cs.wait_event(ev)
        with torch.cuda.stream(cs):
            b.step(0, algo)
        self._bucket_left[k] = -1

    def _grad_ready(self, var_id: int) -> None:
        k = self._bucket_of.get(var_id)
        if k is None or not self._bucket_left or self._bucket_left[k] <= 0:
            return
        self._bucket_left[k] -= 1
        if self._bucket_left[k] == 0:
            self._fire_bucket(k)

    def _ctx(self, r: int, training: bool) -> CallContext:
        leaves = self.leaves[r]
        # CUDA replicas run the layer ops on the native kernels under either policy (same precision contract as the
        # native step program); an engine explicitly requested as "torch" stays on library ops unless mixed_bfloat16
        # asks for bf16 anyway - that is the independent reference the tests compare against.
        native = None
        if self.devices[r].type == "cuda" and self.model._engine_pref != "torch" and \
                os.environ.get("TFDL_NATIVE_OPS", "1") == "1" and not _config.is_op_determinism_enabled():
            native = True
        ctx = CallContext(r, training, lambda v: leaves[id(v)], self.model.compute_dtype, native)
        if native and self._fused is not None and r == 0 and os.environ.get("TFDL_TORCH_INPLACE", "1") == "1":
            ctx.param_lp = self._lp_view
            ctx.grad_sink = self._sink_for
        return ctx

    # bf16 copies of the weights come from the shadow arena that the fused all-reduce + update kernel rewrites every step
    # (no per-layer cast kernels); weight gradients of Conv2D / Dense are written into the gradient arena by the
    # weight-gradient kernels themselves (no dW tensors, no autograd accumulate pass).
    def _lp_view(self, v: Any) -> Optional[torch.Tensor]:
        sh = self._fused.shadow
        if sh is None or not v.trainable:
            return None
        sp = v.spec
        return sh.tensor[sp.offset:sp.offset + sp.numel].view(sp.storage_shape)

    def _sink_for(self, v: Any) -> Any:
        if not v.trainable:
            return None
        return v.grad(0), (lambda _i=id(v): self._grad_ready(_i))

    def weights_changed(self) -> None:
        if self._fused is not None and self._fused.shadow is not None:
            self._fused.refresh_shadow()

    def _to_device(self, a: Any, dev: torch.device) -> torch.Tensor:
        t = torch.from_numpy(np.ascontiguousarray(a)) if not isinstance(a, torch.Tensor) else a
        if dev.type == "cuda":
            return t.pin_memory().to(dev, non_blocking=True) if not t.is_cuda else t
        return t

    def _split(self, batch: Any) -> Tuple[Any, Any]:
        if isinstance(batch, (tuple, list)):
            return batch[0], (batch[1] if len(batch) > 1 else None)
        return batch, None

    def train_step(self, batches: PerReplica) -> None:
        m = self.model
        assert m.loss is not None and m.optimizer is not None
        chk = getattr(self.strategy.cross_device_ops, "check_abort", None)
        if chk is not None:
            chk()  # a peer died: the previous step's collective was released, fail here instead of training on
        if self._graph_eligible():
            x, y = self._split(batches[0])
            n = int(np.asarray(x).shape[0]) if not isinstance(x, torch.Tensor) else x.shape[0]
            if n > 0 and y is not None:
                dev = self.devices[0]
                self._train_step_graphed(self._to_device(x, dev), self._to_device(y, dev))
                return
        self._sync_hyper()
        items: List[Optional[Tuple[torch.Tensor, torch.Tensor]]] = []
        for r, dev in enumerate(self.devices):
            x, y = self._split(batches[r])
            n = x.shape[0] if isinstance(x, torch.Tensor) else int(np.asarray(x).shape[0])
            items.append((self._to_device(x, dev), self._to_device(y, dev)) if n else None)
        self._step_body(items)
        m.optimizer.step_done()

    # ---- whole-step CUDA graph -------------------------------------------------------------------------------
    # The eager step of a transformer is several hundred small launches and the host, not the GPU, sets the step
    # time. For one-replica-per-process CUDA training with SGD the whole step (zero gradients, forward, backward,
    # metric counters, fused all-reduce + update) is captured once per (shapes, learning rate) and replayed; inputs
    # are copied into static buffers. The reference gets the same effect from tf.function tracing (README.md:67).
    def _graph_eligible(self) -> bool:
        if self._graph_off or len(self.devices) != 1 or self.devices[0].type != "cuda":
            return False
        opt = self.model.optimizer
        ok_opt = isinstance(opt, _optimizers.SGD) or (isinstance(opt, _optimizers.Adam) and self._fused is not None)
        if not ok_opt or callable(getattr(opt, "_lr", None)):
            return False
        return os.environ.get("TFDL_TORCH_GRAPH", "1") == "1"

    def _sync_hyper(self) -> None:
        if self._fused is not None:
            opt = self.model.optimizer
            if self._fused_lr != opt.learning_rate:
                self._fused_lr = opt.learning_rate
                for st in [self._fused] + [b for b, _ in self._buckets]:
                    if isinstance(opt, _optimizers.Adam):
                        st.set_hyper(opt.learning_rate, 0.0, False, 1.0, opt.beta_1, opt.beta_2, opt.epsilon,
                                     float(self._adam_steps0))
                    else:
                        st.set_hyper(opt.learning_rate, opt.momentum, opt.nesterov, 1.0)

    def _train_step_graphed(self, xt: torch.Tensor, yt: torch.Tensor) -> None:
        opt = self.model.optimizer
        key = (tuple(xt.shape), xt.dtype, tuple(yt.shape), yt.dtype, float(opt.learning_rate),
               getattr(opt, "momentum", 0.0), getattr(opt, "nesterov", False))
        self._sync_hyper()
        ent = self._graphs.get(key)
        if ent is None:
            seen = self._graph_seen.get(key, 0)
            self._graph_seen[key] = seen + 1
            if seen < 3:  # warm-up (library handles, autotuning, allocator pools) before anything is captured
                self._step_body([(xt, yt)])
                opt.step_done()
                return
            try:
                sx, sy = xt.clone(), yt.clone()
                torch.cuda.synchronize(self.devices[0])
                graph = torch.cuda.CUDAGraph()
                from ..ops import _lib as _nl

                calls0 = _nl.native_calls
                with torch.cuda.graph(graph):
                    self._step_body([(sx, sy)])
                self.launches_per_step = _nl.native_calls - calls0  # native entry points captured in one step
                ent = self._graphs[key] = (graph, sx, sy)
            except Exception as e:  # an op that cannot be captured: stay eager for this engine
                self._graph_off = True
                torch.cuda.synchronize(self.devices[0])
                if os.environ.get("TFDL_TORCH_GRAPH_DEBUG"):
                    print(f"[tfdl] step graph capture failed, running eagerly: {e!r}", file=sys.stderr)
                self._step_body([(xt, yt)])
                opt.step_done()
                return
        graph, sx, sy = ent
        sx.copy_(xt, non_blocking=True)
        sy.copy_(yt, non_blocking=True)
        graph.replay()
        opt.step_done()

    def _step_body(self, per_replica: List[Optional[Tuple[torch.Tensor, torch.Tensor]]]) -> None:
        """Device work of one step: forward / backward per local replica, gradient collective, update."""
        m, s = self.model, self.strategy
        nrep = s.num_replicas_in_sync
        for r, item in enumerate(per_replica):
            arena = m.arenas[r]
            arena.grads.zero_()
            if self._buckets and r == 0:
                self._bucket_left = list(self._bucket_need)
            if item is None:
                continue
            xt, yt = item
            ctx = self._ctx(r, True)
            with nvtx_range("tfdl/forward"):
                pred = m._forward(xt, ctx)
                per = m.loss.per_example(yt, pred)
            with nvtx_range("tfdl/backward"):
                (per.mean() / nrep).backward()
            for fn in ctx.deferred:
                fn()
            m.loss_tracker.add(r, per.detach().sum(), float(per.numel()))
            with torch.no_grad():
                for met in m.compiled_metrics:
                    met.update_state(yt, pred.detach(), r)
        if self._fused is not None:
            with nvtx_range("tfdl/allreduce+sgd"):
                if self._buckets:
                    for k in range(len(self._buckets)):  # buckets whose variables saw no gradient this step
                        if self._bucket_left[k] >= 0:
                            self._fire_bucket(k)
                    for cs in self._comm_streams:
                        torch.cuda.current_stream(self.devices[0]).wait_stream(cs)
                    self._bucket_left = []
                else:
                    self._fused.step(0, self._fused.torch_algo)
            return
        with nvtx_range("tfdl/allreduce"):
            s.cross_device_ops.allreduce_grads([a.grads for a in m.arenas])
        with nvtx_range("tfdl/optimizer"), torch.no_grad():
            for r, arena in enumerate(m.arenas):
                m.optimizer.apply_flat(arena.params, arena.grads, r, None if m._all_trainable else m._masks[r])

    def test_step(self, batches: PerReplica) -> None:
        m = self.model
        assert m.loss is not None
        with torch.no_grad():
            for r, dev in enumerate(self.devices):
                x, y = self._split(batches[r])
                if np.asarray(x).shape[0] == 0:
                    continue
                xt, yt = self._to_device(x, dev), self._to_device(y, dev)
                pred = m._forward(xt, self._ctx(r, False))
                per = m.loss.per_example(yt, pred)
                m.loss_tracker.add(r, per.sum(), float(per.numel()))
                for met in m.compiled_metrics:
                    met.update_state(yt, pred, r)

    def predict_step(self, x: np.ndarray) -> np.ndarray:
        with torch.no_grad():
            xt = self._to_device(x, self.devices[0])
            return self.model._forward(xt, self._ctx(0, False)).float().cpu().numpy()
