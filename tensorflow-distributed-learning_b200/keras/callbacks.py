"""Callbacks. The reference's fit() passes none (tf_dist_example.py:59) but its README assigns the
chief "extra work such as saving checkpoints and writing TensorBoard" (README.md:51); these are
the callbacks that do that work, chief-only, plus the progress line Keras prints on every worker."""
from __future__ import annotations

import json
import os
import shutil
import sys
import time
from typing import Any, Callable, Dict, List, Optional, Sequence, Union


class Callback:
    # True for callbacks that only consume an epoch's logs (history, progress line, scalar writers): fit() may then
    # hand them epoch k's logs while epoch k+1 is already running, so the device never waits for the host at an epoch end.
    # Anything that acts on the model at the epoch boundary (checkpoints, early stopping, LR schedules) keeps False.
    lag_ok = False

    def __init__(self) -> None:
        self.model: Any = None
        self.params: Dict[str, Any] = {}

    def set_model(self, model: Any) -> None:
        self.model = model

    def set_params(self, params: Dict[str, Any]) -> None:
        self.params = params

    def on_train_begin(self, logs=None) -> None: ...
    def on_train_end(self, logs=None) -> None: ...
    def on_epoch_begin(self, epoch: int, logs=None) -> None: ...
    def on_epoch_end(self, epoch: int, logs=None) -> None: ...
    def on_train_batch_begin(self, batch: int, logs=None) -> None: ...
    def on_train_batch_end(self, batch: int, logs=None) -> None: ...

    def initial_epoch(self, requested: int) -> int:
        return requested


class CallbackList:
    def __init__(self, callbacks: Sequence[Callback], model: Any, verbose: Union[int, str], epochs: int,
                 steps: Optional[int]) -> None:
        self.callbacks = list(callbacks)
        if verbose and not any(isinstance(c, ProgbarLogger) for c in self.callbacks):
            self.callbacks.append(ProgbarLogger())
        params = {"epochs": epochs, "steps": steps, "verbose": verbose}
        for c in self.callbacks:
            c.set_model(model)
            c.set_params(params)
        # batch hooks cost a Python call per step; only dispatch to callbacks that override them
        self._batch_begin = [c for c in self.callbacks
                             if type(c).on_train_batch_begin is not Callback.on_train_batch_begin]
        self._batch_end = [c for c in self.callbacks if type(c).on_train_batch_end is not Callback.on_train_batch_end]

    @property
    def has_batch_hooks(self) -> bool:
        return bool(self._batch_begin or self._batch_end)

    @property
    def lag_ok(self) -> bool:
        return all(c.lag_ok for c in self.callbacks)

    def initial_epoch(self, requested: int) -> int:
        for c in self.callbacks:
            requested = c.initial_epoch(requested)
        return requested

    def on_train_begin(self, logs=None) -> None:
        for c in self.callbacks:
            c.on_train_begin(logs)

    def on_train_end(self, logs=None) -> None:
        for c in self.callbacks:
            c.on_train_end(logs)

    def on_epoch_begin(self, epoch, logs=None) -> None:
        for c in self.callbacks:
            c.on_epoch_begin(epoch, logs)

    def on_epoch_end(self, epoch, logs=None) -> None:
        for c in self.callbacks:
            c.on_epoch_end(epoch, logs)

    def on_train_batch_begin(self, batch, logs=None) -> None:
        for c in self._batch_begin:
            c.on_train_batch_begin(batch, logs)

    def on_train_batch_end(self, batch, logs=None) -> None:
        for c in self._batch_end:
            c.on_train_batch_end(batch, logs)


class ProgbarLogger(Callback):
    """One line per epoch on every worker: `Epoch 3/10 - 20/20 - 0.1s - loss: ... - accuracy: ...`."""
    lag_ok = True

    def on_epoch_end(self, epoch: int, logs=None) -> None:
        logs = logs or {}
        vals = " - ".join(f"{k}: {v:.4f}" for k, v in logs.items() if not k.startswith("_"))
        dt = logs.get("_epoch_time", 0.0)
        steps = logs.get("_steps", self.params.get("steps"))
        ms = f" - {1000 * dt / steps:.2f}ms/step" if steps else ""
        print(f"Epoch {epoch + 1}/{self.params.get('epochs')} - {steps}/{self.params.get('steps') or steps} - "
              f"{dt:.2f}s{ms} - {vals}", flush=True)


class LambdaCallback(Callback):
    def __init__(self, on_epoch_end: Optional[Callable] = None, on_train_batch_end: Optional[Callable] = None,
                 on_epoch_begin: Optional[Callable] = None, on_train_begin: Optional[Callable] = None,
                 on_train_end: Optional[Callable] = None) -> None:
        super().__init__()
        self._f = {"ee": on_epoch_end, "be": on_train_batch_end, "eb": on_epoch_begin, "tb": on_train_begin,
                   "te": on_train_end}
        if on_train_batch_end is None:
            self.on_train_batch_end = Callback.on_train_batch_end.__get__(self)  # keep the fast path

    def on_epoch_end(self, epoch, logs=None):
        if self._f["ee"]:
            self._f["ee"](epoch, logs)

    def on_epoch_begin(self, epoch, logs=None):
        if self._f["eb"]:
            self._f["eb"](epoch, logs)

    def on_train_begin(self, logs=None):
        if self._f["tb"]:
            self._f["tb"](logs)

    def on_train_end(self, logs=None):
        if self._f["te"]:
            self._f["te"](logs)


class _BatchHook(Callback):
    def __init__(self, fn: Callable[[int], None]) -> None:
        super().__init__()
        self.fn = fn

    def on_train_batch_end(self, batch, logs=None):
        self.fn(batch)


class LearningRateScheduler(Callback):
    def __init__(self, schedule: Callable[..., float], verbose: int = 0) -> None:
        super().__init__()
        self.schedule, self.verbose = schedule, verbose

    def on_epoch_begin(self, epoch, logs=None):
        opt = self.model.optimizer
        try:
            lr = self.schedule(epoch, opt.learning_rate)
        except TypeError:
            lr = self.schedule(epoch)
        opt.learning_rate = float(lr)
        if hasattr(self.model, "_weights_changed"):
            eng = getattr(self.model, "_engine", None)
            if eng is not None and hasattr(eng, "hyperparams_changed"):
                eng.hyperparams_changed()


class EarlyStopping(Callback):
    def __init__(self, monitor: str = "loss", patience: int = 0, min_delta: float = 0.0, mode: str = "min") -> None:
        super().__init__()
        self.monitor, self.patience, self.min_delta, self.mode = monitor, patience, min_delta, mode
        self.best: Optional[float] = None
        self.wait = 0

    def on_epoch_end(self, epoch, logs=None):
        v = (logs or {}).get(self.monitor)
        if v is None:
            return
        better = self.best is None or (v < self.best - self.min_delta if self.mode == "min"
                                       else v > self.best + self.min_delta)
        if better:
            self.best, self.wait = v, 0
        else:
            self.wait += 1
            if self.wait > self.patience:
                self.model.stop_training = True


class ReduceLROnPlateau(Callback):
    """tf.keras.callbacks.ReduceLROnPlateau: multiply the learning rate by `factor` when `monitor` has not improved for
    `patience` epochs. Every worker sees the same (all-reduced) logs, so every worker makes the same decision."""

    def __init__(self, monitor: str = "val_loss", factor: float = 0.1, patience: int = 10, min_delta: float = 1e-4,
                 cooldown: int = 0, min_lr: float = 0.0, mode: str = "min", verbose: int = 0) -> None:
        super().__init__()
        if not 0.0 < factor < 1.0:
            raise ValueError("ReduceLROnPlateau needs 0 < factor < 1")
        self.monitor, self.factor, self.patience, self.min_delta = monitor, factor, patience, min_delta
        self.cooldown, self.min_lr, self.mode, self.verbose = cooldown, min_lr, mode, verbose
        self.best: Optional[float] = None
        self.wait = 0
        self.cooldown_left = 0

    def on_epoch_end(self, epoch, logs=None):
        logs = logs if logs is not None else {}
        opt = self.model.optimizer
        logs["lr"] = opt.learning_rate
        v = logs.get(self.monitor)
        if v is None:
            return
        if self.cooldown_left > 0:
            self.cooldown_left -= 1
            self.wait = 0
        better = self.best is None or (v < self.best - self.min_delta if self.mode == "min" else v > self.best + self.min_delta)
        if better:
            self.best, self.wait = v, 0
            return
        if self.cooldown_left > 0:
            return
        self.wait += 1
        if self.wait >= self.patience and opt.learning_rate > self.min_lr:
            new_lr = max(opt.learning_rate * self.factor, self.min_lr)
            opt.learning_rate = new_lr
            self.cooldown_left, self.wait = self.cooldown, 0
            if self.verbose:
                print(f"Epoch {epoch + 1}: ReduceLROnPlateau reducing learning rate to {new_lr:g}.")


class ModelCheckpoint(Callback):
    """Saves weights (or, with save_weights_only=False, the whole model) at epoch end. Every worker takes part (the save
    is a collective point), only the chief writes to `filepath`; other workers write nothing (README.md:51)."""

    def __init__(self, filepath: str, save_weights_only: bool = True, save_freq: Union[str, int] = "epoch",
                 monitor: str = "loss", save_best_only: bool = False, mode: str = "min", verbose: int = 0) -> None:
        super().__init__()
        self.filepath, self.save_freq, self.save_weights_only = filepath, save_freq, save_weights_only
        self.monitor, self.save_best_only, self.mode, self.verbose = monitor, save_best_only, mode, verbose
        self.best: Optional[float] = None
        self.saved: List[str] = []
        if isinstance(save_freq, int):
            self.on_train_batch_end = self._batch_end  # type: ignore[method-assign]

    def _path(self, epoch: int, logs: Dict[str, float]) -> str:
        try:
            return self.filepath.format(epoch=epoch + 1, **{k: v for k, v in logs.items() if not k.startswith("_")})
        except (KeyError, IndexError):
            return self.filepath

    def _save(self, epoch: int, logs: Dict[str, float]) -> None:
        if self.save_best_only:
            v = logs.get(self.monitor)
            if v is None:
                return
            if self.best is not None and not (v < self.best if self.mode == "min" else v > self.best):
                return
            self.best = v
        path = self._path(epoch, logs)
        if self.save_weights_only:
            self.model.save_weights(path)
        else:
            self.model.save(path)  # architecture + compile arguments + variables (keras/models.py)
        if self.model.strategy.is_chief:
            self.saved.append(path)
            if self.verbose:
                print(f"Epoch {epoch + 1}: saved weights to {path}")

    def on_epoch_end(self, epoch, logs=None):
        if self.save_freq == "epoch":
            self._save(epoch, logs or {})

    def _batch_end(self, batch, logs=None):
        if self.model._steps_done % int(self.save_freq) == 0:
            self._save(0, logs or {})


class TensorBoard(Callback):
    """Chief-only scalar summaries in TensorBoard's event-file format (native writer)."""
    lag_ok = True

    def __init__(self, log_dir: str = "logs", update_freq: Union[str, int] = "epoch", **_: Any) -> None:
        super().__init__()
        self.log_dir = log_dir
        self._writer: Any = None

    def on_train_begin(self, logs=None):
        if self.model.strategy.is_chief:
            from ..runtime.native import EventWriter

            self._writer = EventWriter(os.path.join(self.log_dir, "train"))

    def on_epoch_end(self, epoch, logs=None):
        if self._writer is None:
            return
        for k, v in (logs or {}).items():
            if k.startswith("_"):
                continue
            self._writer.scalar(f"epoch_{k}", v, epoch)
        if logs and logs.get("_epoch_time") and logs.get("_steps"):
            self._writer.scalar("steps_per_second", logs["_steps"] / logs["_epoch_time"], epoch)
        self._writer.flush()

    def on_train_end(self, logs=None):
        if self._writer is not None:
            self._writer.close()
            self._writer = None


class BackupAndRestore(Callback):
    """Fault tolerance the way TF does it for a fixed-membership cluster: there is no elastic
    resize; when a worker dies everyone is restarted and training resumes from the last backup.
    The chief writes weights + optimizer slots + epoch counter at every epoch end; on_train_begin
    restores them on all workers and fit() continues from the next epoch."""

    def __init__(self, backup_dir: str, delete_checkpoint: bool = True) -> None:
        super().__init__()
        self.backup_dir, self.delete_checkpoint = backup_dir, delete_checkpoint
        self._resume_epoch: Optional[int] = None

    def _prefix(self) -> str:
        return os.path.join(self.backup_dir, "chief", "backup")

    def on_train_begin(self, logs=None):
        from ..checkpoint import load_training_state

        st = load_training_state(self.model, self._prefix())
        self._resume_epoch = st["epoch"] + 1 if st is not None else None

    def initial_epoch(self, requested: int) -> int:
        return max(requested, self._resume_epoch) if self._resume_epoch is not None else requested

    def on_epoch_end(self, epoch, logs=None):
        from ..checkpoint import save_training_state

        save_training_state(self.model, self._prefix(), epoch)

    def on_train_end(self, logs=None):
        self.model.strategy.barrier()
        if self.delete_checkpoint and self.model.strategy.is_chief and not self.model.stop_training:
            shutil.rmtree(os.path.join(self.backup_dir, "chief"), ignore_errors=True)


class CSVLogger(Callback):
    lag_ok = True

    def __init__(self, filename: str) -> None:
        super().__init__()
        self.filename = filename
        self._keys: Optional[List[str]] = None

    def on_epoch_end(self, epoch, logs=None):
        if not self.model.strategy.is_chief:
            return
        logs = {k: v for k, v in (logs or {}).items() if not k.startswith("_")}
        new = self._keys is None
        if new:
            self._keys = sorted(logs)
        with open(self.filename, "w" if new else "a") as f:
            if new:
                f.write(",".join(["epoch"] + self._keys) + "\n")
            f.write(",".join([str(epoch)] + [repr(logs.get(k, "")) for k in self._keys]) + "\n")


class FaultInjection(Callback):
    """Test hook: kill or stall THIS worker at a chosen global step.
    Also driven by the environment: TFDL_FAULT="kill:<rank>:<step>" or "stall:<rank>:<step>:<seconds>"."""

    def __init__(self, spec: Optional[str] = None) -> None:
        super().__init__()
        self.spec = spec if spec is not None else os.environ.get("TFDL_FAULT", "")

    def on_train_batch_end(self, batch, logs=None):
        if not self.spec:
            return
        parts = self.spec.split(":")
        kind, rank, step = parts[0], int(parts[1]), int(parts[2])
        if self.model.strategy.worker_index != rank or self.model._steps_done != step:
            return
        if kind == "kill":
            sys.stderr.write(f"[tfdl fault injection] worker {rank} exits at step {step}\n")
            sys.stderr.flush()
            os._exit(43)
        if kind == "stall":
            time.sleep(float(parts[3]) if len(parts) > 3 else 5.0)
