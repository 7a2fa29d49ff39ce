"""The `evaluator` task of a TF_CONFIG cluster (ref: /root/reference/README.md:38-44,57: an "evaluator" job that does the
validation on the side). It is not a replica of the mirrored strategies - they refuse to run in that role - so it never
takes part in a collective: it watches the directory the chief writes its checkpoints to, loads each new one into its own
copy of the model and evaluates it (tf.keras.utils.SidecarEvaluator).

    resolver = tfdl.cluster.TFConfigClusterResolver.from_env()
    if resolver.task_type == "evaluator":
        model = build_and_compile_cnn_model()          # same architecture, default (single-device) strategy
        tfdl.keras.utils.SidecarEvaluator(model, val_data, checkpoint_dir, log_dir=...).start()
"""
from __future__ import annotations

import os
import time
from typing import Any, Dict, List, Optional, Sequence

from . import callbacks as _cb


class SidecarEvaluator:
    """Evaluate every new checkpoint that appears in `checkpoint_dir`.

    `steps`: batches per evaluation (None = the whole dataset). Only the newest checkpoint is looked at, so one that is
    overwritten before the next poll is skipped (as in TensorFlow). Stops after `max_evaluations` checkpoints, or once a
    checkpoint whose optimizer step counter has reached `until_step` (the training run's total steps) has been evaluated,
    or when `timeout_s` passes without a new checkpoint; with none of them it runs until interrupted. Results are kept in
    `history` (one dict per checkpoint, with the checkpoint path and the optimizer's step counter), handed to
    `callbacks` as epoch-end logs, and written as TensorBoard scalars under `log_dir/validation` when given."""

    def __init__(self, model: Any, data: Any, checkpoint_dir: str, steps: Optional[int] = None,
                 max_evaluations: Optional[int] = None, callbacks: Optional[Sequence[_cb.Callback]] = None,
                 log_dir: Optional[str] = None, poll_interval_s: float = 1.0, timeout_s: Optional[float] = None,
                 until_step: Optional[int] = None) -> None:
        if model.loss is None:
            raise ValueError("SidecarEvaluator needs a compiled model (loss and metrics)")
        self.model, self.data, self.checkpoint_dir, self.steps = model, data, checkpoint_dir, steps
        self.max_evaluations, self.callbacks = max_evaluations, list(callbacks or [])
        self.log_dir, self.poll_interval_s, self.timeout_s = log_dir, poll_interval_s, timeout_s
        self.until_step = until_step
        self.history: List[Dict[str, Any]] = []
        self._seen: Optional[str] = None

    def _complete(self, prefix: str) -> bool:
        return os.path.exists(prefix + ".index") and os.path.exists(prefix + ".data-00000-of-00001")

    def evaluate_once(self) -> Optional[Dict[str, Any]]:
        """Evaluate the newest checkpoint if it has not been evaluated yet; returns its logs, else None."""
        from ..checkpoint import latest_checkpoint

        latest = latest_checkpoint(self.checkpoint_dir)
        if latest is None or latest == self._seen or not self._complete(latest):
            return None
        self.model.load_weights(latest)
        logs = self.model.evaluate(self.data, steps=self.steps, verbose=0, return_dict=True)
        self._seen = latest
        step = int(getattr(self.model.optimizer, "iterations", 0) or 0)
        out = dict(logs, checkpoint=latest, step=step)
        self.history.append(out)
        return out

    def start(self) -> List[Dict[str, Any]]:
        writer = None
        if self.log_dir:
            from ..runtime.native import EventWriter

            writer = EventWriter(os.path.join(self.log_dir, "validation"))
        for c in self.callbacks:
            c.set_model(self.model)
            c.on_train_begin()
        last_new = time.time()
        try:
            while self.max_evaluations is None or len(self.history) < self.max_evaluations:
                out = self.evaluate_once()
                if out is None:
                    if self.timeout_s is not None and time.time() - last_new > self.timeout_s:
                        break
                    time.sleep(self.poll_interval_s)
                    continue
                last_new = time.time()
                scalars = {k: v for k, v in out.items() if k not in ("checkpoint", "step")}
                if writer is not None:
                    for k, v in scalars.items():
                        writer.scalar(f"evaluation_{k}_vs_iterations", float(v), out["step"])
                    writer.flush()
                for c in self.callbacks:
                    c.on_epoch_end(len(self.history) - 1, dict(scalars))
                if getattr(self.model, "stop_training", False):
                    break
                if self.until_step is not None and out["step"] >= self.until_step:
                    break  # that was the training run's last checkpoint
        finally:
            for c in self.callbacks:
                c.on_train_end()
            if writer is not None:
                writer.close()
        return self.history
