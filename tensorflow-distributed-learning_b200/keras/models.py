"""Whole-model persistence: tf.keras.models.save_model / load_model / model_from_json for this framework.

The reference's text has the chief "save checkpoint models" (/root/reference/README.md:51); `ModelCheckpoint` and
`save_weights` cover the variables, this module adds the rest of what `model.save()` keeps in Keras: the architecture
(layer classes and constructor arguments) and the `compile()` arguments, so that a job can be resumed - or a model
served - from the directory alone:

    <path>/model.json                       architecture, compile arguments, format version
    <path>/variables/variables.index|.data  TensorFlow-format checkpoint (Keras object-graph keys, optimizer slots)

The variables file is the same TensorBundle `save_weights` writes (csrc/runtime/bundle.cc), so
`tf.keras.Model.load_weights(<path>/variables/variables)` on the reference's model reads it too.
"""
from __future__ import annotations

import importlib
import json
import os
import shutil
from typing import Any, Dict, Optional

from . import layers as _layers
from . import losses as _losses
from . import metrics as _metrics
from . import optimizers as _optimizers
from .engine import Model, Sequential

FORMAT = "tfdl-keras-model/1"


def resolve_class(class_name: str, module: Optional[str], custom_objects: Optional[Dict[str, Any]], base: type) -> type:
    """custom_objects first, then this package's layers / model families, then the module recorded at save time."""
    if custom_objects and class_name in custom_objects:
        return custom_objects[class_name]
    for home in (_layers, ):
        c = getattr(home, class_name, None)
        if isinstance(c, type) and issubclass(c, base):
            return c
    if class_name == "Sequential":
        return Sequential
    if module:
        try:
            mod_name = module.replace("tensorflow-distributed-learning_b200", "tfdl")
            c = getattr(importlib.import_module(mod_name), class_name, None)
            if isinstance(c, type) and issubclass(c, base):
                return c
        except ImportError:
            pass
    raise ValueError(f"unknown {base.__name__} class {class_name!r}; pass it in custom_objects")


def _compile_config(model: Model, include_optimizer: bool) -> Optional[Dict[str, Any]]:
    if model.optimizer is None and model.loss is None:
        return None
    out: Dict[str, Any] = {"metrics": [m.name for m in model.compiled_metrics]}
    if model.loss is not None:
        out["loss"] = {"class_name": type(model.loss).__name__, "config": model.loss.get_config()}
    if include_optimizer and model.optimizer is not None:
        out["optimizer"] = dict(model.optimizer.get_config(), iterations=int(model.optimizer.iterations))
    return out


def model_to_json(model: Model, include_optimizer: bool = True) -> str:
    cls = type(model)
    return json.dumps({"format": FORMAT, "class_name": cls.__name__, "module": cls.__module__,
                       "config": model.get_config(), "compile": _compile_config(model, include_optimizer),
                       "compute_dtype": str(model.compute_dtype).replace("torch.", "")}, indent=1)


def model_from_json(text: str, custom_objects: Optional[Dict[str, Any]] = None, compile: bool = True) -> Model:  # noqa: A002
    """Rebuild (and optionally compile) a model from `Model.to_json()` under the current strategy scope."""
    spec = json.loads(text)
    if spec.get("format") != FORMAT:
        raise ValueError(f"not a {FORMAT} description (format = {spec.get('format')!r})")
    cls = resolve_class(spec["class_name"], spec.get("module"), custom_objects, Model)
    model = cls.from_config(dict(spec["config"]), custom_objects) if cls is Sequential or issubclass(cls, Sequential) \
        else cls.from_config(dict(spec["config"]))
    comp = spec.get("compile")
    if compile and comp:
        loss = None
        if comp.get("loss"):
            loss_cls = resolve_class(comp["loss"]["class_name"], _losses.__name__, custom_objects, _losses.Loss) \
                if not hasattr(_losses, comp["loss"]["class_name"]) else getattr(_losses, comp["loss"]["class_name"])
            loss = loss_cls.from_config(comp["loss"]["config"])
        opt_cfg = dict(comp.get("optimizer") or {"name": "SGD"})
        iterations = int(opt_cfg.pop("iterations", 0))
        opt = _optimizers.from_config(opt_cfg)
        model.compile(optimizer=opt, loss=loss, metrics=[_metrics.get(m, loss) for m in comp.get("metrics", [])])
        opt.iterations = iterations
    return model


def save_model(model: Model, filepath: str, overwrite: bool = True, include_optimizer: bool = True) -> None:
    """Every worker calls this (saving can involve collectives); only the chief's files stay (README.md:51)."""
    chief = model.strategy.is_chief
    if chief:
        if os.path.exists(os.path.join(filepath, "model.json")) and not overwrite:
            raise FileExistsError(f"{filepath} already holds a saved model (overwrite=False)")
        os.makedirs(os.path.join(filepath, "variables"), exist_ok=True)
    model.save_weights(os.path.join(filepath, "variables", "variables"))
    if chief:
        tmp = os.path.join(filepath, "model.json.tmp")
        with open(tmp, "w") as f:
            f.write(model_to_json(model, include_optimizer))
        os.replace(tmp, os.path.join(filepath, "model.json"))


def load_model(filepath: str, custom_objects: Optional[Dict[str, Any]] = None, compile: bool = True) -> Model:  # noqa: A002
    """tf.keras.models.load_model: architecture, compile arguments, variables and optimizer slots. Call it inside
    `strategy.scope()` to get mirrored variables, like building a model."""
    path = os.path.join(filepath, "model.json")
    if not os.path.exists(path):
        raise FileNotFoundError(f"{filepath} is not a saved model directory (no model.json)")
    with open(path) as f:
        model = model_from_json(f.read(), custom_objects, compile)
    if not model.built:
        raise ValueError("the saved model was never built (no input shape): build it, then load_weights")
    model.load_weights(os.path.join(filepath, "variables", "variables"))
    return model


def remove_model(filepath: str) -> None:
    """Delete a saved-model directory written by save_model."""
    if os.path.exists(os.path.join(filepath, "model.json")):
        shutil.rmtree(filepath)


__all__ = ["Sequential", "Model", "save_model", "load_model", "model_from_json", "model_to_json", "remove_model"]
