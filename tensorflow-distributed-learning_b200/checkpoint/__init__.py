"""Chief-only checkpoints in the Keras TF-checkpoint (TensorBundle) format, or .npz.

The reference only says that the chief saves checkpoints (README.md:51); the north star asks that
names and shapes load into the reference's Keras model (tf_dist_example.py:40-48). So a checkpoint
written here is what `model.save_weights(prefix)` writes in TF 2.x:

  <prefix>.index / <prefix>.data-00000-of-00001   TensorBundle (native writer, csrc/runtime/bundle.cc)
  keys   layer_with_weights-<i>/kernel/.ATTRIBUTES/VARIABLE_VALUE      (HWIO conv kernels, (in,out) dense)
         layer_with_weights-<i>/bias/.ATTRIBUTES/VARIABLE_VALUE
         optimizer/iter|learning_rate|momentum|decay/.ATTRIBUTES/VARIABLE_VALUE
         layer_with_weights-<i>/kernel/.OPTIMIZER_SLOT/optimizer/momentum/.ATTRIBUTES/VARIABLE_VALUE
         _CHECKPOINTABLE_OBJECT_GRAPH                                    (serialized TrackableObjectGraph)
  `checkpoint`                                     text state file naming the latest prefix

Arena storage (OHWI / [out,in]) is permuted back to the Keras layout on save and forth on load.
Under a multi-worker strategy every worker calls save (it is a barrier), only the chief writes.
TensorFlow is not installable on this box, so compatibility is validated with an independent
pure-Python reader in tests/test_checkpoint.py, not against TF itself (flagged in BASELINE.md).
"""
from __future__ import annotations

import os
from typing import Any, Dict, List, Optional, Tuple

import numpy as np

_ATTR = "/.ATTRIBUTES/VARIABLE_VALUE"
_SHORT = {"kernel": "kernel", "bias": "bias", "gamma": "gamma", "beta": "beta", "moving_mean": "moving_mean",
          "moving_variance": "moving_variance", "embeddings": "embeddings"}


def _var_local_name(var_name: str) -> str:
    # "conv2d/kernel:0" -> "kernel"
    return var_name.split("/")[-1].split(":")[0]


def _weight_layers(model: Any) -> List[Any]:
    return [l for l in model.layers if l.weights]


def keras_checkpoint_keys(model: Any) -> List[Tuple[str, Any]]:
    """(checkpoint key, variable) for every model variable, Keras object-graph naming."""
    out = []
    for i, layer in enumerate(_weight_layers(model)):
        for w in layer.weights:
            out.append((f"layer_with_weights-{i}/{_var_local_name(w.name)}{_ATTR}", w))
    return out


def _object_graph(model: Any, slot_keys: Dict[str, List[Tuple[str, str]]], opt_keys: List[Tuple[str, str]]) -> bytes:
    from tensorboard.compat.proto import trackable_object_graph_pb2 as pb

    g = pb.TrackableObjectGraph()
    root = g.nodes.add()
    var_node: Dict[str, int] = {}
    layer_idx = 0
    for li, layer in enumerate(model.layers):
        ln = g.nodes.add()
        lid = len(g.nodes) - 1
        root.children.add(node_id=lid, local_name=f"layer-{li}")
        if layer.weights:
            root.children.add(node_id=lid, local_name=f"layer_with_weights-{layer_idx}")
            for w in layer.weights:
                vn = g.nodes.add()
                vid = len(g.nodes) - 1
                local = _var_local_name(w.name)
                # re-fetch: protobuf repeated-field adds may invalidate earlier references
                g.nodes[lid].children.add(node_id=vid, local_name=local)
                key = f"layer_with_weights-{layer_idx}/{local}{_ATTR}"
                g.nodes[vid].attributes.add(name="VARIABLE_VALUE", full_name=w.name.split(":")[0], checkpoint_key=key)
                var_node[key] = vid
            layer_idx += 1
    if opt_keys:
        g.nodes.add()
        oid = len(g.nodes) - 1
        g.nodes[0].children.add(node_id=oid, local_name="optimizer")
        for local, key in opt_keys:
            g.nodes.add()
            hid = len(g.nodes) - 1
            g.nodes[oid].children.add(node_id=hid, local_name=local)
            g.nodes[hid].attributes.add(name="VARIABLE_VALUE", full_name=f"{model.optimizer.name}/{local}",
                                        checkpoint_key=key)
        for var_key, slots in slot_keys.items():
            for slot_name, skey in slots:
                g.nodes.add()
                sid = len(g.nodes) - 1
                g.nodes[sid].attributes.add(name="VARIABLE_VALUE", full_name=f"{model.optimizer.name}/{slot_name}",
                                            checkpoint_key=skey)
                g.nodes[oid].slot_variables.add(original_variable_node_id=var_node[var_key], slot_name=slot_name,
                                                slot_variable_node_id=sid)
    return g.SerializeToString()


def _slot_view(model: Any, slot: str, w: Any) -> np.ndarray:
    arena = model.optimizer.slots(slot)[0]
    sp = w.spec
    a = arena[sp.offset:sp.offset + sp.numel].view(sp.storage_shape).detach().cpu().numpy()
    return a if sp.to_keras is None else np.ascontiguousarray(np.transpose(a, sp.to_keras))


def _assign_slot(model: Any, slot: str, w: Any, a: np.ndarray) -> None:
    import torch

    sp = w.spec
    st = a if sp.to_storage is None else np.transpose(a, sp.to_storage)
    t = torch.from_numpy(np.ascontiguousarray(st, dtype=np.float32)).reshape(-1)
    for arena in model.optimizer.slots(slot):
        arena[sp.offset:sp.offset + sp.numel].copy_(t)


def _write_bundle(model: Any, prefix: str, include_optimizer: bool, extra: Optional[Dict[str, np.ndarray]] = None) -> None:
    from ..runtime.native import BundleWriter

    w = BundleWriter(prefix)
    keys = keras_checkpoint_keys(model)
    for key, var in keys:
        w.add(key, var.numpy())
    opt_keys: List[Tuple[str, str]] = []
    slot_keys: Dict[str, List[Tuple[str, str]]] = {}
    opt = model.optimizer if include_optimizer else None
    if opt is not None:
        hp: Dict[str, np.ndarray] = {"iter": np.array(opt.iterations, np.int64),
                                     "learning_rate": np.array(opt.learning_rate, np.float32),
                                     "decay": np.array(0.0, np.float32)}
        if hasattr(opt, "momentum"):
            hp["momentum"] = np.array(opt.momentum, np.float32)
        for n in ("beta_1", "beta_2"):
            if hasattr(opt, n):
                hp[n] = np.array(getattr(opt, n), np.float32)
        for local, val in hp.items():
            key = f"optimizer/{local}{_ATTR}"
            w.add(key, val)
            opt_keys.append((local, key))
        if opt._slots:
            for key, var in keys:
                if not var.trainable:
                    continue
                base = key[:-len(_ATTR)]
                for slot in opt.slot_names():
                    skey = f"{base}/.OPTIMIZER_SLOT/optimizer/{slot}{_ATTR}"
                    w.add(skey, _slot_view(model, slot, var))
                    slot_keys.setdefault(key, []).append((slot, skey))
    for k, v in (extra or {}).items():
        w.add(k, v)
    w.add_string("_CHECKPOINTABLE_OBJECT_GRAPH", _object_graph(model, slot_keys, opt_keys))
    w.finish()
    d = os.path.dirname(os.path.abspath(prefix))
    name = os.path.basename(prefix)
    tmp = os.path.join(d, f".checkpoint.tmp{os.getpid()}")
    with open(tmp, "w") as f:
        f.write(f'model_checkpoint_path: "{name}"\nall_model_checkpoint_paths: "{name}"\n')
    os.replace(tmp, os.path.join(d, "checkpoint"))


def _sync_slots(model: Any) -> None:
    """Every worker takes part in a save: optimizer slots that the sliced NVLink all-reduce paths keep
    distributed over the ranks are made whole before the chief reads them."""
    sync = getattr(model.strategy.cross_device_ops, "sync_optimizer_slots", None)
    if sync is not None and model.optimizer is not None and getattr(model.optimizer, "_slots", None):
        sync()


def save_model_weights(model: Any, filepath: str, save_format: Optional[str] = None) -> None:
    """Chief writes; everyone synchronises afterwards so no worker races ahead of the file."""
    model._finalize()
    strategy = model.strategy
    fmt = save_format or ("npz" if filepath.endswith(".npz") else "tf")
    _sync_slots(model)
    if strategy.is_chief:
        if fmt == "npz":
            names = [w.name for w in model.weights]
            tmp = filepath + f".tmp{os.getpid()}.npz"
            np.savez(tmp, __names__=np.array(names), **{f"w{i}": w.numpy() for i, w in enumerate(model.weights)})
            os.replace(tmp, filepath)
        else:
            _write_bundle(model, filepath, include_optimizer=model.optimizer is not None)
    strategy.barrier()


def load_model_weights(model: Any, filepath: str) -> None:
    model._finalize()
    if filepath.endswith(".npz"):
        z = np.load(filepath)
        ws = [z[f"w{i}"] for i in range(len(model.weights))]
        model.set_weights(ws)
        return
    from ..runtime.native import BundleReader

    r = BundleReader(filepath)
    try:
        for key, var in keras_checkpoint_keys(model):
            if key not in r.entries:
                raise KeyError(f"checkpoint {filepath} has no entry {key} (for variable {var.name})")
            if tuple(r.shape(key)) != tuple(var.shape):
                raise ValueError(f"{key}: checkpoint shape {r.shape(key)} does not match variable shape {var.shape}")
            var.assign(r.read(key))
        opt = model.optimizer
        if opt is not None:
            ik = f"optimizer/iter{_ATTR}"
            if ik in r.entries:
                opt.iterations = int(r.read(ik))
            if opt.slot_names():
                if not opt._slots:
                    opt.bind([a.params for a in model.arenas])
                for key, var in keras_checkpoint_keys(model):
                    base = key[:-len(_ATTR)]
                    for slot in opt.slot_names():
                        skey = f"{base}/.OPTIMIZER_SLOT/optimizer/{slot}{_ATTR}"
                        if skey in r.entries and var.trainable:
                            _assign_slot(model, slot, var, r.read(skey))
    finally:
        r.close()


def save_training_state(model: Any, prefix: str, epoch: int) -> None:
    """Backup for BackupAndRestore: weights + optimizer + progress counters."""
    model._finalize()
    s = model.strategy
    _sync_slots(model)
    if s.is_chief:
        ds = getattr(model, "_data_state", None) or {"shuffle": [], "batches": 0}
        shuf = np.array(ds["shuffle"], np.int64).reshape(-1, 2)
        _write_bundle(model, prefix, include_optimizer=True,
                      extra={"_tfdl/epoch": np.array(epoch, np.int64),
                             "_tfdl/steps_done": np.array(model._steps_done, np.int64),
                             # the data iterator's position: seeds of the current pass + steps handed out
                             "_tfdl/iter_batches": np.array(ds["batches"], np.int64),
                             "_tfdl/iter_shuffle": shuf if shuf.size else np.zeros((1, 2), np.int64) - 1})
    s.barrier()


def load_training_state(model: Any, prefix: str) -> Optional[Dict[str, int]]:
    if not os.path.exists(prefix + ".index"):
        return None
    if not model.built:
        return None
    load_model_weights(model, prefix)
    from ..runtime.native import BundleReader

    r = BundleReader(prefix)
    try:
        st = {"epoch": int(r.read("_tfdl/epoch")), "steps_done": int(r.read("_tfdl/steps_done"))}
        if "_tfdl/iter_batches" in r.entries:
            shuf = np.asarray(r.read("_tfdl/iter_shuffle")).reshape(-1, 2)
            shuf = [] if (shuf.size and shuf[0, 0] == -1 and shuf[0, 1] == -1) else [(int(a), int(b)) for a, b in shuf]
            model._resume_data = {"batches": int(r.read("_tfdl/iter_batches")), "shuffle": shuf}
    finally:
        r.close()
    model._steps_done = st["steps_done"]
    model._weights_changed()
    return st


def latest_checkpoint(directory: str) -> Optional[str]:
    """tf.train.latest_checkpoint: reads the `checkpoint` state file."""
    p = os.path.join(directory, "checkpoint")
    if not os.path.exists(p):
        return None
    with open(p) as f:
        for line in f:
            if line.startswith("model_checkpoint_path:"):
                return os.path.join(directory, line.split(":", 1)[1].strip().strip('"'))
    return None
