"""B200-native synchronous data-parallel training framework.

Public surface mirrors what /root/reference/tf_dist_example.py touches in TensorFlow:

    tfdl.distribute   MirroredStrategy, MultiWorkerMirroredStrategy, CollectiveCommunication
    tfdl.data         Dataset, Options, AutoShardPolicy
    tfdl.keras        Sequential, layers, losses, optimizers, metrics, callbacks
    tfdl.cluster      TF_CONFIG resolver
    tfdl.datasets     tfds-style loader (mnist: IDX files or synthetic)
    tfdl.ops          ctypes bindings of the sm_100a kernels
    tfdl.config       set_random_seed, enable_op_determinism
"""
__version__ = "0.1.0"

import importlib as _importlib

_LAZY = {
    "cluster": ".cluster",
    "data": ".data",
    "distribute": ".parallel",
    "parallel": ".parallel",
    "keras": ".keras",
    "ops": ".ops",
    "models": ".models",
    "utils": ".utils",
    "checkpoint": ".checkpoint",
    "runtime": ".runtime",
    "datasets": ".data.datasets",
    "config": ".config",
}


def __getattr__(name):
    if name in _LAZY:
        mod = _importlib.import_module(_LAZY[name], __name__)
        globals()[name] = mod
        return mod
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
