"""tf.keras look-alike - ref: /root/reference/tf_dist_example.py:39-59."""
from . import callbacks, layers, losses, metrics, optimizers
from .engine import History, Model, Sequential, TorchEngine, mixed_precision
from .layers import clear_session
from . import models  # noqa: E402  (needs engine)


class backend:  # tf.keras.backend look-alike
    clear_session = staticmethod(clear_session)


def Input(shape, name=None, **_):  # noqa: N802
    """tf.keras.Input for Sequential models: an InputLayer carrying the per-example shape."""
    return layers.InputLayer(tuple(shape), name=name)


layers.Input = Input  # tf.keras.layers.Input is the same function


class utils:  # tf.keras.utils look-alike
    @staticmethod
    def to_categorical(y, num_classes=None, dtype="float32"):
        import numpy as np

        y = np.asarray(y, dtype=np.int64).reshape(-1)
        n = int(num_classes) if num_classes is not None else int(y.max()) + 1
        out = np.zeros((y.size, n), dtype=dtype)
        out[np.arange(y.size), y] = 1
        return out

    @staticmethod
    def set_random_seed(seed):
        from .. import config

        config.set_random_seed(seed)


from .sidecar import SidecarEvaluator  # noqa: E402

utils.SidecarEvaluator = SidecarEvaluator


__all__ = ["callbacks", "layers", "losses", "metrics", "optimizers", "History", "Model", "Sequential", "TorchEngine",
           "mixed_precision", "backend", "clear_session", "models", "Input", "utils"]
