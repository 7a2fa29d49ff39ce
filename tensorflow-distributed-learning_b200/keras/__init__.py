"""tf.keras look-alike - ref: /root/reference/tf_dist_example.py:39-59."""
from . import callbacks, layers, losses, metrics, optimizers
from .engine import History, Model, Sequential, TorchEngine, mixed_precision
from .layers import clear_session
from . import models  # noqa: E402  (needs engine)


class backend:  # tf.keras.backend look-alike
    clear_session = staticmethod(clear_session)


__all__ = ["callbacks", "layers", "losses", "metrics", "optimizers", "History", "Model", "Sequential", "TorchEngine",
           "mixed_precision", "backend", "clear_session", "models"]
