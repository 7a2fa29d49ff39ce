"""Model zoo: the reference CNN plus the larger configs named in BASELINE.json."""
from .mnist_cnn import build_and_compile_cnn_model, build_cnn_model

__all__ = ["build_and_compile_cnn_model", "build_cnn_model"]
