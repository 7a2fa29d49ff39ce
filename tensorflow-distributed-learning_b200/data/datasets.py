"""tfds-style dataset loader (ref: /root/reference/tf_dist_example.py:15,27-31).

    datasets, info = tfdl.datasets.load(name='mnist', with_info=True, as_supervised=True)
    train = datasets['train']

MNIST is read from IDX files (optionally .gz) under $TFDL_DATA_DIR/mnist or
~/tensorflow_datasets/mnist when they exist. There is no network on the build or GPU boxes, so
when the files are absent a deterministic *synthetic* MNIST-shaped set is generated instead:
uint8 images [N,28,28,1], int64 labels [N] in 0..9, with a class-dependent stroke pattern plus
noise so that a model can actually learn it (loss goes down, accuracy goes up). `info.synthetic`
says which one you got.
"""
from __future__ import annotations

import gzip
import os
import struct
from dataclasses import dataclass, field
from typing import Dict, Optional, Tuple, Union

import numpy as np

from .dataset import Dataset, _ArrayBase, _TensorSliceDataset

_progress_bar = [True]


def disable_progress_bar() -> None:
    """tfds.disable_progress_bar() - there is no download here, so this only records the wish."""
    _progress_bar[0] = False


@dataclass
class SplitInfo:
    num_examples: int


@dataclass
class DatasetInfo:
    name: str
    splits: Dict[str, SplitInfo]
    features: Dict[str, Tuple[Tuple[int, ...], str]]
    num_classes: int
    synthetic: bool
    supervised_keys: Tuple[str, str] = ("image", "label")
    description: str = ""
    extra: Dict[str, object] = field(default_factory=dict)


def _read_idx(path: str) -> np.ndarray:
    op = gzip.open if path.endswith(".gz") else open
    with op(path, "rb") as f:
        magic = struct.unpack(">I", f.read(4))[0]
        dtype_code, ndim = (magic >> 8) & 0xFF, magic & 0xFF
        dims = struct.unpack(">" + "I" * ndim, f.read(4 * ndim))
        dt = {0x08: np.uint8, 0x09: np.int8, 0x0B: ">i2", 0x0C: ">i4", 0x0D: ">f4", 0x0E: ">f8"}[dtype_code]
        return np.frombuffer(f.read(), dtype=dt).reshape(dims)


def _find(dirpath: str, stem: str) -> Optional[str]:
    for name in (stem, stem + ".gz", stem.replace("-idx", ".idx"), stem.replace("-idx", ".idx") + ".gz"):
        p = os.path.join(dirpath, name)
        if os.path.exists(p):
            return p
    return None


def _mnist_from_files(data_dir: str) -> Optional[Dict[str, Tuple[np.ndarray, np.ndarray]]]:
    files = {
        "train": ("train-images-idx3-ubyte", "train-labels-idx1-ubyte"),
        "test": ("t10k-images-idx3-ubyte", "t10k-labels-idx1-ubyte"),
    }
    out = {}
    for split, (fi, fl) in files.items():
        pi, pl = _find(data_dir, fi), _find(data_dir, fl)
        if pi is None or pl is None:
            return None
        out[split] = (_read_idx(pi)[..., None].astype(np.uint8), _read_idx(pl).astype(np.int64))
    return out


def synthetic_mnist(n: int, seed: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """Learnable MNIST-shaped data: each class lights a distinct pair of bars, plus noise."""
    rng = np.random.default_rng(seed)
    labels = rng.integers(0, 10, size=n, dtype=np.int64)
    img = rng.integers(0, 64, size=(n, 28, 28), dtype=np.int64)
    templ = np.zeros((10, 28, 28), dtype=np.int64)
    for c in range(10):
        r = 2 + 2 * c
        templ[c, r:r + 3, 4:24] = 160
        col = 3 + 2 * ((c * 7) % 10)
        templ[c, 4:24, col:col + 2] = 120
    img = np.clip(img + templ[labels], 0, 255).astype(np.uint8)
    return img[..., None], labels


def load(name: str = "mnist", split: Union[None, str] = None, with_info: bool = False,
         as_supervised: bool = False, data_dir: Optional[str] = None, synthetic_size: Tuple[int, int] = (60000, 10000),
         seed: int = 0, **_: object):
    """tfds.load look-alike. Returns a dict of splits (or one split if `split` is given),
    and a DatasetInfo when with_info=True."""
    if name != "mnist":
        raise ValueError(f"dataset {name!r} is not bundled; use Dataset.from_tensor_slices for your own data")
    data_dir = data_dir or os.environ.get("TFDL_DATA_DIR") or os.path.expanduser("~/tensorflow_datasets")
    found = _mnist_from_files(os.path.join(data_dir, "mnist")) or _mnist_from_files(data_dir)
    synthetic = found is None
    if synthetic:
        found = {"train": synthetic_mnist(synthetic_size[0], seed), "test": synthetic_mnist(synthetic_size[1], seed + 1)}
    splits: Dict[str, Dataset] = {}
    for k, (x, y) in found.items():
        tensors = (x, y) if as_supervised else {"image": x, "label": y}
        base = _ArrayBase(tensors, num_files=1, file_bounds=np.zeros(1, dtype=np.int64))
        splits[k] = _TensorSliceDataset(base)
    info = DatasetInfo(
        name="mnist",
        splits={k: SplitInfo(len(v[1])) for k, v in found.items()},
        features={"image": ((28, 28, 1), "uint8"), "label": ((), "int64")},
        num_classes=10, synthetic=synthetic,
        description="MNIST (IDX files)" if not synthetic else "synthetic MNIST-shaped data (no network on this box)")
    out = splits if split is None else splits[split]
    return (out, info) if with_info else out
