"""tf.data-style input pipeline.

Covers every Dataset call the reference makes (/root/reference/tf_dist_example.py:20-37,
README.md:113-129): from_tensor_slices, map, cache, shuffle, batch, with_options,
Options().experimental_distribute.auto_shard_policy, AutoShardPolicy.{OFF,AUTO,FILE,DATA} -
plus repeat/take/skip/shard/prefetch/unbatch/rebatch which the strategies need.

Design (not tf.data's): a pipeline is split into an *element stage* and a *batch stage*.
Element-stage datasets never materialise Python objects per element; they expose
  (a) a random-access `base` that gathers many elements at once (`gather(int64 indices)`), and
  (b) an *index stream* (chunks of int64 indices) that shuffle / shard / take / skip / repeat
      transform lazily.
`batch()` pulls `batch_size` indices at a time and performs ONE vectorised gather per batch, so a
60k-image epoch is ~470 numpy fancy-index calls rather than 60k Python iterations. The shuffle
buffer (tf.data semantics: a sliding reservoir of `buffer_size` elements) runs natively in
libtfdl_runtime.so when it is built, else in numpy.
Elements are numpy arrays (or tuples/dicts of them); conversion to torch happens at the strategy
boundary where batches are staged into pinned memory.
"""
from __future__ import annotations

import enum
import itertools
import threading
import queue
from typing import Any, Callable, Iterable, Iterator, List, Optional, Sequence, Tuple, Union

import numpy as np

# --------------------------------------------------------------------------------------------
# options
# --------------------------------------------------------------------------------------------


class AutoShardPolicy(enum.IntEnum):
    """How a dataset is split between workers (tf.data.experimental.AutoShardPolicy).

    OFF  - no sharding: every worker iterates the whole dataset (the reference's choice,
           tf_dist_example.py:35; workers then differ only by their shuffle order).
    AUTO - FILE if the source is file based with >= one file per worker, else DATA.
    FILE - each worker reads a disjoint subset of the source files.
    DATA - each worker reads everything and keeps elements with index % num_workers == id.
    """

    OFF = -1
    AUTO = 0
    FILE = 1
    DATA = 2


class DistributeOptions:
    def __init__(self) -> None:
        self.auto_shard_policy: AutoShardPolicy = AutoShardPolicy.AUTO
        self.num_devices: Optional[int] = None


class ThreadingOptions:
    def __init__(self) -> None:
        self.private_threadpool_size: Optional[int] = None
        self.max_intra_op_parallelism: Optional[int] = None


class Options:
    """tf.data.Options: only the fields the reference touches carry meaning."""

    def __init__(self) -> None:
        self.experimental_distribute = DistributeOptions()
        self.experimental_threading = ThreadingOptions()
        self.experimental_deterministic: Optional[bool] = None

    @property
    def distribute(self) -> DistributeOptions:  # newer TF spelling
        return self.experimental_distribute

    def merge(self, other: "Options") -> "Options":
        out = Options()
        out.experimental_distribute.auto_shard_policy = other.experimental_distribute.auto_shard_policy
        out.experimental_distribute.num_devices = (other.experimental_distribute.num_devices
                                                   or self.experimental_distribute.num_devices)
        out.experimental_deterministic = (other.experimental_deterministic
                                          if other.experimental_deterministic is not None
                                          else self.experimental_deterministic)
        return out


# --------------------------------------------------------------------------------------------
# structure helpers
# --------------------------------------------------------------------------------------------

def _map_structure(fn: Callable[[Any], Any], s: Any) -> Any:
    if isinstance(s, tuple):
        return tuple(_map_structure(fn, x) for x in s)
    if isinstance(s, list):
        return [_map_structure(fn, x) for x in s]
    if isinstance(s, dict):
        return {k: _map_structure(fn, v) for k, v in s.items()}
    return fn(s)


def _flatten(s: Any) -> List[Any]:
    if isinstance(s, (tuple, list)):
        return [y for x in s for y in _flatten(x)]
    if isinstance(s, dict):
        return [y for k in sorted(s) for y in _flatten(s[k])]
    return [s]


def _to_numpy(x: Any) -> np.ndarray:
    if isinstance(x, np.ndarray):
        return x
    try:
        import torch

        if isinstance(x, torch.Tensor):
            return x.detach().cpu().numpy()
    except ImportError:  # pragma: no cover
        pass
    return np.asarray(x)


def _call(fn: Callable, elem: Any) -> Any:
    return fn(*elem) if isinstance(elem, tuple) else fn(elem)


def _leading(s: Any) -> int:
    return int(_flatten(s)[0].shape[0])


# --------------------------------------------------------------------------------------------
# random-access bases (element stage)
# --------------------------------------------------------------------------------------------


class _Base:
    """Random-access element store."""

    def __len__(self) -> int:  # pragma: no cover - interface
        raise NotImplementedError

    def gather(self, idx: np.ndarray) -> Any:  # pragma: no cover - interface
        raise NotImplementedError

    num_files: int = 0  # > 0 for file-backed sources (FILE sharding)

    def file_of(self, idx: np.ndarray) -> np.ndarray:  # pragma: no cover - file sources only
        raise NotImplementedError


class _ArrayBase(_Base):
    def __init__(self, arrays: Any, num_files: int = 0, file_bounds: Optional[np.ndarray] = None) -> None:
        self.arrays = _map_structure(_to_numpy, arrays)
        flat = _flatten(self.arrays)
        if not flat:
            raise ValueError("from_tensor_slices needs at least one tensor")
        n = {int(a.shape[0]) if a.ndim else -1 for a in flat}
        if len(n) != 1 or -1 in n:
            raise ValueError(f"all tensors must share their first dimension, got {[a.shape for a in flat]}")
        self._n = n.pop()
        self.num_files = num_files
        self._file_bounds = file_bounds

    def __len__(self) -> int:
        return self._n

    def gather(self, idx: np.ndarray) -> Any:
        if idx.size and idx[-1] - idx[0] == idx.size - 1 and np.all(np.diff(idx) == 1):
            lo, hi = int(idx[0]), int(idx[-1]) + 1
            return _map_structure(lambda a: a[lo:hi], self.arrays)  # contiguous: a view, no copy
        return _map_structure(lambda a: a[idx], self.arrays)

    def file_of(self, idx: np.ndarray) -> np.ndarray:
        return np.searchsorted(self._file_bounds, idx, side="right") - 1


class _MappedBase(_Base):
    """fn applied to gathered batches. `vectorized` says fn accepts a leading batch dim."""

    def __init__(self, parent: _Base, fn: Callable, vectorized: bool) -> None:
        self.parent, self.fn, self.vectorized = parent, fn, vectorized
        self.num_files = parent.num_files

    def __len__(self) -> int:
        return len(self.parent)

    def gather(self, idx: np.ndarray) -> Any:
        batch = self.parent.gather(idx)
        if self.vectorized:
            return _map_structure(_to_numpy, _call(self.fn, batch))
        outs = []
        for i in range(idx.size):
            elem = _map_structure(lambda a: a[i], batch)
            outs.append(_map_structure(_to_numpy, _call(self.fn, elem)))
        return _stack(outs)

    def file_of(self, idx: np.ndarray) -> np.ndarray:
        return self.parent.file_of(idx)


def _stack(elems: Sequence[Any]) -> Any:
    first = elems[0]
    if isinstance(first, tuple):
        return tuple(_stack([e[i] for e in elems]) for i in range(len(first)))
    if isinstance(first, list):
        return [_stack([e[i] for e in elems]) for i in range(len(first))]
    if isinstance(first, dict):
        return {k: _stack([e[k] for e in elems]) for k in first}
    return np.stack([np.asarray(e) for e in elems])


def _probe_vectorized(fn: Callable, base: _Base) -> bool:
    """Decide whether `fn` can be applied to a whole batch at once: run it on element 0 and on
    the batch [0, 1] and compare. Anything that raises or disagrees falls back to per-element."""
    n = len(base)
    if n == 0:
        return True
    try:
        idx = np.arange(min(2, n), dtype=np.int64)
        batch = base.gather(idx)
        single = _call(fn, _map_structure(lambda a: a[0], batch))
        multi = _call(fn, batch)
        fs, fm = _flatten(_map_structure(_to_numpy, single)), _flatten(_map_structure(_to_numpy, multi))
        if len(fs) != len(fm):
            return False
        for s, m in zip(fs, fm):
            if m.shape != (idx.size,) + s.shape or m.dtype != s.dtype:
                return False
            if not np.array_equal(m[0], s, equal_nan=True):
                return False
        return True
    except Exception:
        return False


# --------------------------------------------------------------------------------------------
# shuffle buffer over an index stream
# --------------------------------------------------------------------------------------------

def _shuffle_indices_numpy(src: np.ndarray, buffer_size: int, rng: np.random.Generator) -> np.ndarray:
    """tf.data shuffle semantics on indices: keep a buffer of `buffer_size`, emit a uniformly
    random slot, refill the slot from the stream; drain the buffer at the end."""
    n = src.size
    if buffer_size >= n:
        return src[rng.permutation(n)]
    out = np.empty(n, dtype=np.int64)
    buf = src[:buffer_size].copy()
    picks = rng.integers(0, buffer_size, size=n - buffer_size)
    for t in range(n - buffer_size):
        j = picks[t]
        out[t] = buf[j]
        buf[j] = src[buffer_size + t]
    out[n - buffer_size:] = buf[rng.permutation(buffer_size)]
    return out


def _shuffle_indices(src: np.ndarray, buffer_size: int, seed: int) -> np.ndarray:
    try:
        from ..runtime import native as _native

        if _native.available():
            return _native.shuffle_indices(src, buffer_size, seed)
    except Exception:
        pass
    return _shuffle_indices_numpy(src, buffer_size, np.random.default_rng(seed))


# --------------------------------------------------------------------------------------------
# Dataset
# --------------------------------------------------------------------------------------------

_UNKNOWN = -2
_INFINITE = -1


class Dataset:
    """Base class. Subclasses implement `_elements()`; element-stage subclasses additionally
    implement `_base()` / `_index_chunks()`."""

    _options: Optional[Options] = None
    _parent: Optional["Dataset"] = None

    # ---- construction -----------------------------------------------------------------
    @staticmethod
    def from_tensor_slices(tensors: Any) -> "Dataset":
        """Slices along the first dimension (README.md:123-128: NumPy arrays -> Dataset)."""
        return _TensorSliceDataset(_ArrayBase(tensors))

    @staticmethod
    def from_tensors(tensors: Any) -> "Dataset":
        return _TensorSliceDataset(_ArrayBase(_map_structure(lambda a: _to_numpy(a)[None], tensors)))

    @staticmethod
    def range(*args: int) -> "Dataset":
        return _TensorSliceDataset(_ArrayBase(np.arange(*args, dtype=np.int64)))

    @staticmethod
    def from_generator(gen: Callable[[], Iterable[Any]]) -> "Dataset":
        return _GeneratorDataset(gen)

    # ---- transformations ----------------------------------------------------------------
    def map(self, fn: Callable, num_parallel_calls: Any = None, deterministic: Any = None,
            vectorized: Optional[bool] = None) -> "Dataset":
        return _MapDataset(self, fn, vectorized)

    def cache(self, filename: str = "") -> "Dataset":
        return _CacheDataset(self)

    def shuffle(self, buffer_size: int, seed: Optional[int] = None,
                reshuffle_each_iteration: bool = True) -> "Dataset":
        return _ShuffleDataset(self, int(buffer_size), seed, reshuffle_each_iteration)

    def batch(self, batch_size: int, drop_remainder: bool = False, **_: Any) -> "Dataset":
        return _BatchDataset(self, int(batch_size), bool(drop_remainder))

    def unbatch(self) -> "Dataset":
        return _UnbatchDataset(self)

    def rebatch(self, batch_size: int, drop_remainder: bool = False) -> "Dataset":
        return self.unbatch().batch(batch_size, drop_remainder)

    def repeat(self, count: Optional[int] = None) -> "Dataset":
        return _RepeatDataset(self, -1 if count is None else int(count))

    def take(self, count: int) -> "Dataset":
        return _TakeSkipDataset(self, int(count), take=True)

    def skip(self, count: int) -> "Dataset":
        return _TakeSkipDataset(self, int(count), take=False)

    def shard(self, num_shards: int, index: int) -> "Dataset":
        if not 0 <= index < num_shards:
            raise ValueError(f"shard index {index} out of range for {num_shards} shards")
        return _ShardDataset(self, int(num_shards), int(index))

    def prefetch(self, buffer_size: int = 1) -> "Dataset":
        return _PrefetchDataset(self, max(1, int(buffer_size) if buffer_size and buffer_size > 0 else 2))

    def with_options(self, options: Options) -> "Dataset":
        return _OptionsDataset(self, options)

    def apply(self, fn: Callable[["Dataset"], "Dataset"]) -> "Dataset":
        return fn(self)

    def enumerate(self, start: int = 0) -> "Dataset":
        return _GeneratorDataset(lambda: ((np.int64(i), e) for i, e in enumerate(self, start)))

    def concatenate(self, other: "Dataset") -> "Dataset":
        return _GeneratorDataset(lambda: itertools.chain(iter(self), iter(other)))

    def filter(self, predicate: Callable[..., Any]) -> "Dataset":
        """Elements for which `predicate(*element)` is true (a generic, element-at-a-time stage)."""
        def keep(e: Any) -> bool:
            return bool(np.asarray(predicate(*e) if isinstance(e, tuple) else predicate(e)))

        return _GeneratorDataset(lambda: (e for e in self if keep(e)))

    @staticmethod
    def zip(*datasets: Any) -> "Dataset":
        """tf.data.Dataset.zip((a, b)) / zip(a, b): tuples of corresponding elements, as long as the shortest input."""
        if len(datasets) == 1 and isinstance(datasets[0], (tuple, list)):
            datasets = tuple(datasets[0])
        return _GeneratorDataset(lambda: zip(*[iter(d) for d in datasets]))

    def reduce(self, initial_state: Any, reduce_func: Callable[[Any, Any], Any]) -> Any:
        state = initial_state
        for e in self:
            state = reduce_func(state, e)
        return state

    # ---- introspection --------------------------------------------------------------------
    def options(self) -> Options:
        node, chain = self, []
        while node is not None:
            if node._options is not None:
                chain.append(node._options)
            node = node._parent
        out = Options()
        for o in reversed(chain):
            out = out.merge(o)
        return out

    def cardinality(self) -> int:
        return _UNKNOWN

    def __len__(self) -> int:
        c = self.cardinality()
        if c < 0:
            raise TypeError("dataset length is infinite or unknown")
        return c

    def __iter__(self) -> Iterator[Any]:
        return self._elements()

    def as_numpy_iterator(self) -> Iterator[Any]:
        return iter(self)

    @property
    def element_spec(self) -> Any:
        for e in self:
            return _map_structure(lambda a: (np.asarray(a).shape, np.asarray(a).dtype), e)
        return None

    # ---- subclass hooks ---------------------------------------------------------------------
    def _elements(self) -> Iterator[Any]:  # pragma: no cover - interface
        raise NotImplementedError

    def _is_indexable(self) -> bool:
        return False

    def _base(self) -> _Base:
        raise TypeError(f"{type(self).__name__} is not an element-stage dataset")

    def _index_chunks(self) -> Iterator[np.ndarray]:
        raise TypeError(f"{type(self).__name__} is not an element-stage dataset")

    def _identity_count(self) -> Optional[int]:
        """n when the index stream of this (indexable) dataset is exactly 0..n-1, else None."""
        return None

    def _with_parent(self, parent: "Dataset") -> "Dataset":
        """Shallow copy of this node on top of a different parent (pipeline rewrites)."""
        import copy

        c = copy.copy(self)
        c._parent = parent
        return c


class _IndexableDataset(Dataset):
    _CHUNK = 8192

    def _is_indexable(self) -> bool:
        return True

    def _elements(self) -> Iterator[Any]:
        base = self._base()
        for chunk in self._index_chunks():
            if chunk.size == 0:
                continue
            batch = base.gather(chunk)
            for i in range(chunk.size):
                yield _map_structure(lambda a: a[i], batch)


class _TensorSliceDataset(_IndexableDataset):
    def __init__(self, base: _Base) -> None:
        self._b = base

    def _base(self) -> _Base:
        return self._b

    def _index_chunks(self) -> Iterator[np.ndarray]:
        n = len(self._b)
        for lo in range(0, n, self._CHUNK):
            yield np.arange(lo, min(n, lo + self._CHUNK), dtype=np.int64)

    def _identity_count(self) -> Optional[int]:
        return len(self._b)

    def cardinality(self) -> int:
        return len(self._b)


class _GeneratorDataset(Dataset):
    def __init__(self, gen: Callable[[], Iterable[Any]]) -> None:
        self._gen = gen

    def _elements(self) -> Iterator[Any]:
        for e in self._gen():
            yield _map_structure(_to_numpy, e)


class _MapDataset(Dataset):
    def __init__(self, parent: Dataset, fn: Callable, vectorized: Optional[bool]) -> None:
        self._parent, self._fn, self._vec = parent, fn, vectorized
        self._mapped: Optional[_MappedBase] = None

    def _is_indexable(self) -> bool:
        return self._parent._is_indexable()

    def _base(self) -> _Base:
        if self._mapped is None or self._mapped.parent is not self._parent._base():
            pb = self._parent._base()
            vec = self._vec if self._vec is not None else _probe_vectorized(self._fn, pb)
            self._mapped = _MappedBase(pb, self._fn, vec)
        return self._mapped

    def _index_chunks(self) -> Iterator[np.ndarray]:
        return self._parent._index_chunks()

    def _identity_count(self) -> Optional[int]:
        return self._parent._identity_count() if self._is_indexable() else None

    def _elements(self) -> Iterator[Any]:
        if self._is_indexable():
            return _IndexableDataset._elements(self)  # type: ignore[arg-type]
        return (_map_structure(_to_numpy, _call(self._fn, e)) for e in self._parent)

    def cardinality(self) -> int:
        return self._parent.cardinality()


class _CacheDataset(_IndexableDataset):
    """First complete pass materialises the parent's elements (in the parent's order) in host
    memory; later passes are served from that copy - the parent pipeline (e.g. the scale map of
    tf_dist_example.py:22-31) never runs again."""

    def __init__(self, parent: Dataset) -> None:
        self._parent = parent
        self._cached: Optional[_ArrayBase] = None
        self._lock = threading.Lock()

    def _materialise(self) -> _ArrayBase:
        with self._lock:
            if self._cached is None:
                p = self._parent
                if p._is_indexable():
                    chunks = list(p._index_chunks())
                    idx = np.concatenate(chunks) if chunks else np.zeros(0, np.int64)
                    pb = p._base()
                    data = pb.gather(idx)
                    data = _map_structure(np.ascontiguousarray, data)
                    bounds = None
                    if pb.num_files and idx.size and np.all(np.diff(idx) >= 0):
                        f = pb.file_of(idx)
                        bounds = np.searchsorted(f, np.arange(pb.num_files), side="left")
                    self._cached = _ArrayBase(data, pb.num_files if bounds is not None else 0, bounds)
                else:
                    elems = list(p)
                    if not elems:
                        raise ValueError("cannot cache an empty dataset")
                    self._cached = _ArrayBase(_stack(elems))
            return self._cached

    def _base(self) -> _Base:
        return self._materialise()

    def _index_chunks(self) -> Iterator[np.ndarray]:
        n = len(self._materialise())
        for lo in range(0, n, self._CHUNK):
            yield np.arange(lo, min(n, lo + self._CHUNK), dtype=np.int64)

    def _identity_count(self) -> Optional[int]:
        return len(self._materialise())

    def cardinality(self) -> int:
        c = self._parent.cardinality()
        return c if c >= 0 else len(self._materialise())


AUTOTUNE = -1  # tf.data.AUTOTUNE: accepted wherever TensorFlow takes it (map / prefetch); this pipeline sizes itself

_global_seed = [None]


def set_global_seed(seed: Optional[int]) -> None:
    _global_seed[0] = seed


class _ShuffleDataset(Dataset):
    def __init__(self, parent: Dataset, buffer_size: int, seed: Optional[int], reshuffle: bool) -> None:
        if buffer_size < 1:
            raise ValueError("buffer_size must be >= 1")
        self._parent, self._buf, self._seed, self._reshuffle = parent, buffer_size, seed, reshuffle
        self._epoch = 0
        self._entropy = int(np.random.SeedSequence().entropy % (1 << 62))

    def _next_seed(self) -> int:
        base = self._seed if self._seed is not None else (
            _global_seed[0] if _global_seed[0] is not None else self._entropy)
        e = self._epoch if self._reshuffle else 0
        self._epoch += 1
        return int((base * 1000003 + e * 7919 + 12345) % (1 << 62))

    def _is_indexable(self) -> bool:
        return self._parent._is_indexable()

    def _base(self) -> _Base:
        return self._parent._base()

    def _index_chunks(self) -> Iterator[np.ndarray]:
        seed = self._next_seed()
        ident = self._parent._identity_count()
        if ident is not None:
            src: Any = ident  # the parent's stream is 0..n-1: nothing to materialise
        else:
            chunks = list(self._parent._index_chunks())
            src = np.concatenate(chunks) if chunks else np.zeros(0, np.int64)
        try:
            from ..runtime import native as _native

            stream = _native.shuffle_stream(src, self._buf, seed) if _native.available() else None
        except Exception:
            stream = None
        if stream is None and ident is not None:
            src = np.arange(ident, dtype=np.int64)
        if stream is not None:  # produced on demand: a 20-step fit() draws 1,280 indices, not all 60,000
            yield from stream
            return
        out = _shuffle_indices(src, self._buf, seed)
        step = _IndexableDataset._CHUNK
        for lo in range(0, out.size, step):
            yield out[lo:lo + step]

    def _elements(self) -> Iterator[Any]:
        if self._is_indexable():
            return _IndexableDataset._elements(self)  # type: ignore[arg-type]
        return self._generic()

    def _generic(self) -> Iterator[Any]:
        rng = np.random.default_rng(self._next_seed())
        buf: List[Any] = []
        for e in self._parent:
            if len(buf) < self._buf:
                buf.append(e)
                continue
            j = int(rng.integers(0, self._buf))
            yield buf[j]
            buf[j] = e
        for j in rng.permutation(len(buf)):
            yield buf[int(j)]

    def cardinality(self) -> int:
        return self._parent.cardinality()


class _ShardDataset(Dataset):
    def __init__(self, parent: Dataset, n: int, i: int) -> None:
        self._parent, self._n, self._i = parent, n, i

    def _is_indexable(self) -> bool:
        return self._parent._is_indexable()

    def _base(self) -> _Base:
        return self._parent._base()

    def _index_chunks(self) -> Iterator[np.ndarray]:
        pos = 0
        for c in self._parent._index_chunks():
            first = (self._i - pos) % self._n
            yield c[first::self._n]
            pos += c.size

    def _elements(self) -> Iterator[Any]:
        if self._is_indexable():
            return _IndexableDataset._elements(self)  # type: ignore[arg-type]
        return itertools.islice(iter(self._parent), self._i, None, self._n)

    def cardinality(self) -> int:
        c = self._parent.cardinality()
        return c if c < 0 else (c - self._i + self._n - 1) // self._n


class _FileShardDataset(_ShardDataset):
    """FILE policy: keep elements whose source file index % n == i."""

    def _index_chunks(self) -> Iterator[np.ndarray]:
        base = self._parent._base()
        for c in self._parent._index_chunks():
            yield c[(base.file_of(c) % self._n) == self._i]

    def cardinality(self) -> int:
        return _UNKNOWN


class _TakeSkipDataset(Dataset):
    def __init__(self, parent: Dataset, count: int, take: bool) -> None:
        self._parent, self._count, self._take = parent, count, take

    def _is_indexable(self) -> bool:
        return self._parent._is_indexable()

    def _base(self) -> _Base:
        return self._parent._base()

    def _index_chunks(self) -> Iterator[np.ndarray]:
        left = self._count
        for c in self._parent._index_chunks():
            if self._take:
                if self._count >= 0 and left <= 0:
                    return
                if self._count >= 0 and c.size > left:
                    c = c[:left]
                left -= c.size
                yield c
            else:
                if left >= c.size:
                    left -= c.size
                    continue
                yield c[left:]
                left = 0

    def _elements(self) -> Iterator[Any]:
        if self._is_indexable():
            return _IndexableDataset._elements(self)  # type: ignore[arg-type]
        it = iter(self._parent)
        if self._take:
            return itertools.islice(it, None if self._count < 0 else self._count)
        return itertools.islice(it, self._count, None)

    def cardinality(self) -> int:
        c = self._parent.cardinality()
        if self._take:
            if self._count < 0:
                return c
            return self._count if c == _INFINITE else (c if c < 0 else min(c, self._count))
        return c if c < 0 else max(0, c - self._count)


class _RepeatDataset(Dataset):
    def __init__(self, parent: Dataset, count: int) -> None:
        self._parent, self._count = parent, count

    def _is_indexable(self) -> bool:
        return self._parent._is_indexable()

    def _base(self) -> _Base:
        return self._parent._base()

    def _index_chunks(self) -> Iterator[np.ndarray]:
        it = itertools.count() if self._count < 0 else range(self._count)
        for _ in it:
            empty = True
            for c in self._parent._index_chunks():
                empty = empty and c.size == 0
                yield c
            if empty:
                return

    def _elements(self) -> Iterator[Any]:
        if self._is_indexable():
            return _IndexableDataset._elements(self)  # type: ignore[arg-type]
        return self._generic()

    def _generic(self) -> Iterator[Any]:
        it = itertools.count() if self._count < 0 else range(self._count)
        for _ in it:
            empty = True
            for e in self._parent:
                empty = False
                yield e
            if empty:
                return

    def cardinality(self) -> int:
        c = self._parent.cardinality()
        if self._count < 0:
            return _INFINITE if c != 0 else 0
        return c if c < 0 else c * self._count


class _BatchDataset(Dataset):
    """Element stage -> batch stage. One vectorised gather per batch on the fast path."""

    def __init__(self, parent: Dataset, batch_size: int, drop_remainder: bool) -> None:
        if batch_size < 1:
            raise ValueError("batch_size must be >= 1")
        self._parent, self._bs, self._drop = parent, batch_size, drop_remainder

    @property
    def batch_size(self) -> int:
        return self._bs

    def index_batches(self) -> Iterator[np.ndarray]:
        """The row numbers of every batch, without gathering anything (indexable parents only). Consumers that can
        gather for themselves - the native executor fills its pinned ring straight from the dataset arrays - use this
        instead of iterating the dataset."""
        p = self._parent
        pending: List[np.ndarray] = []
        have = 0
        for c in p._index_chunks():
            pending.append(c)
            have += c.size
            if have < self._bs:
                continue
            buf = np.concatenate(pending) if len(pending) > 1 else pending[0]
            lo = 0
            while buf.size - lo >= self._bs:
                yield buf[lo:lo + self._bs]
                lo += self._bs
            pending, have = ([buf[lo:]], buf.size - lo) if lo < buf.size else ([], 0)
        if have and not self._drop:
            yield np.concatenate(pending)

    def _elements(self) -> Iterator[Any]:
        p = self._parent
        if p._is_indexable():
            base = p._base()
            for idx in self.index_batches():
                yield base.gather(idx)
            return
        buf_e: List[Any] = []
        for e in p:
            buf_e.append(e)
            if len(buf_e) == self._bs:
                yield _stack(buf_e)
                buf_e = []
        if buf_e and not self._drop:
            yield _stack(buf_e)

    def cardinality(self) -> int:
        c = self._parent.cardinality()
        if c < 0:
            return c
        return c // self._bs if self._drop else (c + self._bs - 1) // self._bs


class _UnbatchDataset(Dataset):
    def __init__(self, parent: Dataset) -> None:
        self._parent = parent

    def _elements(self) -> Iterator[Any]:
        for b in self._parent:
            for i in range(_leading(b)):
                yield _map_structure(lambda a: a[i], b)


class _PrefetchDataset(Dataset):
    """Background thread fills a bounded queue (overlaps host batch assembly with the step)."""

    def __init__(self, parent: Dataset, depth: int) -> None:
        self._parent, self._depth = parent, depth

    def _elements(self) -> Iterator[Any]:
        q: "queue.Queue[Any]" = queue.Queue(self._depth)
        stop = threading.Event()
        END, ERR = object(), object()

        def work() -> None:
            try:
                for e in self._parent:
                    while not stop.is_set():
                        try:
                            q.put(e, timeout=0.1)
                            break
                        except queue.Full:
                            continue
                    if stop.is_set():
                        return
                q.put(END)
            except BaseException as ex:  # propagate into the consumer
                q.put((ERR, ex))

        t = threading.Thread(target=work, daemon=True, name="tfdl-prefetch")
        t.start()
        try:
            while True:
                e = q.get()
                if e is END:
                    return
                if isinstance(e, tuple) and len(e) == 2 and e[0] is ERR:
                    raise e[1]
                yield e
        finally:
            stop.set()

    def cardinality(self) -> int:
        return self._parent.cardinality()


class _OptionsDataset(Dataset):
    def __init__(self, parent: Dataset, options: Options) -> None:
        if not isinstance(options, Options):
            raise TypeError("with_options expects a tfdl.data.Options")
        self._parent, self._options = parent, options

    def _is_indexable(self) -> bool:
        return self._parent._is_indexable()

    def _base(self) -> _Base:
        return self._parent._base()

    def _index_chunks(self) -> Iterator[np.ndarray]:
        return self._parent._index_chunks()

    def _identity_count(self) -> Optional[int]:
        return self._parent._identity_count() if self._is_indexable() else None

    def _elements(self) -> Iterator[Any]:
        return iter(self._parent)

    def cardinality(self) -> int:
        return self._parent.cardinality()


# --------------------------------------------------------------------------------------------
# pipeline rewrites used by the strategies
# --------------------------------------------------------------------------------------------

def index_stream(ds: Dataset) -> Optional[Tuple[_Base, Iterator[np.ndarray]]]:
    """(element store, iterator over the row numbers of each batch) when `ds` is a batch over an indexable pipeline
    with nothing but options / prefetch above it; None otherwise (the caller then iterates `ds` normally)."""
    node: Optional[Dataset] = ds
    while isinstance(node, (_OptionsDataset, _PrefetchDataset)):
        node = node._parent
    if not isinstance(node, _BatchDataset) or node._parent is None or not node._parent._is_indexable():
        return None
    return node._parent._base(), node.index_batches()


def shuffle_nodes(ds: Dataset) -> List["_ShuffleDataset"]:
    """The shuffle stages of a pipeline, outermost first (their seeds are what makes an iterator reproducible)."""
    out: List[_ShuffleDataset] = []
    node: Optional[Dataset] = ds
    while node is not None:
        if isinstance(node, _ShuffleDataset):
            out.append(node)
        node = getattr(node, "_parent", None)
    return out


def iterator_state(ds: Dataset) -> List[Tuple[int, int]]:
    """What a NEW iterator over `ds` would be seeded with: (entropy, pass counter) per shuffle stage. Saved with a
    training backup so that a restarted run can rebuild the same order and skip what was already consumed."""
    return [(int(n._entropy), int(n._epoch)) for n in shuffle_nodes(ds)]


def restore_iterator_state(ds: Dataset, state: Sequence[Tuple[int, int]]) -> None:
    nodes = shuffle_nodes(ds)
    if len(nodes) != len(state):
        raise ValueError(f"saved iterator state has {len(state)} shuffle stages, the dataset has {len(nodes)}")
    for n, (entropy, epoch) in zip(nodes, state):
        n._entropy, n._epoch = int(entropy), int(epoch)


def find_batch_node(ds: Dataset) -> Optional[_BatchDataset]:
    node: Optional[Dataset] = ds
    while node is not None:
        if isinstance(node, _BatchDataset):
            return node
        node = node._parent
    return None


def _rebuild_above(ds: Dataset, stop: Dataset, new_stop: Dataset) -> Dataset:
    """Re-create the chain of nodes from `ds` down to (excluding) `stop` on top of `new_stop`."""
    chain = []
    node: Optional[Dataset] = ds
    while node is not stop:
        assert node is not None
        chain.append(node)
        node = node._parent
    out = new_stop
    for n in reversed(chain):
        out = n._with_parent(out)
    return out


def auto_shard(ds: Dataset, num_workers: int, worker_index: int,
               policy: Optional[AutoShardPolicy] = None) -> Dataset:
    """Apply the dataset's auto-shard policy for `num_workers` workers
    (tf_dist_example.py:34-37 selects OFF; README.md:113-129 discusses why)."""
    policy = ds.options().experimental_distribute.auto_shard_policy if policy is None else policy
    if num_workers <= 1 or policy == AutoShardPolicy.OFF:
        return ds
    batch = find_batch_node(ds)
    target = batch._parent if batch is not None else ds
    assert target is not None
    file_ok = target._is_indexable() and target._base().num_files >= num_workers
    if policy == AutoShardPolicy.FILE and not file_ok:
        raise ValueError(
            "AutoShardPolicy.FILE needs a file-based source with at least one file per worker; "
            "use DATA or OFF (tf.data raises the same complaint)")
    use_file = policy == AutoShardPolicy.FILE or (policy == AutoShardPolicy.AUTO and file_ok)
    if use_file:
        sharded: Dataset = _FileShardDataset(target, num_workers, worker_index)
    else:
        # Every worker must run the same number of steps per epoch: a worker whose shard is one element longer would
        # enter one more gradient collective than its peers. With a known element count the tail that does not divide
        # by the number of workers (at most num_workers - 1 elements) is left out of the epoch, so all shards are equal.
        c = target.cardinality()
        even: Dataset = target
        if c >= 0 and c % num_workers:
            even = _TakeSkipDataset(target, c - c % num_workers, True)
        sharded = _ShardDataset(even, num_workers, worker_index)
    if batch is None:
        return sharded
    return _rebuild_above(ds, target, sharded)


def rebatch_for_replicas(ds: Dataset, num_replicas_in_sync: int) -> Tuple[Dataset, int]:
    """The dataset is batched by the GLOBAL batch size (tf_dist_example.py:18,33); the strategy
    re-batches it so each replica step consumes global/num_replicas elements. Returns the
    rewritten dataset (yielding per-replica batches) and the per-replica batch size."""
    batch = find_batch_node(ds)
    if batch is None:
        raise ValueError("distribute_dataset expects a batched dataset (call .batch(GLOBAL_BATCH_SIZE) first)")
    g = batch.batch_size
    if g % num_replicas_in_sync:
        raise ValueError(f"global batch size {g} is not divisible by {num_replicas_in_sync} replicas in sync")
    per = g // num_replicas_in_sync
    if num_replicas_in_sync == 1:
        return ds, per
    assert batch._parent is not None
    new_batch = _BatchDataset(batch._parent, per, batch._drop)
    return _rebuild_above(ds, batch, new_batch), per
