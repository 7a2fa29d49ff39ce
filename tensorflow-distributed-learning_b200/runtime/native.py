"""ctypes bindings of libtfdl_runtime.so (see include/tfdl_runtime.h)."""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from ..ops import _lib

_configured = False


def available() -> bool:
    return os.path.exists(os.path.join(_lib.native_dir(), "libtfdl_runtime.so"))


def lib() -> C.CDLL:
    global _configured
    L = _lib.runtime()
    if not _configured:
        vp, i64p = C.c_void_p, C.POINTER(C.c_int64)
        L.tfdl_rt_crc32c.restype = C.c_uint32
        L.tfdl_rt_crc32c.argtypes = [C.c_uint32, vp, C.c_uint64]
        L.tfdl_rt_crc32c_mask.restype = C.c_uint32
        L.tfdl_rt_crc32c_mask.argtypes = [C.c_uint32]
        L.tfdl_rt_shuffle_indices.argtypes = [vp, C.c_int64, C.c_int64, C.c_uint64, vp]
        L.tfdl_rt_gather_rows.argtypes = [vp, C.c_int64, C.c_int64, vp, C.c_int64, vp, C.c_int]
        L.tfdl_rt_shuffler_create.restype = vp
        L.tfdl_rt_shuffler_create.argtypes = [vp, C.c_int64, C.c_int64, C.c_uint64]
        L.tfdl_rt_shuffler_next.restype = C.c_int64
        L.tfdl_rt_shuffler_next.argtypes = [vp, vp, C.c_int64]
        L.tfdl_rt_shuffler_destroy.restype = None
        L.tfdl_rt_shuffler_destroy.argtypes = [vp]
        L.tfdl_rt_pack_records.argtypes = [vp, vp, C.c_int64, C.c_int64, C.c_int64, C.c_int, vp, C.c_int64, C.c_int]
        L.tfdl_rt_gather_rows_u8_to_f32.argtypes = [vp, C.c_int64, C.c_int64, vp, C.c_int64, C.c_float, vp, C.c_int]
        L.tfdl_rt_gather_records.argtypes = [vp, C.c_int64, C.c_int64, vp, C.c_int, vp, C.c_int64, C.c_int64, vp,
                                             C.c_int64, C.c_int]
        L.tfdl_rt_gather_records_begin.restype = vp
        L.tfdl_rt_gather_records_begin.argtypes = [vp, C.c_int64, C.c_int64, vp, C.c_int, vp, C.c_int64, C.c_int64, vp,
                                                   C.c_int64]
        L.tfdl_rt_gather_records_ready.restype = C.c_int64
        L.tfdl_rt_gather_records_ready.argtypes = [vp]
        L.tfdl_rt_gather_records_wait.argtypes = [vp, C.c_int64]
        L.tfdl_rt_gather_records_end.restype = None
        L.tfdl_rt_gather_records_end.argtypes = [vp]
        L.tfdl_rt_group_create.restype = vp
        L.tfdl_rt_group_create.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_char_p), C.c_int, C.c_int, C.c_int]
        L.tfdl_rt_group_destroy.argtypes = [vp]
        L.tfdl_rt_group_destroy.restype = None
        for f in ("tfdl_rt_group_rank", "tfdl_rt_group_world", "tfdl_rt_group_failed_rank", "tfdl_rt_barrier"):
            getattr(L, f).argtypes = [vp]
        L.tfdl_rt_group_set_abort_word.argtypes = [vp, vp]
        L.tfdl_rt_group_set_abort_word.restype = None
        L.tfdl_rt_group_set_timeout.argtypes = [vp, C.c_int]
        L.tfdl_rt_group_set_timeout.restype = None
        L.tfdl_rt_group_bytes_sent.argtypes = [vp]
        L.tfdl_rt_group_bytes_sent.restype = C.c_uint64
        L.tfdl_rt_allgather.argtypes = [vp, vp, C.c_uint64, vp]
        L.tfdl_rt_broadcast0.argtypes = [vp, vp, C.c_uint64]
        L.tfdl_rt_ring_allreduce_f32.argtypes = [vp, vp, C.c_int64]
        L.tfdl_rt_ring_allreduce_f64.argtypes = [vp, vp, C.c_int64]
        L.tfdl_rt_bundle_writer_create.restype = vp
        L.tfdl_rt_bundle_writer_create.argtypes = [C.c_char_p]
        L.tfdl_rt_bundle_writer_add.argtypes = [vp, C.c_char_p, C.c_int, C.c_int, i64p, vp, C.c_uint64]
        L.tfdl_rt_bundle_writer_add_string.argtypes = [vp, C.c_char_p, C.c_char_p, C.c_uint64]
        L.tfdl_rt_bundle_writer_finish.argtypes = [vp]
        L.tfdl_rt_bundle_writer_abort.argtypes = [vp]
        L.tfdl_rt_bundle_writer_abort.restype = None
        L.tfdl_rt_bundle_reader_open.restype = vp
        L.tfdl_rt_bundle_reader_open.argtypes = [C.c_char_p]
        L.tfdl_rt_bundle_reader_close.argtypes = [vp]
        L.tfdl_rt_bundle_reader_close.restype = None
        L.tfdl_rt_bundle_reader_num_entries.argtypes = [vp]
        L.tfdl_rt_bundle_reader_entry.argtypes = [vp, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_int),
                                                  C.POINTER(C.c_int), i64p, C.POINTER(C.c_uint64)]
        L.tfdl_rt_bundle_reader_read.argtypes = [vp, C.c_int, vp, C.c_uint64, C.POINTER(C.c_uint64)]
        L.tfdl_rt_events_open.restype = vp
        L.tfdl_rt_events_open.argtypes = [C.c_char_p, C.c_char_p]
        L.tfdl_rt_events_scalar.argtypes = [vp, C.c_char_p, C.c_double, C.c_int64]
        L.tfdl_rt_events_flush.argtypes = [vp]
        L.tfdl_rt_events_path.argtypes = [vp]
        L.tfdl_rt_events_path.restype = C.c_char_p
        L.tfdl_rt_events_close.argtypes = [vp]
        L.tfdl_rt_events_close.restype = None
        _configured = True
    return L


class CollectiveError(RuntimeError):
    """A collective failed (peer lost, timeout); the message names the rank."""


def _np_ptr(a: np.ndarray) -> C.c_void_p:
    return C.c_void_p(a.ctypes.data)


def crc32c(data: bytes, crc: int = 0) -> int:
    return int(lib().tfdl_rt_crc32c(crc, data, len(data)))


def masked_crc32c(data: bytes) -> int:
    return int(lib().tfdl_rt_crc32c_mask(crc32c(data)))


def shuffle_indices(src: np.ndarray, buffer_size: int, seed: int) -> np.ndarray:
    src = np.ascontiguousarray(src, dtype=np.int64)
    out = np.empty_like(src)
    _lib.check_rt(lib().tfdl_rt_shuffle_indices(_np_ptr(src), src.size, int(buffer_size), int(seed) & (2 ** 64 - 1),
                                               _np_ptr(out)), "shuffle_indices")
    return out


def shuffle_stream(src, buffer_size: int, seed: int, chunk: int = 8192):
    """Generator over chunks of the shuffled index stream (tf.data shuffle-buffer semantics), produced on demand.
    ``src``: the index stream to shuffle, or an int n for the identity stream 0..n-1 (nothing is materialised)."""
    if isinstance(src, (int, np.integer)):
        h = lib().tfdl_rt_shuffler_create(None, int(src), int(buffer_size), C.c_uint64(seed & ((1 << 64) - 1)))
    else:
        src = np.ascontiguousarray(src, dtype=np.int64)
        h = lib().tfdl_rt_shuffler_create(_np_ptr(src), src.size, int(buffer_size), C.c_uint64(seed & ((1 << 64) - 1)))
    if not h:
        raise RuntimeError(_lib.runtime_error() if hasattr(_lib, "runtime_error") else "shuffler_create failed")
    # bound now: a generator that is still open at interpreter exit is finalised after this module's globals are gone
    nxt, destroy, h = lib().tfdl_rt_shuffler_next, lib().tfdl_rt_shuffler_destroy, C.c_void_p(h)
    try:
        first = True
        while True:
            # a small first chunk: the consumer's first batches should not wait for 8192 draws
            out = np.empty(min(chunk, 2048) if first else chunk, dtype=np.int64)
            first = False
            n = int(nxt(h, C.c_void_p(out.ctypes.data), out.size))
            if n == 0:
                return
            yield out[:n]
    finally:
        destroy(h)


def gather_rows(src: np.ndarray, idx: np.ndarray, out: Optional[np.ndarray] = None, threads: int = 4) -> np.ndarray:
    src = np.ascontiguousarray(src)
    idx = np.ascontiguousarray(idx, dtype=np.int64)
    row = int(np.prod(src.shape[1:], dtype=np.int64)) * src.itemsize
    if out is None:
        out = np.empty((idx.size,) + src.shape[1:], dtype=src.dtype)
    _lib.check_rt(lib().tfdl_rt_gather_rows(_np_ptr(src), src.shape[0], row, _np_ptr(idx), idx.size, _np_ptr(out),
                                           threads), "gather_rows")
    return out


def pack_records(xs: list, ys: list, ring_ptr: int, rec_bytes: int, threads: int = 4) -> None:
    """Copy k (x, labels) batches into the record ring at ``ring_ptr`` ([x bytes | int32 labels] per record).
    Every x must be C-contiguous float32 of one size, every y C-contiguous int32 or int64 of one size."""
    k = len(xs)
    px = (C.c_void_p * k)(*[a.ctypes.data for a in xs])
    py = (C.c_void_p * k)(*[a.ctypes.data for a in ys])
    _lib.check_rt(lib().tfdl_rt_pack_records(px, py, k, xs[0].nbytes, ys[0].size, ys[0].dtype.itemsize, ring_ptr,
                                             rec_bytes, threads), "pack_records")


def gather_records(x: np.ndarray, y: np.ndarray, idx: np.ndarray, k: int, B: int, ring_ptr: int, rec_bytes: int,
                   threads: int = 4) -> None:
    """k records [x[idx[r*B:(r+1)*B]] | int32 labels] gathered from dataset arrays into the ring at ``ring_ptr``:
    shuffle + batch + staging of tf_dist_example.py:31-33 in one multi-threaded pass. x: C-contiguous rows of any
    dtype; y: C-contiguous uint8 / int32 / int64 vector; idx: k*B int64 row numbers."""
    row = int(np.prod(x.shape[1:], dtype=np.int64)) * x.itemsize
    _lib.check_rt(lib().tfdl_rt_gather_records(_np_ptr(x), x.shape[0], row, _np_ptr(y), y.dtype.itemsize, _np_ptr(idx),
                                              k, B, C.c_void_p(ring_ptr), rec_bytes, threads), "gather_records")


class GatherJob:
    """gather_records running on the runtime's worker pool while the caller does something else (the prefetch of the
    index path of fit()). ``ready()``: leading records complete; ``wait(n)``: returns once records [0, n) are in the
    ring (the calling thread copies unclaimed parts itself rather than sleep); ``end()``: complete and free. The job
    holds references to x and y; the ring must outlive it."""

    def __init__(self, x: np.ndarray, y: np.ndarray, idx: np.ndarray, k: int, B: int, ring_ptr: int, rec_bytes: int) -> None:
        row = int(np.prod(x.shape[1:], dtype=np.int64)) * x.itemsize
        self._keep = (x, y)
        self.k = k
        self._h = lib().tfdl_rt_gather_records_begin(_np_ptr(x), x.shape[0], row, _np_ptr(y), y.dtype.itemsize, _np_ptr(idx),
                                                     k, B, C.c_void_p(ring_ptr), rec_bytes)
        if not self._h:
            raise RuntimeError("gather_records: " + lib().tfdl_rt_last_error().decode())

    def ready(self) -> int:
        return int(lib().tfdl_rt_gather_records_ready(self._h))

    def wait(self, n: int) -> None:
        lib().tfdl_rt_gather_records_wait(self._h, n)

    def end(self) -> None:
        if self._h:
            lib().tfdl_rt_gather_records_end(self._h)
            self._h = None
            self._keep = ()

    def __del__(self) -> None:  # pragma: no cover
        try:
            self.end()
        except Exception:
            pass


def gather_rows_u8_to_f32(src: np.ndarray, idx: np.ndarray, scale: float, out_ptr: int, threads: int = 4) -> None:
    """Gather uint8 rows, convert to float32 * scale, write to a raw (e.g. pinned) pointer."""
    assert src.dtype == np.uint8 and src.flags.c_contiguous
    idx = np.ascontiguousarray(idx, dtype=np.int64)
    row = int(np.prod(src.shape[1:], dtype=np.int64))
    _lib.check_rt(lib().tfdl_rt_gather_rows_u8_to_f32(_np_ptr(src), src.shape[0], row, _np_ptr(idx), idx.size,
                                                     C.c_float(scale), C.c_void_p(out_ptr), threads),
                  "gather_rows_u8_to_f32")


class Group:
    """Process group over TCP: rendezvous on the TF_CONFIG addresses, barrier, small-blob
    all-gather / broadcast, ring all-reduce, heartbeat-based failure detection."""

    def __init__(self, rank: int, addrs: Sequence[str], timeout_s: float = 60.0, hb_interval_s: float = 0.2,
                 hb_timeout_s: float = 5.0) -> None:
        self.rank, self.world, self.addrs = rank, len(addrs), list(addrs)
        arr = (C.c_char_p * len(addrs))(*[a.encode() for a in addrs])
        self._h = lib().tfdl_rt_group_create(rank, len(addrs), arr, int(timeout_s * 1000), int(hb_interval_s * 1000),
                                             int(hb_timeout_s * 1000))
        if not self._h:
            raise CollectiveError(lib().tfdl_rt_last_error().decode())

    def _ck(self, rc: int) -> None:
        if rc != 0:
            raise CollectiveError(lib().tfdl_rt_last_error().decode())

    def close(self) -> None:
        if getattr(self, "_h", None):
            lib().tfdl_rt_group_destroy(self._h)
            self._h = None

    def __del__(self) -> None:  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def set_abort_word(self, ptr: int) -> None:
        """Attach (ptr != 0) or detach (0) the pinned, device-mapped uint32 that the heartbeat thread sets to
        failed_rank + 1 when a peer dies; GPU kernels poll it inside their spin loops."""
        if self._h:
            lib().tfdl_rt_group_set_abort_word(self._h, C.c_void_p(ptr))

    @property
    def failed_rank(self) -> int:
        return int(lib().tfdl_rt_group_failed_rank(self._h))

    def set_timeout(self, seconds: float) -> None:
        lib().tfdl_rt_group_set_timeout(self._h, int(seconds * 1000))

    @property
    def bytes_sent(self) -> int:
        return int(lib().tfdl_rt_group_bytes_sent(self._h))

    def barrier(self) -> None:
        self._ck(lib().tfdl_rt_barrier(self._h))

    def allgather_bytes(self, blob: bytes) -> List[bytes]:
        n = len(blob)
        out = C.create_string_buffer(max(1, n * self.world))
        self._ck(lib().tfdl_rt_allgather(self._h, blob, n, out))
        return [out.raw[i * n:(i + 1) * n] for i in range(self.world)]

    def allgather_object(self, obj) -> list:
        import pickle

        blob = pickle.dumps(obj)
        sizes = self.allgather_bytes(len(blob).to_bytes(8, "little"))
        m = max(int.from_bytes(s, "little") for s in sizes)
        parts = self.allgather_bytes(blob.ljust(m, b"\0"))
        return [pickle.loads(p[:int.from_bytes(s, "little")]) for p, s in zip(parts, sizes)]

    def broadcast_array(self, a: np.ndarray) -> np.ndarray:
        """In-place broadcast of rank 0's contents."""
        assert a.flags.c_contiguous
        self._ck(lib().tfdl_rt_broadcast0(self._h, _np_ptr(a), a.nbytes))
        return a

    def allreduce_(self, a: np.ndarray) -> np.ndarray:
        """In-place sum over all ranks (float32 / float64)."""
        assert a.flags.c_contiguous
        if a.dtype == np.float32:
            self._ck(lib().tfdl_rt_ring_allreduce_f32(self._h, _np_ptr(a), a.size))
        elif a.dtype == np.float64:
            self._ck(lib().tfdl_rt_ring_allreduce_f64(self._h, _np_ptr(a), a.size))
        else:
            raise TypeError(f"ring all-reduce supports float32/float64, got {a.dtype}")
        return a

    def allreduce_ptr_f32(self, ptr: int, n: int) -> None:
        self._ck(lib().tfdl_rt_ring_allreduce_f32(self._h, C.c_void_p(ptr), n))


# TensorFlow DataType enum values used in bundles
TF_DTYPES: Dict[str, int] = {"float32": 1, "float64": 2, "int32": 3, "uint8": 4, "int16": 5, "int8": 6, "int64": 9,
                             "bool": 10, "float16": 19}
TF_DTYPES_INV = {v: k for k, v in TF_DTYPES.items()}
DT_STRING = 7


class BundleWriter:
    def __init__(self, prefix: str) -> None:
        os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
        self._h = lib().tfdl_rt_bundle_writer_create(prefix.encode())
        self.prefix = prefix

    def add(self, key: str, a: np.ndarray) -> None:
        a = np.asarray(a)
        a = a if a.ndim == 0 else np.ascontiguousarray(a)
        if str(a.dtype) not in TF_DTYPES:
            raise TypeError(f"dtype {a.dtype} is not supported in checkpoints")
        dims = (C.c_int64 * max(1, a.ndim))(*a.shape)
        _lib.check_rt(lib().tfdl_rt_bundle_writer_add(self._h, key.encode(), TF_DTYPES[str(a.dtype)], a.ndim, dims,
                                                     _np_ptr(a), a.nbytes), "bundle add")

    def add_string(self, key: str, data: bytes) -> None:
        _lib.check_rt(lib().tfdl_rt_bundle_writer_add_string(self._h, key.encode(), data, len(data)), "bundle add_string")

    def finish(self) -> None:
        h, self._h = self._h, None
        _lib.check_rt(lib().tfdl_rt_bundle_writer_finish(h), "bundle finish")

    def __del__(self) -> None:  # pragma: no cover
        try:
            if getattr(self, "_h", None):
                lib().tfdl_rt_bundle_writer_abort(self._h)
        except Exception:
            pass


class BundleReader:
    def __init__(self, prefix: str) -> None:
        self._h = lib().tfdl_rt_bundle_reader_open(prefix.encode())
        if not self._h:
            raise IOError(lib().tfdl_rt_last_error().decode())
        self.entries: Dict[str, Tuple[int, int, Tuple[int, ...], int]] = {}
        n = lib().tfdl_rt_bundle_reader_num_entries(self._h)
        key = C.create_string_buffer(1024)
        dt, nd, nb = C.c_int(), C.c_int(), C.c_uint64()
        dims = (C.c_int64 * 8)()
        for i in range(n):
            _lib.check_rt(lib().tfdl_rt_bundle_reader_entry(self._h, i, key, 1024, C.byref(dt), C.byref(nd), dims,
                                                           C.byref(nb)), "bundle entry")
            self.entries[key.value.decode()] = (i, dt.value, tuple(dims[d] for d in range(nd.value)), nb.value)

    def keys(self) -> List[str]:
        return list(self.entries)

    def shape(self, key: str) -> Tuple[int, ...]:
        return self.entries[key][2]

    def dtype(self, key: str) -> str:
        d = self.entries[key][1]
        return "string" if d == DT_STRING else TF_DTYPES_INV[d]

    def read(self, key: str):
        i, dt, shape, nbytes = self.entries[key]
        buf = C.create_string_buffer(max(1, nbytes))
        wr = C.c_uint64()
        _lib.check_rt(lib().tfdl_rt_bundle_reader_read(self._h, i, buf, nbytes, C.byref(wr)), f"bundle read {key}")
        if dt == DT_STRING:
            return buf.raw[:wr.value]
        return np.frombuffer(buf.raw[:wr.value], dtype=TF_DTYPES_INV[dt]).reshape(shape).copy()

    def close(self) -> None:
        if getattr(self, "_h", None):
            lib().tfdl_rt_bundle_reader_close(self._h)
            self._h = None

    def __del__(self) -> None:  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass


class EventWriter:
    """TensorBoard scalar event file."""

    def __init__(self, logdir: str, suffix: str = "") -> None:
        os.makedirs(logdir, exist_ok=True)
        self._h = lib().tfdl_rt_events_open(logdir.encode(), suffix.encode())
        if not self._h:
            raise IOError(lib().tfdl_rt_last_error().decode())
        self.path = lib().tfdl_rt_events_path(self._h).decode()

    def scalar(self, tag: str, value: float, step: int) -> None:
        lib().tfdl_rt_events_scalar(self._h, tag.encode(), float(value), int(step))

    def flush(self) -> None:
        lib().tfdl_rt_events_flush(self._h)

    def close(self) -> None:
        if getattr(self, "_h", None):
            lib().tfdl_rt_events_close(self._h)
            self._h = None

    def __del__(self) -> None:  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass
