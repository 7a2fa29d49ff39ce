"""ctypes mirrors of the structs in include/tfdl_kernels.h (keep field order identical)."""
from __future__ import annotations

import ctypes as C

MAX_RANKS = 8
AR_ONESHOT, AR_TWOSHOT, AR_MULTIMEM, AR_PUSH, AR_PUSH2 = 0, 1, 2, 3, 4
AR_NAMES = {AR_ONESHOT: "oneshot", AR_TWOSHOT: "twoshot", AR_MULTIMEM: "multimem", AR_PUSH: "push", AR_PUSH2: "push2"}


class Hyper(C.Structure):
    _fields_ = [("lr", C.c_float), ("momentum", C.c_float), ("nesterov", C.c_float), ("scale", C.c_float),
                ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float), ("t_offset", C.c_float)]


class FusedArParams(C.Structure):
    _fields_ = [
        ("world", C.c_int), ("rank", C.c_int),
        ("n", C.c_int64),
        ("params", C.c_void_p * MAX_RANKS),
        ("shadow", C.c_void_p * MAX_RANKS),
        ("grads", (C.c_void_p * MAX_RANKS) * 2),
        ("mom", C.c_void_p),
        ("flags", C.c_void_p * MAX_RANKS),
        ("pad_stride", C.c_uint32),
        ("max_blocks", C.c_int),
        ("mc_params", C.c_void_p),
        ("mc_shadow", C.c_void_p),
        ("mc_grads", C.c_void_p * 2),
        ("epoch", C.c_void_p),
        ("done", C.c_void_p),
        ("hyper", C.c_void_p),
        ("abort_word", C.c_void_p),
        ("parity", C.c_int),
        ("inbox", C.c_void_p * MAX_RANKS),
        ("inbox_bytes", C.c_int64),
        ("mom2", C.c_void_p),
    ]


class CnnDims(C.Structure):
    _fields_ = [(n, C.c_int) for n in
                ("B", "H0", "W0", "C1", "P1h", "P1w", "C2", "P2h", "P2w", "F", "D1", "NC", "nrep")]


class CnnStep(C.Structure):
    _fields_ = [
        ("dims", CnnDims),
        ("x", C.c_void_p), ("labels", C.c_void_p),
        ("params", C.c_void_p), ("shadow", C.c_void_p), ("grads", C.c_void_p),
        ("off_w1c", C.c_int64), ("off_b1c", C.c_int64), ("off_w2c", C.c_int64), ("off_b2c", C.c_int64),
        ("off_w1d", C.c_int64), ("off_b1d", C.c_int64), ("off_w2d", C.c_int64), ("off_b2d", C.c_int64),
        ("a1", C.c_void_p), ("a2", C.c_void_p), ("dh", C.c_void_p), ("dy2", C.c_void_p), ("g1", C.c_void_p),
        ("idx1", C.c_void_p), ("idx2", C.c_void_p),
        ("acc", C.c_void_p), ("metrics", C.c_void_p), ("step_out", C.c_void_p),
        ("dense1_ksplit", C.c_int),
        ("use_tma_conv", C.c_int),
        ("use_dense_block", C.c_int),
        ("acc_next", C.c_void_p),
        ("sync", C.c_void_p),
        ("use_fused_conv", C.c_int),
    ]


def make_dims(B: int, H0: int = 28, W0: int = 28, C1: int = 32, C2: int = 64, D1: int = 128, NC: int = 10,
              nrep: int = 1) -> CnnDims:
    P1h, P1w = (H0 - 2) // 2, (W0 - 2) // 2
    P2h, P2w = (P1h - 2) // 2, (P1w - 2) // 2
    return CnnDims(B, H0, W0, C1, P1h, P1w, C2, P2h, P2w, P2h * P2w * C2, D1, NC, nrep)
