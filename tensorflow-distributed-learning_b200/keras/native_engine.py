"""Native (sm_100a) training engine for the reference CNN topology.

`try_build(model)` recognises the layer stack of /root/reference/tf_dist_example.py:40-52
    Conv2D(C1,3,relu) - MaxPooling2D() - Conv2D(C2,3,relu) - MaxPooling2D() - Flatten - Dense(D1,relu) - Dense(NC)
compiled with SparseCategoricalCrossentropy(from_logits=True) + SGD (any lr / momentum / nesterov) or Adam, and returns an
engine whose training step is seven kernels of this library - input staging, conv1+conv2 forward, the fused dense block,
conv2 wgrad, conv2 dgrad + conv1 wgrad (csrc/kernels/cnn_tma.cu, dense_block.cu; tcgen05 for every GEMM-shaped op) and the
two buckets' fused NVLink all-reduce + update kernels (csrc/kernels/allreduce.cu) - launched up to 16 steps per CUDA graph
by the C++ executor (csrc/kernels/program.cu): Python hands it a ring of pinned batches (or row numbers into the dataset
arrays, gathered by one C++ call) and does not run again until the chunk is queued.

`NativeCnnEngine` drives ONE replica (one GPU of a one-process-per-GPU launch, or replica r of an in-process group);
`MirroredNativeEngine` drives all GPUs of a single-process MirroredStrategy() through one NativeCnnEngine per GPU.

Anything else (other topologies, evaluate/predict, a ragged last batch) runs on the TorchEngine, which shares the same
parameter arena and calls the same library's layer kernels (implicit-GEMM Conv2D, pooling, GEMM). On a machine with CUDA
devices the native library must be present: a missing .so is an error, never a silent fallback.
"""
from __future__ import annotations

import ctypes as C
import os
import time
from typing import Any, List, Optional, Tuple

import numpy as np
import torch

from ..utils.trace import nvtx_range

from ..ops import _lib
from ..runtime import native as _rt
from ..ops.allreduce import FusedState, P2PFusedAllReduce, SymmetricBuffer
from ..ops.structs import AR_NAMES, CnnDims, CnnStep, FusedArParams, make_dims
from ..parallel.values import PerReplica
from . import layers as L
from . import losses as _losses
from . import metrics as _metrics
from . import optimizers as _optimizers


def _match(model: Any) -> Optional[Tuple[int, int, int, int, int, int]]:
    """Returns (H0, W0, C1, C2, D1, NC) if the model is the supported topology."""
    from .engine import Sequential

    if not isinstance(model, Sequential) or not model.built:
        return None
    ls = [l for l in model._layers() if not isinstance(l, L.InputLayer)]
    if len(ls) != 7:
        return None
    c1, p1, c2, p2, fl, d1, d2 = ls
    ok = (isinstance(c1, L.Conv2D) and isinstance(p1, L.MaxPooling2D) and isinstance(c2, L.Conv2D)
          and isinstance(p2, L.MaxPooling2D) and isinstance(fl, L.Flatten) and isinstance(d1, L.Dense)
          and isinstance(d2, L.Dense))
    if not ok:
        return None
    for c in (c1, c2):
        if c.kernel_size != (3, 3) or c.strides != (1, 1) or c.padding != "valid" or c.activation != "relu" or not c.use_bias:
            return None
    for p in (p1, p2):
        if p.pool_size != (2, 2) or p.strides != (2, 2) or p.padding != "valid":
            return None
    if d1.activation != "relu" or d2.activation not in (None, "linear") or not d1.use_bias or not d2.use_bias:
        return None
    ish = model._input_shape
    if ish is None or len(ish) != 3 or ish[2] != 1:
        return None
    if not isinstance(model.loss, _losses.SparseCategoricalCrossentropy) or not model.loss.from_logits \
            or model.loss.ignore_class is not None:
        return None
    opt = model.optimizer
    if not (isinstance(opt, _optimizers.SGD) or (isinstance(opt, _optimizers.Adam) and not opt.weight_decay)):
        return None  # SGD / momentum / Nesterov and Adam are fused into the all-reduce kernels
    for m in model.compiled_metrics:
        if not isinstance(m, _metrics.SparseCategoricalAccuracy):
            return None
    H0, W0 = ish[0], ish[1]
    C1, C2, D1, NC = c1.filters, c2.filters, d1.units, d2.units
    if C1 % 32 or C1 > 64 or C2 % 32 or C2 > 128 or D1 % 16 or D1 > 256 or not (2 <= NC <= 64):
        return None
    if ((H0 - 2) // 2) * C1 > 1024 * 99 or ((W0 - 2) // 2) * C1 > 1024:
        return None
    if not all(w.trainable for w in model.weights):
        return None
    return H0, W0, C1, C2, D1, NC


def try_build(model: Any) -> Any:
    """Native engine for `model`, or None if its topology has no native step program."""
    _lib.kernels()  # raises NativeError when the library has not been built: no silent fallback on a GPU box
    s = model.strategy
    if any(d.type != "cuda" for d in s.local_devices):
        return None
    multi = s.num_local_replicas > 1
    if multi and type(s.cross_device_ops).__name__ != "MirroredPeerAllReduce":
        return None  # several local replicas without peer-mapped arenas: the portable engine
    shape = _match(model)
    if shape is None:
        return None
    return MirroredNativeEngine(model, shape) if multi else NativeCnnEngine(model, shape)


class NativeCnnEngine:
    CHUNK = int(os.environ.get("TFDL_FIT_CHUNK", "64"))

    def __init__(self, model: Any, shape: Tuple[int, int, int, int, int, int], replica: int = 0) -> None:
        self.model = model
        self.strategy = model.strategy
        self.replica = replica  # which local replica (device, arena) this engine drives
        self.device = self.strategy.local_devices[replica]
        self.shape = shape
        self.lib = _lib.kernels()
        self._prog: Optional[C.c_void_p] = None
        self._B: Optional[int] = None
        self._torch_engine: Any = None
        self._keep: List[Any] = []
        self._pinned: dict = {}            # host ranges page-locked for the zero-staging input path
        self.prof = {"gather": 0.0, "enqueue": 0.0, "mark": 0.0, "read_wait": 0.0, "ring_wait": 0.0}  # host seconds
        self._ptr_keep: List[Any] = [None, None]
        torch.cuda.set_device(self.device)
        # graphs cannot be captured on the legacy default stream: the engine owns its stream
        self.stream = torch.cuda.Stream(self.device)
        arena = model.arenas[replica]
        ops = self.strategy.cross_device_ops
        if type(ops).__name__ == "MirroredPeerAllReduce":
            ops = ops.replica_ops[replica]  # this replica's rank of the in-process group
        if not isinstance(ops, P2PFusedAllReduce):
            if self.strategy.num_workers != 1:
                raise RuntimeError("a multi-process GPU group must be created through MirroredStrategy / "
                                   "MultiWorkerMirroredStrategy so that its arenas are peer-mapped")
            ops = P2PFusedAllReduce(self.device, single=True)
            for t in (arena.params, arena.grads):
                b = SymmetricBuffer.__new__(SymmetricBuffer)
                b.numel, b.dtype, b.device, b.world, b.rank = t.numel(), t.dtype, t.device, 1, 0
                b.multicast_ptr, b.tensor, b.ptrs, b.handle = 0, t, [t.data_ptr()], None
                ops._arenas.append(b)
        self.ops = ops
        opt = model.optimizer
        self._adam = isinstance(opt, _optimizers.Adam)
        self.state: FusedState = ops.fused_state(arena.params, arena.grads, momentum=bool(getattr(opt, "momentum", 0.0)),
                                                 adam=self._adam)
        self._adam_steps0 = int(opt.iterations)  # Adam's t before this state's first kernel
        if self.strategy.num_local_replicas == 1:  # the optimizer's slots ARE the kernel's buffers
            if self._adam:
                old = dict(opt._slots) if opt._slots else {}
                for name, buf in (("m", self.state.mom), ("v", self.state.mom2)):
                    if old.get(name):
                        buf.copy_(old[name][0])
                opt._slots = {"m": [self.state.mom], "v": [self.state.mom2]}
            elif opt.momentum:
                opt._slots = {"momentum": [self.state.mom]}
        # Gradient buckets (reverse layer order, like the reference's backward): the dense layers' gradients
        # - 91% of the arena - are final right after the fused dense block, so their all-reduce + update runs
        # on its own stream while the conv backward kernels execute; the conv bucket follows at the end.
        self.buckets: List[FusedState] = [self.state]
        self.overlap = os.environ.get("TFDL_AR_OVERLAP", "1") == "1"
        if self.overlap:
            cut = self.model.weights[4].spec.offset  # first dense variable
            dense_blocks = int(os.environ.get("TFDL_DENSE_AR_BLOCKS", "64"))  # CTAs of the overlapped dense-bucket kernel
            self.buckets = [self.state.bucket(0, cut), self.state.bucket(cut, self.state.n - cut, max_blocks=dense_blocks)]
        self.hyperparams_changed()
        self.state.refresh_shadow()
        torch.cuda.synchronize(self.device)
        self.metrics_dev = torch.zeros(3, dtype=torch.float32, device=self.device)
        self._metrics_dirty = False  # steps have run since the counters were last read or closed
        lib = self.lib
        lib.tfdl_cnn_program_run.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
        lib.tfdl_cnn_program_steps_done.restype = C.c_int64
        lib.tfdl_cnn_program_steps_done.argtypes = [C.c_void_p]
        lib.tfdl_cnn_program_step.argtypes = [C.c_void_p, C.c_void_p]
        lib.tfdl_cnn_program_destroy.argtypes = [C.c_void_p]

    # ------------------------------------------------------------------------------------------
    @property
    def algo_name(self) -> str:
        return "+".join(b.algo_name for b in self.buckets)

    @property
    def launches_per_step(self) -> int:
        """Kernels of this library launched per training step (graph nodes), as counted by the executor."""
        self.lib.tfdl_cnn_program_launches_per_step.restype = C.c_int64
        self.lib.tfdl_cnn_program_launches_per_step.argtypes = [C.c_void_p]
        return int(self.lib.tfdl_cnn_program_launches_per_step(self._prog)) if self._prog else 0

    def hyperparams_changed(self) -> None:
        opt = self.model.optimizer
        self._lr = opt.learning_rate
        # the whole-arena state is what the unbucketed program (batch sizes the fused dense block does not take) steps
        for b in list(getattr(self, "buckets", [])) + [self.state]:
            if self._adam:
                b.set_hyper(self._lr, 0.0, False, 1.0, opt.beta_1, opt.beta_2, opt.epsilon, float(self._adam_steps0))
            else:
                b.set_hyper(self._lr, opt.momentum, opt.nesterov, 1.0)

    def weights_changed(self) -> None:
        self.state.refresh_shadow()

    def _offsets(self) -> List[int]:
        return [w.spec.offset for w in self.model.weights]

    def _build_program(self, B: int) -> None:
        if self._prog is not None:
            torch.cuda.synchronize(self.device)
            self.lib.tfdl_cnn_program_destroy(self._prog)
            self._prog, self._keep = None, []
        H0, W0, C1, C2, D1, NC = self.shape
        d = make_dims(B, H0, W0, C1, C2, D1, NC, nrep=self.strategy.num_replicas_in_sync)
        dev = self.device
        bf = torch.bfloat16
        z = lambda *s, dt=bf: torch.zeros(s, dtype=dt, device=dev)  # noqa: E731
        bufs = dict(a1=z(B, d.P1h, d.P1w, C1), a2=z(B, d.F), dh=z(B, D1), dy2=z(B, d.P1h, d.P1w, C2),
                    g1=z(B, d.P1h, d.P1w, C1), idx1=z(B, d.P1h, d.P1w, C1, dt=torch.uint8),
                    idx2=z(B, d.F, dt=torch.uint8))
        accs = [z(B, D1, dt=torch.float32) for _ in range(2)]  # split-K accumulator, one per step parity
        sync = z(4, dt=torch.int32)
        # one device record [x | labels] per parity, so a step's inputs arrive with ONE host-to-device copy
        nx = B * H0 * W0 * 4
        self.in_dev = [z(nx + B * 4, dt=torch.uint8) for _ in range(2)]
        self.x_dev = [t[:nx].view(torch.float32).view(B, H0, W0) for t in self.in_dev]
        self.y_dev = [t[nx:].view(torch.int32) for t in self.in_dev]
        self.step_out = [z(2, dt=torch.float32) for _ in range(2)]
        offs = self._offsets()
        steps = (CnnStep * 2)()
        for par in range(2):
            s = steps[par]
            s.dims = d
            s.x, s.labels = self.x_dev[par].data_ptr(), self.y_dev[par].data_ptr()
            s.params = self.state.params.tensor.data_ptr()
            s.shadow = self.state.shadow.tensor.data_ptr()
            s.grads = self.state.grads[par].tensor.data_ptr()
            (s.off_w1c, s.off_b1c, s.off_w2c, s.off_b2c, s.off_w1d, s.off_b1d, s.off_w2d, s.off_b2d) = offs
            for k, t in bufs.items():
                setattr(s, k, t.data_ptr())
            fused = bool(self.lib.tfdl_cnn_dense_block_supported(C.byref(d))) and \
                os.environ.get("TFDL_DENSE_IMPL", "fused") == "fused"
            s.use_dense_block = int(fused)
            # the fused dense block clears the OTHER parity's accumulator; the unfused head re-zeroes its own
            s.acc = accs[par].data_ptr() if fused else accs[0].data_ptr()
            s.acc_next = accs[par ^ 1].data_ptr()
            s.sync = sync.data_ptr()
            s.metrics = self.metrics_dev.data_ptr()
            s.step_out = self.step_out[par].data_ptr()
            s.dense1_ksplit = max(1, d.F // 64)
            s.use_tma_conv = int(bool(self.lib.tfdl_cnn_tma_supported(C.byref(d))) and
                                 os.environ.get("TFDL_CONV_IMPL", "tma") == "tma")
            s.use_fused_conv = int(bool(s.use_tma_conv) and os.environ.get("TFDL_CONV_FUSE", "1") == "1")
        last = self.buckets[0]
        ars = (FusedArParams * 2)(last.struct(0), last.struct(1))
        prog = C.c_void_p()
        use_graph = int(os.environ.get("TFDL_USE_GRAPH", "1"))
        # which states the C++ executor steps (sync_optimizer_slots gathers momentum slices of the sliced ones only)
        for st_ in [self.state] + list(self.buckets):
            st_.last_algo = None
        if len(self.buckets) == 2 and steps[0].use_dense_block:
            for b_ in self.buckets:
                b_.last_algo = b_.algo
            early = self.buckets[1]
            ars_early = (FusedArParams * 2)(early.struct(0), early.struct(1))
            _lib.check(self.lib.tfdl_cnn_program_create2(steps, ars, last.algo, ars_early, early.algo, use_graph,
                                                         C.byref(prog)), "cnn_program_create2")
        else:
            full = self.state
            full.last_algo = full.algo
            ars = (FusedArParams * 2)(full.struct(0), full.struct(1))
            ars_early = None
            _lib.check(self.lib.tfdl_cnn_program_create(steps, ars, full.algo, use_graph, C.byref(prog)),
                       "cnn_program_create")
        self._prog, self._B, self._dims = prog, B, d
        self._keep = [bufs, steps, ars, ars_early, accs, sync]
        self.rec_bytes = B * H0 * W0 * 4 + B * 4
        n_rec = max(2, self.CHUNK)
        # two pinned rings: Python packs chunk k+1 while the executor is still consuming chunk k
        self.rings = [torch.empty((n_rec, self.rec_bytes), dtype=torch.uint8).pin_memory() for _ in range(2)]
        self.host_outs = [torch.zeros((n_rec, 2), dtype=torch.float32).pin_memory() for _ in range(2)]
        self.ring_done = [torch.cuda.Event(), torch.cuda.Event()]
        self._ring_i = 0
        self._prefetch = os.environ.get("TFDL_FIT_PREFETCH", "1") == "1"
        self.ring, self.host_out = self.rings[0], self.host_outs[0]
        # the executor alternates parity with its own step counter, which restarts at 0: both gradient
        # buffers must be clean and the device epoch is only used for flag values
        for g in self.state.grads:
            g.tensor.zero_()
        torch.cuda.synchronize(dev)
        # every graph variant (1/2/4/8 steps x parity) is captured now, not inside the first fit()
        self.lib.tfdl_cnn_program_prepare.argtypes = [C.c_void_p, C.c_void_p]
        _lib.check(self.lib.tfdl_cnn_program_prepare(self._prog, self.stream.cuda_stream), "cnn_program_prepare")
        torch.cuda.synchronize(dev)

    # ------------------------------------------------------------------------------------------
    def pack_record(self, slot: int, x: np.ndarray, y: np.ndarray) -> None:
        rec = self.ring[slot].numpy()
        B, (H0, W0) = self._B, self.shape[:2]
        nx = B * H0 * W0 * 4
        rec[:nx].view(np.float32)[:] = np.asarray(x, dtype=np.float32).reshape(-1)
        rec[nx:].view(np.int32)[:] = np.asarray(y).reshape(-1).astype(np.int32, copy=False)

    # ---- zero-staging input path ------------------------------------------------------------------------
    # Batches that are contiguous views of one long-lived dataset array (from_tensor_slices(...).batch(), no
    # shuffle) are DMA'd straight from that array: it is page-locked once with cudaHostRegister and the executor
    # gets one pointer per step. Anything else (shuffled gathers, mapped batches, odd dtypes) is staged in the ring.
    ZC_MAX_BYTES = int(os.environ.get("TFDL_PIN_MAX_MB", "4096")) << 20

    def _owner(self, a: np.ndarray) -> Optional[np.ndarray]:
        b = a
        while isinstance(b.base, np.ndarray):
            b = b.base
        return b if b.base is None or not isinstance(b.base, np.ndarray) else None

    def _pinned_range(self, a: np.ndarray) -> bool:
        """True when `a` lies inside a host range this engine has page-locked (registering its owner on first use)."""
        if not a.flags.c_contiguous:
            return False
        own = self._owner(a)
        if own is None or own.nbytes < 8 * a.nbytes or not own.flags.c_contiguous:
            return False  # a per-batch temporary: registering it would cost more than copying it
        key = own.ctypes.data
        ent = self._pinned.get(key)
        if ent is None:
            if own.nbytes > self.ZC_MAX_BYTES or os.environ.get("TFDL_ZERO_COPY", "1") != "1":
                return False
            self.lib.tfdl_host_register.argtypes = [C.c_void_p, C.c_uint64]
            if self.lib.tfdl_host_register(C.c_void_p(key), C.c_uint64(own.nbytes)) != 0:
                self._pinned[key] = (None, 0)  # not registrable (e.g. read-only mapping): remember, use the ring
                return False
            ent = self._pinned[key] = (own, own.nbytes)  # the reference keeps the memory alive while registered
        return ent[0] is not None and key <= a.ctypes.data and a.ctypes.data + a.nbytes <= key + ent[1]

    def _zero_copy_ok(self, xs: List[np.ndarray], ys: List[np.ndarray]) -> bool:
        if not xs or xs[0].dtype != np.float32 or ys[0].dtype != np.int32:
            return False
        return all(self._pinned_range(a) for a in xs) and all(self._pinned_range(a) for a in ys)

    def _run_ptrs(self, xs: List[np.ndarray], ys: List[np.ndarray], r: int) -> None:
        k = len(xs)
        px = (C.c_void_p * k)(*[a.ctypes.data for a in xs])
        py = (C.c_void_p * k)(*[a.ctypes.data for a in ys])
        if self.model.optimizer.learning_rate != self._lr:
            self.hyperparams_changed()
        self.stream.wait_stream(torch.cuda.current_stream(self.device))
        self.lib.tfdl_cnn_program_run_ptrs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                                       C.c_void_p]
        with nvtx_range(f"tfdl/native_chunk[{k}]"):
            _lib.check(self.lib.tfdl_cnn_program_run_ptrs(self._prog, px, py, k, self.host_out.data_ptr(),
                                                          self.stream.cuda_stream), "cnn_program_run_ptrs")
        self._metrics_dirty = True
        self._ptr_keep[r] = (px, py, xs, ys)  # the executor returns before the copies ran: keep the views alive
        if self.replica == 0:
            self.model.optimizer.iterations += k

    def _pack(self, xs: List[np.ndarray], ys: List[np.ndarray]) -> None:
        """Stage a chunk of batches in the pinned ring: one multi-threaded C++ call when the arrays are plain
        float32 / int labels (the normal case), per-record numpy copies otherwise."""
        fast = all(a.dtype == np.float32 and a.flags.c_contiguous for a in xs) and \
            all(a.dtype == ys[0].dtype and a.flags.c_contiguous for a in ys) and ys[0].dtype in (np.int32, np.int64)
        if fast and xs:
            _rt.pack_records(xs, ys, self.ring.data_ptr(), self.rec_bytes)
        else:
            for k, (xb, yb) in enumerate(zip(xs, ys)):
                self.pack_record(k, xb, yb)

    def run_pinned(self, records: torch.Tensor, n_steps: int, host_out: Optional[torch.Tensor] = None,
                   first: int = 0) -> None:
        """Drive n_steps steps from a pinned [n_records, rec_bytes] uint8 tensor (asynchronous), starting at record
        `first`; their (loss, accuracy) pairs go to host_out[first:first + n_steps]."""
        assert records.is_pinned() and records.shape[1] == self.rec_bytes and 0 <= first < records.shape[0]
        if self.model.optimizer.learning_rate != self._lr:
            self.hyperparams_changed()
        out_ptr = host_out.data_ptr() + 8 * first if host_out is not None else None
        self.stream.wait_stream(torch.cuda.current_stream(self.device))
        with nvtx_range(f"tfdl/native_chunk[{n_steps}]"):
            _lib.check(self.lib.tfdl_cnn_program_run(self._prog, records.data_ptr() + first * self.rec_bytes,
                                                     records.shape[0] - first, n_steps, out_ptr,
                                                     self.stream.cuda_stream), "cnn_program_run")
        self._metrics_dirty = True
        if self.replica == 0:
            self.model.optimizer.iterations += n_steps

    def run_device(self, records: torch.Tensor, n_steps: int) -> None:
        """Drive n_steps steps from records already on the device ([n_records, rec_bytes] uint8)."""
        assert records.is_cuda and records.shape[1] == self.rec_bytes
        self.lib.tfdl_cnn_program_run_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]
        self.stream.wait_stream(torch.cuda.current_stream(self.device))
        _lib.check(self.lib.tfdl_cnn_program_run_device(self._prog, records.data_ptr(), records.shape[0], n_steps,
                                                        self.stream.cuda_stream), "cnn_program_run_device")
        self._metrics_dirty = True
        if self.replica == 0:
            self.model.optimizer.iterations += n_steps

    def prepare(self, batch_size: int) -> None:
        """Build (or rebuild) the step program for a per-replica batch size."""
        if self._B != batch_size:
            self._build_program(batch_size)

    def _split(self, batch: Any) -> Tuple[np.ndarray, np.ndarray]:
        x, y = batch[0], batch[1]
        return np.asarray(x), np.asarray(y)

    def train_chunk(self, batches: List[PerReplica]) -> None:
        """Train on consecutive batches without returning to Python between steps."""
        i = 0
        while i < len(batches):
            x, y = self._split(batches[i][0])
            B = x.shape[0]
            if B == 0:
                i += 1
                continue
            if B % 16 != 0:
                self._fallback_step(batches[i])
                i += 1
                continue
            if self._B != B:
                self._build_program(B)
            r = self._ring_i
            self._ring_i ^= 1
            self.ring, self.host_out = self.rings[r], self.host_outs[r]
            self.ring_done[r].synchronize()  # the run that last read this ring has finished (no-op if never used)
            k = 0
            xs, ys = [], []
            while i + k < len(batches) and k < self.ring.shape[0]:
                xb, yb = self._split(batches[i + k][0])
                if xb.shape[0] != B:
                    break
                xs.append(xb)
                ys.append(yb)
                k += 1
            if self._zero_copy_ok(xs, ys):
                self._run_ptrs(xs, ys, r)
            else:
                self._pack(xs, ys)
                self.run_pinned(self.ring, k, self.host_out)
            self.ring_done[r].record(self.stream)
            self._last_k = k
            i += k

    def train_step(self, batches: PerReplica) -> None:
        self.train_chunk([batches])

    # ---- index path: the executor's ring is filled straight from the dataset arrays ---------------------------------
    def index_arrays(self, base: Any) -> Optional[Tuple[np.ndarray, np.ndarray]]:
        """(images, labels) when `base` is a plain array store this engine can gather from itself: float32 images of the
        model's input size and integer labels (what `.map(scale).cache()` of tf_dist_example.py:31 leaves)."""
        arrays = getattr(base, "arrays", None)
        if not isinstance(arrays, (tuple, list)) or len(arrays) != 2:
            return None
        x, y = arrays
        if not (isinstance(x, np.ndarray) and isinstance(y, np.ndarray)):
            return None
        H0, W0 = self.shape[:2]
        if x.dtype != np.float32 or not x.flags.c_contiguous or int(np.prod(x.shape[1:])) != H0 * W0:
            return None
        if y.ndim != 1 or y.dtype not in (np.uint8, np.int32, np.int64) or not y.flags.c_contiguous:
            return None
        return x, y

    def train_index_chunk(self, x: np.ndarray, y: np.ndarray, idxs: List[np.ndarray]) -> None:
        """Train on len(idxs) consecutive batches given as row numbers into (x, y): ONE multi-threaded gather builds the
        records in the pinned ring (shuffle + batch + staging in a single pass), the executor runs them.

        When the device has nothing queued (the start of a fit) the gather runs as a background job on the runtime's
        worker pool - the prefetch of this input path: the first two steps are launched as soon as their records are in
        the ring, and every further launch takes whatever has been gathered by then, so the copy of the later batches
        and the launch work of the earlier ones overlap each other and the running steps."""
        i = 0
        while i < len(idxs):
            B = int(idxs[i].size)
            if B == 0:
                i += 1
                continue
            if B % 16 != 0:  # ragged last batch: the portable step
                self._fallback_step(PerReplica([(x[idxs[i]], y[idxs[i]])]))
                i += 1
                continue
            if self._B != B:
                self._build_program(B)
            r = self._ring_i
            self._ring_i ^= 1
            self.ring, self.host_out = self.rings[r], self.host_outs[r]
            tw = time.perf_counter()
            self.ring_done[r].synchronize()
            self.prof["ring_wait"] += time.perf_counter() - tw
            k, cap = 0, self.ring.shape[0]
            while i + k < len(idxs) and k < cap and idxs[i + k].size == B:
                k += 1
            t0 = time.perf_counter()
            group = idxs[i:i + k]
            if all(int(g[-1]) - int(g[0]) == B - 1 and int(g[B // 2]) - int(g[0]) == B // 2 for g in group) and \
                    y.dtype == np.int32:
                # consecutive rows (an unshuffled dataset): no gather at all - the executor reads the dataset array in place
                xs = [x[int(g[0]):int(g[0]) + B] for g in group]
                ys = [y[int(g[0]):int(g[0]) + B] for g in group]
                if self._zero_copy_ok(xs, ys):
                    self._run_ptrs(xs, ys, r)
                    self.ring_done[r].record(self.stream)
                    self._last_k = k
                    i += k
                    continue
            flat = np.ascontiguousarray(np.concatenate(group), dtype=np.int64)
            if k > 2 and self._prefetch and self.stream.query():
                self._run_prefetched(x, y, flat, k, B, t0)
            else:  # the device is busy anyway: gather everything behind the running steps, one launch sequence
                _rt.gather_records(x, y, flat, k, B, self.ring.data_ptr(), self.rec_bytes, threads=8)
                t1 = time.perf_counter()
                self.run_pinned(self.ring, k, self.host_out)
                self.prof["gather"] += t1 - t0
                self.prof["enqueue"] += time.perf_counter() - t1
            self.ring_done[r].record(self.stream)
            self._last_k = k
            i += k

    def _run_prefetched(self, x: np.ndarray, y: np.ndarray, flat: np.ndarray, k: int, B: int, t0: float) -> None:
        """k steps whose records are gathered by a background job while the first ones are already being launched."""
        job = _rt.GatherJob(x, y, flat, k, B, self.ring.data_ptr(), self.rec_bytes)
        try:
            done = 0
            while done < k:
                # the first launch is two steps (the device starts ~50 us after the gather began); later ones wait for
                # at least four more records so that a slow gather does not turn into a launch per step
                job.wait(min(k, done + (2 if done == 0 else 4)))
                n = job.ready() - done
                if done + n < k:
                    if done == 0:
                        n = 4 if n >= 4 else 2
                    elif k - (done + n) <= 2:  # nearly everything: take the rest in the same launch sequence
                        job.wait(k)
                        n = k - done
                    else:
                        n = 1 << (n.bit_length() - 1)  # whole power-of-two graphs; the tail stays in one piece
                t1 = time.perf_counter()
                self.run_pinned(self.ring, n, self.host_out, first=done)
                self.prof["gather"] += t1 - t0
                t0 = time.perf_counter()
                self.prof["enqueue"] += t0 - t1
                done += n
        finally:
            job.end()

    # ---- epoch totals without stopping the device ----------------------------------------------------------------
    def epoch_mark(self, reduce: bool = True) -> Tuple[int, Any]:
        """Close an epoch on the device: stream-ordered after the epoch's last step, the (loss sum, correct, examples)
        counters are summed over replicas (NVLink kernel), copied to a pinned slot and cleared. Returns a handle for
        epoch_read(); the host does not wait here, so the next epoch's batches can be staged while this one runs."""
        if not hasattr(self, "_ep_pinned"):
            self._ep_pinned = [torch.zeros(3, dtype=torch.float64).pin_memory() for _ in range(4)]
            self._ep_i = 0
        slot = self._ep_i
        self._ep_i = (self._ep_i + 1) % len(self._ep_pinned)
        tm = time.perf_counter()
        with torch.cuda.stream(self.stream):
            if self.ops.world > 1 and not reduce:  # in-process group: the caller adds the replicas' totals on the host
                lib = self.lib
                lib.tfdl_metrics_snapshot.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
                _lib.check(lib.tfdl_metrics_snapshot(_lib.ptr(self.metrics_dev), 3, _lib.ptr(self._ep_pinned[slot]), 1,
                                                     _lib.cur_stream()), "metrics_snapshot")
            else:
                self.ops.snapshot_sum_to_host(self.metrics_dev, self._ep_pinned[slot])
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self._metrics_dirty = False
        self.prof["mark"] += time.perf_counter() - tm
        return slot, ev

    def epoch_read(self, handle: Tuple[int, Any]) -> Tuple[float, float, float]:
        slot, ev = handle
        tw = time.perf_counter()
        ev.synchronize()
        self.ops.check_abort()
        self.prof["read_wait"] += time.perf_counter() - tw
        t = self._ep_pinned[slot]
        return float(t[0]), float(t[1]), float(t[2])

    def last_step_out(self) -> Tuple[float, float]:
        self.stream.synchronize()
        self.ops.check_abort()
        k = max(1, getattr(self, "_last_k", 1))
        return float(self.host_out[k - 1, 0]), float(self.host_out[k - 1, 1])

    # ------------------------------------------------------------------------------------------
    def flush_metrics(self) -> None:
        """Fold the device-side (loss sum, correct, examples) counters into the Keras metric objects."""
        if not self._metrics_dirty:
            return
        self._metrics_dirty = False
        self.stream.synchronize()
        self.ops.check_abort()
        m = self.metrics_dev.cpu()
        if float(m[2]) == 0.0:
            return
        self.metrics_dev.zero_()
        self.model.loss_tracker.add(self.replica, float(m[0]), float(m[2]))
        for met in self.model.compiled_metrics:
            met.add(self.replica, float(m[1]), float(m[2]))

    def _torch(self) -> Any:
        if self._torch_engine is None:
            from .engine import TorchEngine

            self._torch_engine = TorchEngine(self.model)
        return self._torch_engine

    def _fallback_step(self, batches: PerReplica) -> None:
        """Ragged batch: one portable step on the shared arena, then resynchronise the native state."""
        self.flush_metrics()
        self.stream.synchronize()
        self._torch().train_step(batches)
        torch.cuda.current_stream(self.device).synchronize()
        for g in self.state.grads:
            g.tensor.zero_()
        self.state.refresh_shadow()

    def test_step(self, batches: PerReplica) -> None:
        self._torch().test_step(batches)

    def predict_step(self, x: np.ndarray) -> np.ndarray:
        return self._torch().predict_step(x)

    def __del__(self) -> None:  # pragma: no cover
        try:
            if self._prog is not None:
                torch.cuda.synchronize(self.device)
                self.lib.tfdl_cnn_program_destroy(self._prog)
            for key, (own, _) in list(self._pinned.items()):
                if own is not None:
                    self.lib.tfdl_host_unregister(C.c_void_p(key))
        except Exception:
            pass


class MirroredNativeEngine:
    """MirroredStrategy() in ONE process over several GPUs (tf_dist_example.py:13; README.md:15-19): one native step
    program per GPU, all driven from this thread. Each program is the one the one-process-per-GPU launch runs - its
    all-reduce kernels push gradient packets into the other GPUs' inboxes over NVLink (peer access instead of
    symmetric memory) - so the replicas stay bit-identical. A step's per-worker batch is split in replica order."""

    CHUNK = NativeCnnEngine.CHUNK

    def __init__(self, model: Any, shape: Tuple[int, int, int, int, int, int]) -> None:
        self.model, self.strategy, self.shape = model, model.strategy, shape
        self.devices = list(self.strategy.local_devices)
        self.replicas: List[NativeCnnEngine] = []
        for r, d in enumerate(self.devices):
            with torch.cuda.device(d):
                self.replicas.append(NativeCnnEngine(model, shape, replica=r))
        if isinstance(model.optimizer, _optimizers.Adam):
            model.optimizer._slots = {"m": [rep.state.mom for rep in self.replicas],
                                      "v": [rep.state.mom2 for rep in self.replicas]}
        elif model.optimizer.momentum:
            model.optimizer._slots = {"momentum": [rep.state.mom for rep in self.replicas]}
        self._torch_engine: Any = None
        self.prof = self.replicas[0].prof

    # ---- facts ----------------------------------------------------------------------------------------------
    @property
    def algo_name(self) -> str:
        return self.replicas[0].algo_name

    @property
    def launches_per_step(self) -> int:
        return sum(rep.launches_per_step for rep in self.replicas)

    @property
    def stream(self) -> Any:
        return self.replicas[0].stream

    def _each(self, fn) -> list:
        out = []
        for rep in self.replicas:
            with torch.cuda.device(rep.device):
                out.append(fn(rep))
        return out

    def hyperparams_changed(self) -> None:
        self._each(lambda rep: rep.hyperparams_changed())

    def weights_changed(self) -> None:
        self._each(lambda rep: rep.weights_changed())

    def prepare(self, batch_size: int) -> None:
        self._each(lambda rep: rep.prepare(batch_size))

    # ---- training -------------------------------------------------------------------------------------------
    def train_chunk(self, batches: List[PerReplica]) -> None:
        """batches[i][r] is replica r's batch of step i. Every replica's executor gets the whole chunk before the next
        replica's is queued: the launches are asynchronous, and a replica whose kernels wait for a peer's packets is
        served as soon as that peer's launches follow."""
        sizes = {int(np.asarray(b[r][0]).shape[0]) for b in batches for r in range(len(self.replicas))}
        if len(sizes) != 1 or next(iter(sizes)) % 16 != 0:
            for b in batches:  # ragged or uneven step: the portable engine, then resynchronise the native state
                self._fallback_step(b)
            return
        for r, rep in enumerate(self.replicas):
            with torch.cuda.device(rep.device):
                rep.train_chunk([PerReplica([b[r]]) for b in batches])

    def train_step(self, batches: PerReplica) -> None:
        self.train_chunk([batches])

    def index_arrays(self, base: Any) -> Any:
        return self.replicas[0].index_arrays(base)

    def train_index_chunk(self, x: np.ndarray, y: np.ndarray, idxs: List[np.ndarray]) -> None:
        W = len(self.replicas)
        if any(i.size % W or (i.size // W) % 16 for i in idxs):
            for i in idxs:  # ragged last step
                parts = np.array_split(i, W)
                self._fallback_step(PerReplica([(x[p], y[p]) for p in parts]))
            return
        for r, rep in enumerate(self.replicas):
            with torch.cuda.device(rep.device):
                rep.train_index_chunk(x, y, [i[r * (i.size // W):(r + 1) * (i.size // W)] for i in idxs])

    def run_device(self, records: List[torch.Tensor], n_steps: int) -> None:
        """Benchmark path: records[r] is replica r's device-resident pool."""
        for rep, rec in zip(self.replicas, records):
            with torch.cuda.device(rep.device):
                rep.run_device(rec, n_steps)

    @property
    def rec_bytes(self) -> int:
        return self.replicas[0].rec_bytes

    # ---- metrics --------------------------------------------------------------------------------------------
    def epoch_mark(self) -> Any:
        return self._each(lambda rep: rep.epoch_mark(reduce=False))

    def epoch_read(self, handles: Any) -> Tuple[float, float, float]:
        tot = [0.0, 0.0, 0.0]
        for rep, h in zip(self.replicas, handles):
            with torch.cuda.device(rep.device):
                vals = rep.epoch_read(h)
            tot = [a + b for a, b in zip(tot, vals)]
        return tot[0], tot[1], tot[2]

    def flush_metrics(self) -> None:
        self._each(lambda rep: rep.flush_metrics())

    def last_step_out(self) -> Tuple[float, float]:
        outs = self._each(lambda rep: rep.last_step_out())
        return float(np.mean([o[0] for o in outs])), float(np.mean([o[1] for o in outs]))

    # ---- everything that has no native program --------------------------------------------------------------
    def _torch(self) -> Any:
        if self._torch_engine is None:
            from .engine import TorchEngine

            self._torch_engine = TorchEngine(self.model)
        return self._torch_engine

    def _fallback_step(self, batches: PerReplica) -> None:
        self.flush_metrics()
        self._each(lambda rep: rep.stream.synchronize())
        self._torch().train_step(batches)

        def resync(rep: NativeCnnEngine) -> None:
            torch.cuda.current_stream(rep.device).synchronize()
            for g in rep.state.grads:
                g.tensor.zero_()
            rep.state.refresh_shadow()

        self._each(resync)
        self._each(lambda rep: torch.cuda.synchronize(rep.device))

    def test_step(self, batches: PerReplica) -> None:
        self._torch().test_step(batches)

    def predict_step(self, x: np.ndarray) -> np.ndarray:
        return self._torch().predict_step(x)
