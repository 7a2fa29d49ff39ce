"""Fused all-reduce + SGD over NVLink peer memory: Python side (buffers, handles, dispatch).

Stands in for `NcclAllReduce` / `CollectiveCommunication.NCCL|AUTO` of the reference
(/root/reference/README.md:17,21,23; tf_dist_example.py:12). torch.distributed's symmetric-memory
allocator is used ONLY to allocate peer-mapped buffers and exchange their handles (including the
NVLS multicast mapping); every byte of gradient traffic is moved by the hand-written kernels in
csrc/kernels/allreduce.cu.

Buffers per rank (all symmetric when world > 1):
    params   fp32 master weights            grads[0], grads[1]   fp32 gradient arenas (parity double buffer)
    shadow   bf16 copy of params            flags                3 barrier pads of MAX_BLOCKS*world u32
    inbox    packet inboxes of the push paths (2 parities x world regions)
Device-local: epoch / done counters, hyper-parameters {lr, momentum, nesterov, scale, beta1, beta2, eps, t_offset},
momentum / Adam slot arenas. Host (pinned, device-mapped): the abort word the heartbeat thread sets when a peer dies.

Algorithm choice (`AUTO`, `pick_algo`): small buckets are PUSHED - every rank writes {value, epoch} packets into
its peers' inboxes and polls its own, one NVLink one-way trip and no barrier; mid-size buckets take the same
packets in two hops (slices to their owner, new weights back: push2); large arenas go through the NVSwitch
(multimem: in-switch reduction, one multicast store of the fp32 weights, bf16 shadow cast locally) from 4 GPUs up or,
on 2 GPUs and without a multicast mapping, two-shot peer loads/stores. The pull-style one-shot kernel (one barrier, then peer loads)
remains for ragged element counts and for comparison. Override with
TFDL_AR_ALGO=oneshot|twoshot|multimem|push|push2; thresholds TFDL_AR_PUSH_MAX / TFDL_AR_PUSH2_MAX (bytes).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Any, Dict, List, Optional, Sequence

import numpy as np
import torch

from ..parallel.cross_device_ops import CrossDeviceOps
from . import _lib
from .structs import (AR_MULTIMEM, AR_NAMES, AR_ONESHOT, AR_PUSH, AR_PUSH2, AR_TWOSHOT, MAX_RANKS, FusedArParams,
                      Hyper)

MAX_BLOCKS = 148
ONESHOT_MAX_BYTES = int(os.environ.get("TFDL_AR_ONESHOT_MAX", 4 * 1024 * 1024))  # pull one-shot vs sliced crossover
PUSH_MAX_BYTES = int(os.environ.get("TFDL_AR_PUSH_MAX", 2 * 1024 * 1024))        # bytes x (W-1) sent per rank, one hop
PUSH2_MAX_BYTES = int(os.environ.get("TFDL_AR_PUSH2_MAX", 8 * 1024 * 1024))      # arena bytes, two hops
# NVLS pays off from 4 GPUs up: with two ranks the in-switch reduction saves nothing and the two-shot kernel's plain peer
# loads / stores are faster (256 MB, 2 GPUs: fused 483 vs 318 GB/s, plain sum 635 vs 385; 4 GPUs: on par; 8: multimem ahead)
MULTIMEM_MIN_WORLD = int(os.environ.get("TFDL_AR_MULTIMEM_MIN_WORLD", "4"))
_ALGOS = {"oneshot": AR_ONESHOT, "twoshot": AR_TWOSHOT, "multimem": AR_MULTIMEM, "push": AR_PUSH, "push2": AR_PUSH2}


def multimem_ok(world: int, has_multicast: bool) -> bool:
    """NVLS multicast reduction: needs the multicast mapping, pays off from MULTIMEM_MIN_WORLD GPUs, and is left out under
    op determinism (the order in which the switch adds the ranks' values is not specified; every other kernel here adds
    them in rank order)."""
    from .. import config as _config

    return bool(has_multicast and world >= MULTIMEM_MIN_WORLD and not _config.is_op_determinism_enabled())


def pick_algo(nbytes: int, world: int, has_multicast: bool) -> int:
    """The reference's AUTO: choose by message size, world size and hardware (README.md:21)."""
    forced = os.environ.get("TFDL_AR_ALGO", "").lower()
    if forced:
        a = _ALGOS[forced]
        if a == AR_MULTIMEM and not has_multicast:
            raise RuntimeError("TFDL_AR_ALGO=multimem but the symmetric allocation has no multicast mapping")
        return a if world > 1 else AR_ONESHOT
    if world == 1:
        return AR_ONESHOT
    if nbytes % 16:
        return AR_ONESHOT if nbytes * (world - 1) <= ONESHOT_MAX_BYTES else AR_TWOSHOT  # ragged tail: scalar path
    if nbytes * (world - 1) <= PUSH_MAX_BYTES:
        return AR_PUSH
    if nbytes <= PUSH2_MAX_BYTES:
        return AR_PUSH2
    return AR_MULTIMEM if multimem_ok(world, has_multicast) else AR_TWOSHOT


class SymmetricBuffer:
    """One tensor per rank, peer-mapped. world == 1: a plain device tensor."""

    def __init__(self, numel: int, dtype: torch.dtype, device: torch.device, world: int, rank: int) -> None:
        self.numel, self.dtype, self.device, self.world, self.rank = numel, dtype, device, world, rank
        self.multicast_ptr = 0
        if world == 1:
            self.tensor = torch.zeros(numel, dtype=dtype, device=device)
            self.ptrs = [self.tensor.data_ptr()]
            self.handle = None
            return
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem

        group = dist.group.WORLD
        try:
            symm_mem.enable_symm_mem_for_group(group.group_name)
        except Exception:
            pass
        self.tensor = symm_mem.empty(numel, dtype=dtype, device=device)
        self.tensor.zero_()
        self.handle = symm_mem.rendezvous(self.tensor, group=group)
        self.ptrs = [int(p) for p in self.handle.buffer_ptrs]
        try:
            self.multicast_ptr = int(self.handle.multicast_ptr or 0)
        except Exception:
            self.multicast_ptr = 0

    def barrier(self) -> None:
        if self.handle is not None:
            torch.cuda.synchronize(self.device)
            self.handle.barrier()


class P2PFusedAllReduce(CrossDeviceOps):
    """Cross-device ops for one-process-per-GPU groups (and the degenerate single-GPU case)."""

    name = "p2p-fused"

    def __init__(self, device: torch.device, control_group: Any = None, single: bool = False) -> None:
        self.device = device
        self.control = control_group
        self.dist = None
        if not single:  # single=True: a private one-GPU instance even if a process group exists
            try:
                import torch.distributed as dist

                self.dist = dist if dist.is_initialized() else None
            except Exception:  # pragma: no cover
                self.dist = None
        self.world = self.dist.get_world_size() if self.dist else 1
        self.rank = self.dist.get_rank() if self.dist else 0
        if self.world > MAX_RANKS:
            raise RuntimeError(f"the fused NVLink all-reduce supports up to {MAX_RANKS} GPUs per group, got {self.world}")
        _lib.kernels()  # fail loudly on a GPU box without the native library
        # Abort word: one uint32 in pinned, device-mapped host memory. Every spin loop of the collective kernels polls
        # it; the control group's heartbeat thread stores failed_rank + 1 there when a peer dies (csrc/runtime/
        # control.cc), so survivors leave their kernels within the heartbeat timeout and check_abort() raises.
        self._abort = torch.zeros(1, dtype=torch.int32).pin_memory() if device.type == "cuda" else None
        if self._abort is not None and self.control is not None and hasattr(self.control, "set_abort_word"):
            self.control.set_abort_word(self._abort.data_ptr())
        self._arenas: List[SymmetricBuffer] = []
        self._states: Dict[int, "FusedState"] = {}
        self._plain: Dict[Any, "FusedState"] = {}
        self._fused_states: List["FusedState"] = []  # every updating state / bucket (sync_optimizer_slots)
        self._scalars: Optional[list] = None

    def make_buffer(self, numel: int, dtype: torch.dtype) -> SymmetricBuffer:
        """One peer-mapped buffer per rank (collective: every rank makes the same sequence of calls)."""
        return SymmetricBuffer(numel, dtype, self.device, self.world, self.rank)

    def group_barrier(self) -> None:
        """Host-side rendezvous of the group's ranks (before a collective that reads what peers wrote eagerly)."""
        torch.cuda.synchronize(self.device)
        if self.dist is not None:
            self.dist.barrier()

    # ---- arena allocation hook used by parallel.values.ParamArena -----------------------------
    def alloc_arena(self, numel: int, device: torch.device) -> torch.Tensor:
        buf = self.make_buffer(numel, torch.float32)
        self._arenas.append(buf)
        return buf.tensor

    def _find(self, t: torch.Tensor) -> SymmetricBuffer:
        for b in self._arenas:
            if b.tensor.data_ptr() == t.data_ptr():
                return b
        raise RuntimeError("tensor was not allocated through P2PFusedAllReduce.alloc_arena")

    # ---- fused state for a (params, grads) pair ---------------------------------------------
    def fused_state(self, params: torch.Tensor, grads: torch.Tensor, momentum: bool, adam: bool = False) -> "FusedState":
        key = params.data_ptr()
        if key not in self._states:
            self._states[key] = FusedState(self, self._find(params), self._find(grads), momentum or adam, adam=adam)
        st = self._states[key]
        if (momentum or adam) and st.mom is None:
            st.mom = torch.zeros(st.n, dtype=torch.float32, device=self.device)
        if adam and st.mom2 is None:
            st.mom2 = torch.zeros(st.n, dtype=torch.float32, device=self.device)
        if (momentum or adam) and st._structs[0].mom is None or (adam and st._structs[0].mom2 is None):
            st._structs = [st._make_struct(par) for par in (0, 1)]
        return st

    # ---- CrossDeviceOps interface (used by the torch engine on GPU and by tests) ----------------
    def allreduce_grads(self, grads: Sequence[torch.Tensor]) -> None:
        assert len(grads) == 1, "one replica per process on the P2P path"
        if self.world == 1:
            return
        g = grads[0]
        key = g.data_ptr()
        if key not in self._plain:
            self._plain[key] = FusedState(self, None, self._find(g), False, plain=True)
        self._plain[key].allreduce_sum()

    def allreduce_scalars(self, vals: np.ndarray) -> np.ndarray:
        """Metric totals (sync-on-read SUM, README.md:67): a one-CTA NVLink kernel adds the ranks' float64 slots in rank
        order, so every rank reads bit-identical totals. No NCCL call."""
        if self.world == 1:
            return vals
        self.check_abort()
        v = np.ascontiguousarray(vals, dtype=np.float64).reshape(-1)
        if v.size > 256:  # not a metric read any more: generic path
            t = torch.from_numpy(v.copy()).to(self.device)
            self.dist.all_reduce(t)
            return t.cpu().numpy().reshape(np.shape(vals))
        if self._scalars is None:
            buf = self.make_buffer(512, torch.float64)
            st = FusedState(self, None, buf, False, plain=True)
            self._scalars = [buf, st, torch.zeros(256, dtype=torch.float64, device=self.device), 0]
        buf, st, out, calls = self._scalars
        par = calls & 1  # == the kernel's epoch parity: this state is used by nothing else
        self._scalars[3] = calls + 1
        buf.tensor[par * 256:par * 256 + v.size].copy_(torch.from_numpy(v))
        _lib.check(_lib.kernels().tfdl_allreduce_small_f64(C.byref(st.struct(0)), _lib.ptr(out), int(v.size),
                                                           _lib.cur_stream()), "allreduce_small_f64")
        res = out[:v.size].cpu().numpy().reshape(np.shape(vals))
        self.check_abort()  # a kernel released by the abort word returns garbage: never hand that out as a total
        return res

    def allreduce_scalars_device(self, vals: torch.Tensor) -> torch.Tensor:
        """Sum of <= 256 float64 values over the ranks with input and result on the device, ordered on the current
        stream - no host round trip (epoch totals are read one epoch late, keras/engine.py). Same kernel and slot
        scheme as allreduce_scalars; the two must not be interleaved on different streams."""
        if self.world == 1:
            return vals
        n = int(vals.numel())
        assert n <= 256 and vals.dtype == torch.float64 and vals.is_cuda
        if self._scalars is None:
            buf = self.make_buffer(512, torch.float64)
            st = FusedState(self, None, buf, False, plain=True)
            self._scalars = [buf, st, torch.zeros(256, dtype=torch.float64, device=self.device), 0]
        buf, st, out, calls = self._scalars
        par = calls & 1
        self._scalars[3] = calls + 1
        buf.tensor[par * 256:par * 256 + n].copy_(vals.reshape(-1))
        _lib.check(_lib.kernels().tfdl_allreduce_small_f64(C.byref(st.struct(0)), _lib.ptr(out), n, _lib.cur_stream()),
                   "allreduce_small_f64")
        return out[:n].clone()

    def snapshot_sum_to_host(self, metrics: torch.Tensor, host_slot: torch.Tensor) -> None:
        """Stream-ordered, no host wait: the fp32 device counters `metrics` are converted to float64, cleared, summed over
        the ranks (NVLink kernel, rank order: identical totals everywhere) and written into `host_slot`, a pinned float64
        tensor the device writes through its mapped address. Two launches (one on a single GPU)."""
        n = int(metrics.numel())
        lib = _lib.kernels()
        lib.tfdl_metrics_snapshot.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
        if self.world == 1:
            _lib.check(lib.tfdl_metrics_snapshot(_lib.ptr(metrics), n, _lib.ptr(host_slot), 1, _lib.cur_stream()),
                       "metrics_snapshot")
            return
        if self._scalars is None:
            buf = self.make_buffer(512, torch.float64)
            st = FusedState(self, None, buf, False, plain=True)
            self._scalars = [buf, st, torch.zeros(256, dtype=torch.float64, device=self.device), 0]
        buf, st, _out, calls = self._scalars
        par = calls & 1
        self._scalars[3] = calls + 1
        _lib.check(lib.tfdl_metrics_snapshot(_lib.ptr(metrics), n, C.c_void_p(buf.tensor.data_ptr() + par * 256 * 8), 1,
                                             _lib.cur_stream()), "metrics_snapshot")
        _lib.check(lib.tfdl_allreduce_small_f64(C.byref(st.struct(0)), _lib.ptr(host_slot), n, _lib.cur_stream()),
                   "allreduce_small_f64")

    def sync_optimizer_slots(self) -> None:
        """Make every rank's momentum buffer complete (collective: every rank calls it, e.g. before a checkpoint).
        The sliced paths (two-shot, multimem, push2) update the momentum of a slice only on the rank that owns the
        slice, so between steps rank r holds slice r of each such bucket; here the slices are gathered - each rank
        contributes its slice in an otherwise zero buffer and the plain NVLink sum reassembles the whole."""
        if self.world == 1:
            return
        for st in list(self._fused_states):
            if st.mom is None or st.last_algo not in (AR_TWOSHOT, AR_MULTIMEM, AR_PUSH2):
                continue
            n = st.n
            per = (((n + 3) // 4 + self.world - 1) // self.world) * 4
            lo, hi = min(n, per * self.rank), min(n, per * (self.rank + 1))
            if st._gather_tmp is None:
                buf = self.make_buffer(n, torch.float32)
                st._gather_tmp = (buf, FusedState(self, None, buf, False, plain=True))
            buf, plain = st._gather_tmp
            for slot in (st.mom, st.mom2):
                if slot is None:
                    continue
                view = slot[st.offset:st.offset + n]
                buf.tensor.zero_()
                buf.tensor[lo:hi].copy_(view[lo:hi])
                self.group_barrier()
                plain.allreduce_sum()
                view.copy_(buf.tensor)
                torch.cuda.synchronize(self.device)
        torch.cuda.synchronize(self.device)

    def broadcast_params(self, params: Sequence[torch.Tensor]) -> None:
        """Initial values from the chief (tf_dist_example.py:56-57): every other rank copies rank 0's arena over NVLink
        with peer loads (csrc/kernels/allreduce.cu:broadcast_kernel)."""
        if self.world == 1:
            return
        for t in params:
            try:
                sym = self._find(t)
            except RuntimeError:  # not a symmetric arena: library fallback
                self.dist.broadcast(t, src=0)
                continue
            key = ("bcast", t.data_ptr())
            if key not in self._plain:
                self._plain[key] = FusedState(self, None, sym, False, plain=True)
            _lib.check(_lib.kernels().tfdl_broadcast_f32(C.byref(self._plain[key].struct(0)), 0, _lib.cur_stream()),
                       "broadcast_f32")
        torch.cuda.synchronize(self.device)

    def barrier(self) -> None:
        if self.world > 1:
            torch.cuda.synchronize(self.device)
            self.dist.barrier()
        if self.control is not None:
            self.control.barrier()

    def check_abort(self) -> None:
        """Raise if a peer has been declared dead (cheap: one read of pinned host memory). Called by the engines
        whenever they are about to wait for, or have just waited for, the device."""
        if self._abort is not None and int(self._abort[0]) != 0:
            from ..runtime.native import CollectiveError

            raise CollectiveError(f"collective aborted: worker rank {int(self._abort[0]) - 1} failed (heartbeat lost); "
                                  "restart all workers and resume from the last checkpoint")

    def close(self) -> None:
        if self.control is not None:
            if hasattr(self.control, "set_abort_word"):
                self.control.set_abort_word(0)
            self.control.close()


class FusedState:
    """Everything the fused kernel needs for one arena: extra buffers, counters, the C struct."""

    def __init__(self, ops: P2PFusedAllReduce, params: Optional[SymmetricBuffer], grads0: SymmetricBuffer,
                 momentum: bool, plain: bool = False, adam: bool = False) -> None:
        self.ops = ops
        dev, W, R = ops.device, ops.world, ops.rank
        n = grads0.numel
        assert plain or n % 4 == 0, "parameter arenas are padded to 16 bytes (parallel/values.py)"
        self.n = n
        self.params, self.grads = params, [grads0]
        self.shadow: Optional[SymmetricBuffer] = None
        if not plain:
            self.grads.append(ops.make_buffer(n, torch.float32))
            self.shadow = ops.make_buffer(n, torch.bfloat16)
        self.mom = torch.zeros(n, dtype=torch.float32, device=dev) if momentum else None
        self.mom2 = torch.zeros(n, dtype=torch.float32, device=dev) if adam else None  # Adam's second-moment slot
        self.pad_stride = MAX_BLOCKS * max(W, 1)
        self.flags = ops.make_buffer(3 * self.pad_stride, torch.int32) if W > 1 else None
        self.counters = torch.zeros(4, dtype=torch.int32, device=dev)  # epoch, done, abort, pad
        self.hyper = torch.zeros(8, dtype=torch.float32, device=dev)
        self.has_multicast = bool(W > 1 and grads0.multicast_ptr and (plain or (
            params is not None and params.multicast_ptr and self.grads[1].multicast_ptr)))
        self.algo = pick_algo(n * 4, W, self.has_multicast) if not plain else (
            AR_MULTIMEM if multimem_ok(W, self.has_multicast) else AR_TWOSHOT)
        self.offset = 0
        self.max_blocks = MAX_BLOCKS
        self.inbox: Optional[SymmetricBuffer] = None
        self.last_algo: Optional[int] = None
        self._gather_tmp: Any = None
        if not plain:
            ops._fused_states.append(self)
        self._structs = [self._make_struct(par) for par in (0, 1)]
        if W > 1:
            grads0.barrier()  # every rank has zeroed its flags before anyone signals
        if not plain:
            self.ensure_inbox(self.algo)

    def bucket(self, offset: int, length: int, max_blocks: int = MAX_BLOCKS) -> "FusedState":
        """A view of elements [offset, offset+length) of the same buffers with its OWN flag pad, step counter
        and hyper-parameter words, so that it can be reduced by a separate kernel on a separate stream
        (bucketed overlap with backward). offset/length must be multiples of 4 elements (16 bytes)."""
        assert offset % 4 == 0 and length % 4 == 0 and offset + length <= self.n
        b = FusedState.__new__(FusedState)
        b.ops, b.n, b.offset, b.max_blocks = self.ops, length, offset, max_blocks
        b.params, b.grads, b.shadow, b.mom, b.mom2 = self.params, self.grads, self.shadow, self.mom, self.mom2
        dev, W, R = self.ops.device, self.ops.world, self.ops.rank
        b.pad_stride = self.pad_stride
        b.flags = self.ops.make_buffer(3 * b.pad_stride, torch.int32) if W > 1 else None
        b.counters = torch.zeros(4, dtype=torch.int32, device=dev)
        b.hyper = torch.zeros(8, dtype=torch.float32, device=dev)
        b.has_multicast = self.has_multicast
        b.algo = pick_algo(length * 4, W, b.has_multicast)
        b.inbox = None
        b.last_algo, b._gather_tmp = None, None
        self.ops._fused_states.append(b)
        b._structs = [b._make_struct(par) for par in (0, 1)]
        if W > 1:
            b.flags.barrier()
        b.ensure_inbox(b.algo)
        return b

    def ensure_inbox(self, algo: int) -> None:
        """Packet inbox for the push paths (collective: every rank calls it in the same order). Zeroed once - epoch 0
        is never sent - and never cleared again: stale packets carry an older epoch."""
        W = self.ops.world
        if W == 1 or algo not in (AR_PUSH, AR_PUSH2):
            return
        lib = _lib.kernels()
        lib.tfdl_fused_ar_inbox_bytes.restype = C.c_int64
        lib.tfdl_fused_ar_inbox_bytes.argtypes = [C.c_int, C.c_int, C.c_int64]
        need = int(lib.tfdl_fused_ar_inbox_bytes(algo, W, self.n))
        if self.inbox is not None and self.inbox.numel * 4 >= need:
            return
        torch.cuda.synchronize(self.ops.device)
        self.inbox = self.ops.make_buffer((need + 3) // 4, torch.int32)
        self.inbox.barrier()  # nobody's first packet can land in memory that is zeroed afterwards
        self._structs = [self._make_struct(par) for par in (0, 1)]

    def _make_struct(self, parity: int) -> FusedArParams:
        p = FusedArParams()
        W, R = self.ops.world, self.ops.rank
        p.world, p.rank, p.n = W, R, self.n
        o4, o2 = self.offset * 4, self.offset * 2  # byte offsets of this bucket in fp32 / bf16 buffers
        for r in range(W):
            if self.params is not None:
                p.params[r] = self.params.ptrs[r] + o4
            if self.shadow is not None:
                p.shadow[r] = self.shadow.ptrs[r] + o2
            for k, g in enumerate(self.grads):
                p.grads[k][r] = g.ptrs[r] + o4
            if self.flags is not None:
                p.flags[r] = self.flags.ptrs[r]
        p.mom = self.mom.data_ptr() + o4 if self.mom is not None else None
        p.mom2 = self.mom2.data_ptr() + o4 if self.mom2 is not None else None
        p.pad_stride = self.pad_stride
        p.max_blocks = self.max_blocks
        if self.has_multicast:
            p.mc_params = self.params.multicast_ptr + o4 if self.params is not None else None
            p.mc_shadow = (self.shadow.multicast_ptr + o2 if self.shadow.multicast_ptr else None) \
                if self.shadow is not None else None
            for k, g in enumerate(self.grads):
                p.mc_grads[k] = g.multicast_ptr + o4
        if self.inbox is not None:
            for r in range(W):
                p.inbox[r] = self.inbox.ptrs[r]
            p.inbox_bytes = self.inbox.numel * 4
        base = self.counters.data_ptr()
        p.epoch, p.done = base, base + 4
        p.abort_word = self.ops._abort.data_ptr() if self.ops._abort is not None else base + 8
        p.hyper = self.hyper.data_ptr()
        p.parity = parity
        return p

    def struct(self, parity: int) -> FusedArParams:
        return self._structs[parity]

    def set_hyper(self, lr: float, momentum: float = 0.0, nesterov: bool = False, scale: float = 1.0, beta1: float = 0.0,
                  beta2: float = 0.0, eps: float = 0.0, t_offset: float = 0.0) -> None:
        """SGD / momentum / Nesterov, or - with beta2 > 0 - Adam (both slot arenas must exist). t_offset: optimizer steps
        taken before this state's first kernel (the kernel's own step counter supplies the rest of Adam's t)."""
        assert beta2 == 0.0 or (self.mom is not None and self.mom2 is not None), "Adam needs the m and v slot arenas"
        self.hyper.copy_(torch.tensor([lr, momentum, 1.0 if nesterov else 0.0, scale, beta1, beta2, eps, t_offset],
                                      dtype=torch.float32))

    def refresh_shadow(self) -> None:
        """fp32 master -> bf16 shadow (after set_weights / load_weights / broadcast)."""
        assert self.params is not None and self.shadow is not None
        _lib.check(_lib.kernels().tfdl_cast_f32_to_bf16(_lib.ptr(self.params.tensor), _lib.ptr(self.shadow.tensor),
                                                        C.c_int64(self.n), _lib.cur_stream()), "cast_f32_to_bf16")

    def step(self, parity: int, algo: Optional[int] = None) -> None:
        """One fused all-reduce + update on the current stream (eager; the training path captures it)."""
        if algo is not None and algo != self.algo:
            self.ensure_inbox(algo)
        self.last_algo = self.algo if algo is None else algo
        _lib.check(_lib.kernels().tfdl_fused_allreduce_sgd(C.byref(self._structs[parity]),
                                                           self.algo if algo is None else algo, _lib.cur_stream()),
                   "fused_allreduce_sgd")

    def allreduce_sum(self, algo: Optional[int] = None) -> None:
        a = self.algo if algo is None else algo
        if a not in (AR_MULTIMEM, AR_TWOSHOT):
            a = AR_MULTIMEM if multimem_ok(self.ops.world, self.has_multicast) else AR_TWOSHOT
        _lib.check(_lib.kernels().tfdl_allreduce_sum(C.byref(self._structs[0]), a, _lib.cur_stream()), "allreduce_sum")

    @property
    def algo_name(self) -> str:
        return AR_NAMES[self.algo]


# ---- single process, several GPUs (MirroredStrategy() in a plain python process; tf_dist_example.py:13, README.md:15-19) ----
class PeerBuffer:
    """One rank's view of a buffer that exists once per GPU of an in-process group: the same fields a SymmetricBuffer
    has, with plain cudaMalloc'ed tensors and peer access between the devices instead of symmetric memory (one
    process sees every device's pointers directly; there is no multicast mapping)."""

    def __init__(self, tensors: List[torch.Tensor], rank: int) -> None:
        self.tensor = tensors[rank]
        self.numel, self.dtype, self.device = self.tensor.numel(), self.tensor.dtype, self.tensor.device
        self.world, self.rank = len(tensors), rank
        self.ptrs = [t.data_ptr() for t in tensors]
        self.multicast_ptr = 0
        self.handle = None
        self._devices = [t.device for t in tensors]

    def barrier(self) -> None:
        for d in self._devices:
            torch.cuda.synchronize(d)


class _PeerPool:
    """Allocations shared by the per-device ops objects of one in-process group: request number k of every rank names
    the same logical buffer (every rank makes the same sequence of requests)."""

    def __init__(self, devices: Sequence[torch.device]) -> None:
        self.devices = list(devices)
        self.bufs: Dict[int, List[torch.Tensor]] = {}
        lib = _lib.kernels()
        for a in self.devices:
            for b in self.devices:
                if a != b:
                    _lib.check(lib.tfdl_enable_peer_access(a.index, b.index), "enable_peer_access")

    def get(self, rank: int, seq: int, numel: int, dtype: torch.dtype) -> PeerBuffer:
        if seq not in self.bufs:
            self.bufs[seq] = [torch.zeros(numel, dtype=dtype, device=d) for d in self.devices]
            for d in self.devices:
                torch.cuda.synchronize(d)
        ts = self.bufs[seq]
        assert ts[0].numel() == numel and ts[0].dtype == dtype, "ranks of an in-process group diverged in their allocations"
        return PeerBuffer(ts, rank)


class InProcessPeerOps(P2PFusedAllReduce):
    """Rank `rank` of a group whose ranks are the GPUs of ONE process. Same fused kernels and the same FusedState as the
    one-process-per-GPU path; only allocation (peer access instead of symmetric memory) and the host rendezvous differ."""

    name = "p2p-fused(in-process)"

    def __init__(self, pool: _PeerPool, rank: int) -> None:
        self.device = pool.devices[rank]
        self.control, self.dist = None, None
        self.world, self.rank = len(pool.devices), rank
        self.pool, self._seq = pool, 0
        self._abort = torch.zeros(1, dtype=torch.int32).pin_memory()
        self._arenas, self._states, self._plain, self._fused_states, self._scalars = [], {}, {}, [], None

    def make_buffer(self, numel: int, dtype: torch.dtype) -> PeerBuffer:  # type: ignore[override]
        seq = self._seq
        self._seq += 1
        with torch.cuda.device(self.device):
            return self.pool.get(self.rank, seq, numel, dtype)

    def group_barrier(self) -> None:
        for d in self.pool.devices:
            torch.cuda.synchronize(d)

    def barrier(self) -> None:
        self.group_barrier()

    def close(self) -> None:
        pass


class MirroredPeerAllReduce(CrossDeviceOps):
    """Cross-device ops of a single-process MirroredStrategy over several GPUs: every parameter / gradient arena is
    allocated through the matching per-device InProcessPeerOps so that the native step programs of the replicas can push
    gradients into each other's memory. Generic engines (torch ops) reduce with a local sum."""

    name = "p2p-fused(in-process)"

    def __init__(self, devices: Sequence[torch.device]) -> None:
        self.devices = list(devices)
        self.pool = _PeerPool(self.devices)
        self.replica_ops = [InProcessPeerOps(self.pool, r) for r in range(len(self.devices))]

    @staticmethod
    def possible(devices: Sequence[torch.device]) -> bool:
        if len(devices) < 2 or len(devices) > MAX_RANKS or any(d.type != "cuda" for d in devices):
            return False
        if len({d.index for d in devices}) != len(devices) or not _lib.kernels_available():
            return False
        return all(torch.cuda.can_device_access_peer(a.index, b.index) for a in devices for b in devices if a != b)

    def alloc_arena(self, numel: int, device: torch.device) -> torch.Tensor:
        r = [d.index for d in self.devices].index(device.index)
        return self.replica_ops[r].alloc_arena(numel, device)

    def allreduce_grads(self, grads: Sequence[torch.Tensor]) -> None:
        if len(grads) == 1:
            return
        acc = grads[0]
        for g in grads[1:]:
            acc.add_(g.to(acc.device, non_blocking=True))
        for g in grads[1:]:
            g.copy_(acc, non_blocking=True)

    def barrier(self) -> None:
        for d in self.devices:
            torch.cuda.synchronize(d)

    def sync_optimizer_slots(self) -> None:
        """Momentum of a sliced bucket lives on the slice's owner: copy every owner's slice to the other replicas."""
        self.barrier()
        W = len(self.devices)
        per_rank = [ops._fused_states for ops in self.replica_ops]
        for i, st0 in enumerate(per_rank[0]):
            if st0.mom is None or st0.last_algo not in (AR_TWOSHOT, AR_MULTIMEM, AR_PUSH2):
                continue
            n = st0.n
            per = (((n + 3) // 4 + W - 1) // W) * 4
            for r in range(W):
                lo, hi = min(n, per * r), min(n, per * (r + 1))
                src = per_rank[r][i]
                for q in range(W):
                    if q != r and hi > lo:
                        dst = per_rank[q][i]
                        dst.mom[dst.offset + lo:dst.offset + hi].copy_(src.mom[src.offset + lo:src.offset + hi])
                        if src.mom2 is not None:
                            dst.mom2[dst.offset + lo:dst.offset + hi].copy_(src.mom2[src.offset + lo:src.offset + hi])
        self.barrier()
