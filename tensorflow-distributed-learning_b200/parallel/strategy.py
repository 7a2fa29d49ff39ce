"""Distribution strategies: MirroredStrategy and MultiWorkerMirroredStrategy.

Reference surface (/root/reference/tf_dist_example.py:12-13,36,56; README.md:13-29):

    strategy = MultiWorkerMirroredStrategy(CollectiveCommunication.AUTO)   # reads TF_CONFIG
    strategy = MirroredStrategy()                                          # all local GPUs
    with strategy.scope(): model = build_and_compile()
    dist_dataset = strategy.experimental_distribute_dataset(dataset)

Semantics kept: synchronous data parallelism; one full model copy per replica; variables created
inside scope() are mirrored and start from the chief's initial values; the dataset is batched by the
GLOBAL batch size and each replica consumes global/num_replicas_in_sync; gradients are summed over
all replicas every step before anyone applies them (README.md:67).

Execution model (B200-first, differs from TF): the fused all-reduce + update kernels exchange gradients between the
replicas' arenas over NVLink in both launch forms. A MirroredStrategy constructed inside a torchrun-style launch
(WORLD_SIZE/RANK/LOCAL_RANK in the environment) is one replica of a one-process-per-GPU group (symmetric memory).
Constructed in a plain process it drives all its `devices` from one Python thread: several GPUs form an in-process
peer group (peer access) running one native step program each; CPU devices (devices=["cpu:0","cpu:1"], the CPU tests)
and GPUs without peer access reduce with a local sum.
"""
from __future__ import annotations

import contextlib
import os
import threading
from typing import Any, Callable, Iterator, List, Optional, Sequence, Tuple, Union

import numpy as np
import torch

from ..cluster import TFConfigClusterResolver
from ..data import dataset as _ds
from .cross_device_ops import (CollectiveCommunication, CrossDeviceOps, LocalAllReduce, TcpRingAllReduce,
                               TorchDistAllReduce)
from .values import PerReplica

_tls = threading.local()


def _stack() -> List["Strategy"]:
    if not hasattr(_tls, "stack"):
        _tls.stack = []
    return _tls.stack


def get_strategy() -> "Strategy":
    """Strategy of the innermost active scope(), or the default single-replica strategy."""
    s = _stack()
    return s[-1] if s else _default_strategy()


def has_strategy() -> bool:
    return bool(_stack())


_default: Optional["Strategy"] = None


def _default_strategy() -> "Strategy":
    global _default
    if _default is None:
        _default = OneDeviceStrategy()
    return _default


class ReduceOp:
    SUM = "SUM"
    MEAN = "MEAN"


class DistributedDataset:
    """What experimental_distribute_dataset returns: iterating yields, per step, a PerReplica of
    (per-replica) batches for this worker's local replicas."""

    def __init__(self, dataset: _ds.Dataset, num_local: int, per_replica_batch: int, global_batch: int) -> None:
        self.dataset = dataset
        self.num_local = num_local
        self.per_replica_batch = per_replica_batch
        self.global_batch = global_batch

    def __iter__(self) -> Iterator[PerReplica]:
        it = iter(self.dataset)
        while True:
            group = []
            for _ in range(self.num_local):
                try:
                    group.append(next(it))
                except StopIteration:
                    break
            if not group:
                return
            if len(group) < self.num_local:
                # ragged tail: replicas without data get an empty batch (weight 0 in the global mean)
                proto = group[0]
                empty = _ds._map_structure(lambda a: a[:0], proto)
                group += [empty] * (self.num_local - len(group))
            yield PerReplica(group)

    def cardinality(self) -> int:
        c = self.dataset.cardinality()
        return c if c < 0 else (c + self.num_local - 1) // self.num_local

    def index_stream(self):
        """(element store, iterator of per-step row numbers) for a single local replica over an indexable pipeline,
        else None. Lets an engine gather its batches itself (data/dataset.py:index_stream)."""
        st = _ds.index_stream(self.dataset)
        if st is None or self.num_local == 1:
            return st
        base, it = st

        def grouped():  # one step = one per-replica batch for each local replica, in replica order
            buf = []
            for idx in it:
                buf.append(idx)
                if len(buf) == self.num_local:
                    yield np.concatenate(buf)
                    buf = []
            if buf:
                yield np.concatenate(buf)

        return base, grouped()


class Strategy:
    """Common machinery; see MirroredStrategy / MultiWorkerMirroredStrategy."""

    def __init__(self, devices: Sequence[torch.device], resolver: TFConfigClusterResolver,
                 cross_device_ops: CrossDeviceOps, num_workers: int, worker_index: int) -> None:
        self.local_devices = list(devices)
        self.cluster_resolver = resolver
        self.cross_device_ops = cross_device_ops
        self.num_workers = num_workers
        self.worker_index = worker_index
        self._closed = False

    # ---- facts ----------------------------------------------------------------------------
    @property
    def num_local_replicas(self) -> int:
        return len(self.local_devices)

    @property
    def num_replicas_in_sync(self) -> int:
        return self.num_local_replicas * self.num_workers

    @property
    def is_chief(self) -> bool:
        """The chief alone writes checkpoints / TensorBoard (README.md:51)."""
        return self.worker_index == 0 if self.cluster_resolver.is_single_worker else self.cluster_resolver.is_chief

    @property
    def extended(self) -> "Strategy":
        return self

    @property
    def should_checkpoint(self) -> bool:
        return self.is_chief

    # ---- scope ----------------------------------------------------------------------------
    @contextlib.contextmanager
    def scope(self):
        """Everything that creates variables (model build AND compile, tf_dist_example.py:56-57)
        runs under this; layers / optimizers / metrics pick the strategy up from the thread-local
        stack and create mirrored / sync-on-read variables."""
        _stack().append(self)
        try:
            yield self
        finally:
            _stack().pop()

    # ---- data -----------------------------------------------------------------------------
    def experimental_distribute_dataset(self, dataset: _ds.Dataset, options: Any = None) -> DistributedDataset:
        """Shard across workers per the dataset's AutoShardPolicy, then re-batch the GLOBAL batch
        into per-replica batches (tf_dist_example.py:18,33-37)."""
        if isinstance(dataset, DistributedDataset):
            return dataset
        batch = _ds.find_batch_node(dataset)
        if batch is None:
            raise ValueError("experimental_distribute_dataset expects a dataset batched by the global batch size")
        sharded = _ds.auto_shard(dataset, self.num_workers, self.worker_index)
        rebatched, per = _ds.rebatch_for_replicas(sharded, self.num_replicas_in_sync)
        return DistributedDataset(rebatched, self.num_local_replicas, per, batch.batch_size)

    def distribute_datasets_from_function(self, fn: Callable[[Any], _ds.Dataset]) -> DistributedDataset:
        class _Ctx:
            num_input_pipelines = self.num_workers
            input_pipeline_id = self.worker_index
            num_replicas_in_sync = self.num_replicas_in_sync

            def get_per_replica_batch_size(_s, global_batch_size: int) -> int:
                return global_batch_size // self.num_replicas_in_sync

        ds = fn(_Ctx())
        b = _ds.find_batch_node(ds)
        per = b.batch_size if b is not None else 0
        return DistributedDataset(ds, self.num_local_replicas, per, per * self.num_replicas_in_sync)

    # ---- compute --------------------------------------------------------------------------
    def run(self, fn: Callable[..., Any], args: Tuple[Any, ...] = (), kwargs: Optional[dict] = None) -> PerReplica:
        """Call fn once per local replica with that replica's slice of every PerReplica argument."""
        kwargs = kwargs or {}
        outs = []
        for r in range(self.num_local_replicas):
            a = [x[r] if isinstance(x, PerReplica) else x for x in args]
            k = {n: (v[r] if isinstance(v, PerReplica) else v) for n, v in kwargs.items()}
            with replica_context(self, r):
                outs.append(fn(*a, **k))
        return PerReplica(outs)

    def reduce(self, reduce_op: str, value: Any, axis: Optional[int] = None) -> Any:
        """Reduce a PerReplica of scalars/tensors over ALL replicas in sync."""
        vals = list(value) if isinstance(value, PerReplica) else [value]
        arrs = [np.asarray(v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else v, dtype=np.float64)
                for v in vals]
        if axis is not None:
            arrs = [a.sum(axis=axis) if reduce_op == ReduceOp.SUM else a.mean(axis=axis) for a in arrs]
        local = np.sum(np.stack(arrs), axis=0)
        tot = self.cross_device_ops.allreduce_scalars(np.atleast_1d(local)).reshape(np.shape(local))
        if reduce_op == ReduceOp.MEAN:
            tot = tot / self.num_replicas_in_sync
        return tot

    def experimental_local_results(self, value: Any) -> Tuple[Any, ...]:
        return tuple(value) if isinstance(value, PerReplica) else (value,)

    def barrier(self) -> None:
        self.cross_device_ops.barrier()

    def close(self) -> None:
        """Shut the worker's service down when training is over (README.md:68)."""
        if not self._closed:
            self._closed = True
            self.cross_device_ops.close()

    def __repr__(self) -> str:
        return (f"<{type(self).__name__} workers={self.num_workers} worker_index={self.worker_index} "
                f"local_devices={[str(d) for d in self.local_devices]} comm={self.cross_device_ops.name}>")


@contextlib.contextmanager
def replica_context(strategy: Strategy, replica_id: int):
    prev = getattr(_tls, "replica", None)
    _tls.replica = (strategy, replica_id)
    try:
        yield
    finally:
        _tls.replica = prev


def get_replica_context() -> Optional[Tuple[Strategy, int]]:
    return getattr(_tls, "replica", None)


def _parse_devices(devices: Optional[Sequence[Union[str, torch.device]]]) -> List[torch.device]:
    if devices is None:
        n = torch.cuda.device_count() if torch.cuda.is_available() else 0
        return [torch.device("cuda", i) for i in range(n)] or [torch.device("cpu")]
    out = []
    for d in devices:
        if isinstance(d, torch.device):
            out.append(d)
            continue
        s = str(d).lower().replace("/device:", "").replace("/", "")
        s = s.replace("gpu", "cuda")
        kind, _, idx = s.partition(":")
        if kind == "cpu":
            out.append(torch.device("cpu"))  # several CPU replicas are allowed (tests)
        else:
            out.append(torch.device(kind, int(idx or 0)))
    return out


def _torchrun_env() -> Optional[Tuple[int, int, int]]:
    try:
        w = int(os.environ.get("WORLD_SIZE", "1"))
        if w > 1:
            return int(os.environ["RANK"]), w, int(os.environ.get("LOCAL_RANK", os.environ["RANK"]))
    except (KeyError, ValueError):
        pass
    return None


def _ensure_process_group(rank: int, world: int, device: torch.device, master: Optional[str] = None) -> None:
    """torch.distributed is the plumbing for the GPU process group (symmetric-memory allocation and
    handle exchange); the gradient collective itself never goes through it."""
    import torch.distributed as dist

    if dist.is_initialized():
        return
    if master:
        host, _, port = master.rpartition(":")
        os.environ.setdefault("MASTER_ADDR", host)
        os.environ.setdefault("MASTER_PORT", str(int(port) + 1))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    if device.type == "cuda":
        torch.cuda.set_device(device)
        dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world, device_id=device)
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world)


class OneDeviceStrategy(Strategy):
    """Default when no scope is active: one replica on the default device."""

    def __init__(self, device: Optional[Union[str, torch.device]] = None) -> None:
        dev = _parse_devices([device])[0] if device is not None else (
            torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu"))
        super().__init__([dev], TFConfigClusterResolver.single_worker(), LocalAllReduce(), 1, 0)


class MirroredStrategy(Strategy):
    """Synchronous data parallelism on one host (README.md:15-19; tf_dist_example.py:13).

    `devices` defaults to every visible GPU (CPU if none). Under a one-process-per-GPU launch
    (torchrun) each process holds one replica and the replicas of all processes are in sync.
    """

    def __init__(self, devices: Optional[Sequence[Union[str, torch.device]]] = None,
                 cross_device_ops: Optional[CrossDeviceOps] = None) -> None:
        from .cross_device_ops import NcclAllReduce

        if isinstance(cross_device_ops, NcclAllReduce):
            cross_device_ops = None  # TensorFlow's default, and the default here: the NVLink kernels where possible
        env = _torchrun_env() if devices is None else None
        if env is not None:
            rank, world, local_rank = env
            if torch.cuda.is_available():
                dev = torch.device("cuda", local_rank % torch.cuda.device_count())
            else:
                dev = torch.device("cpu")
            if dev.type == "cuda":
                from ..utils.affinity import bind_to_device

                bind_to_device(dev, int(os.environ.get("LOCAL_WORLD_SIZE", world)))  # before any helper thread exists
            _ensure_process_group(rank, world, dev)
            ops = cross_device_ops or _make_group_ops(dev, CollectiveCommunication.AUTO, _launcher_control_group(rank, world, dev))
            super().__init__([dev], TFConfigClusterResolver.single_worker(), ops, world, rank)
            return
        devs = _parse_devices(devices)
        ops = cross_device_ops
        if ops is None and len(devs) > 1 and all(d.type == "cuda" for d in devs) and \
                os.environ.get("TFDL_MIRRORED_NATIVE", "1") == "1":
            # one process driving several GPUs (the form the reference names, tf_dist_example.py:13): the replicas' native
            # step programs exchange gradients over NVLink peer memory, like the one-process-per-GPU launch
            from ..ops.allreduce import MirroredPeerAllReduce

            if MirroredPeerAllReduce.possible(devs):
                ops = MirroredPeerAllReduce(devs)
        super().__init__(devs, TFConfigClusterResolver.single_worker(), ops or LocalAllReduce(), 1, 0)


class MultiWorkerMirroredStrategy(Strategy):
    """Synchronous data parallelism over the TF_CONFIG cluster (README.md:21-29; the path the
    reference script runs, tf_dist_example.py:12).

    Each worker process holds a full copy of every variable. `communication` chooses the gradient
    collective: RING = TCP ring over the cluster ports (CPU workers), NCCL = the fused NVLink
    kernels, AUTO = NVLink kernels when every worker drives a CUDA device, else RING. With a single
    worker it degenerates to MirroredStrategy (README.md:34).
    """

    def __init__(self, communication: CollectiveCommunication = CollectiveCommunication.AUTO,
                 cluster_resolver: Optional[TFConfigClusterResolver] = None, *,
                 tf_config: Union[None, str, dict] = None, devices: Optional[Sequence[Union[str, torch.device]]] = None,
                 timeout_s: float = 120.0, heartbeat_timeout_s: float = 10.0, communication_options: Any = None) -> None:
        if communication_options is not None:  # tf.distribute.MultiWorkerMirroredStrategy(communication_options=...)
            communication = communication_options.implementation
            if communication_options.timeout_seconds:
                timeout_s = float(communication_options.timeout_seconds)
        if isinstance(communication, str):
            communication = CollectiveCommunication[communication.upper()]
        self.communication = communication
        res = cluster_resolver or TFConfigClusterResolver.from_env(tf_config)
        if not res.is_training_task:
            raise ValueError(f"task type {res.task_type!r} is not a training replica; MultiWorkerMirroredStrategy "
                             "runs on 'chief' and 'worker' tasks only ('ps'/'evaluator' belong to ParameterServerStrategy)")
        devs = _parse_devices(devices)
        if res.is_single_worker:
            super().__init__(devs, res, LocalAllReduce(), 1, 0)
            return
        from ..runtime import native

        group = native.Group(res.id_in_cluster, res.worker_addresses, timeout_s=timeout_s,
                             hb_timeout_s=heartbeat_timeout_s)
        self.group = group
        all_cuda = all(d.type == "cuda" for d in devs)
        # AUTO chooses by hardware and topology (README.md:21): the NVLink peer-memory kernels need every worker to
        # drive one CUDA device of the SAME host (one NVLink / NVSwitch domain); workers spread over machines - the
        # reference's own 172.16.16.5 / .6 deployment - get the TCP ring.
        facts = group.allgather_object((all_cuda and len(devs) == 1, _host_identity()))
        one_host = len({h for _, h in facts}) == 1
        use_p2p = communication != CollectiveCommunication.RING and all(k for k, _ in facts) and one_host
        if communication == CollectiveCommunication.NCCL and not use_p2p:
            group.close()
            why = ("the workers run on different hosts (the NVLink peer-memory kernels need one NVLink domain; use RING "
                   "or AUTO across machines)" if not one_host else
                   "every worker must drive exactly one CUDA device")
            raise ValueError(f"CollectiveCommunication.NCCL is not possible here: {why}")
        if use_p2p:
            from ..utils.affinity import bind_to_device

            bind_to_device(devs[0], res.num_workers)  # all workers are on this host: share its sockets by GPU locality
            _ensure_process_group(res.id_in_cluster, res.num_workers, devs[0], master=res.leader_address)
            ops: CrossDeviceOps = _make_group_ops(devs[0], communication, group)
        else:
            ops = TcpRingAllReduce(group)
        super().__init__(devs, res, ops, res.num_workers, res.id_in_cluster)


def _launcher_control_group(rank: int, world: int, dev: torch.device) -> Any:
    """Under a torchrun-style launch there is no TF_CONFIG and so no cluster service - but the GPU collectives still
    need failure detection (a dead peer must release the survivors' kernels). For a single-node launch a control group
    with its heartbeat thread is opened on MASTER_ADDR, ports MASTER_PORT+17+rank. TFDL_HEARTBEAT=0 switches it off; if
    the ports cannot be had, training goes on without it (a lost peer then ends in the kernels' spin limit)."""
    if dev.type != "cuda" or os.environ.get("TFDL_HEARTBEAT", "1") == "0":
        return None
    if int(os.environ.get("LOCAL_WORLD_SIZE", str(world))) != world:
        return None  # several nodes: ranks do not share MASTER_ADDR's host
    try:
        from ..runtime import native

        host = os.environ.get("MASTER_ADDR", "127.0.0.1")
        base = int(os.environ.get("MASTER_PORT", "29500")) + 17
        addrs = [f"{host}:{base + r}" for r in range(world)]
        return native.Group(rank, addrs, timeout_s=float(os.environ.get("TFDL_HEARTBEAT_CONNECT_S", "20")),
                            hb_timeout_s=float(os.environ.get("TFDL_HEARTBEAT_TIMEOUT_S", "10")))
    except Exception as e:  # noqa: BLE001
        import sys

        print(f"[tfdl] no heartbeat for this launch ({type(e).__name__}: {e}); a lost worker will not be detected early",
              file=sys.stderr)
        return None


def _host_identity() -> str:
    """Something equal for processes of one machine and different across machines: the kernel's boot id (falls back to
    the host name). TFDL_HOST_ID overrides it (tests; containers that share a boot id but not an NVLink domain)."""
    env = os.environ.get("TFDL_HOST_ID")
    if env:
        return env
    try:
        with open("/proc/sys/kernel/random/boot_id") as f:
            return f.read().strip()
    except OSError:
        import socket

        return socket.gethostname()


def _make_group_ops(dev: torch.device, communication: CollectiveCommunication, group: Any) -> CrossDeviceOps:
    if dev.type == "cuda":
        from ..ops.allreduce import P2PFusedAllReduce

        return P2PFusedAllReduce(dev, control_group=group)
    return TorchDistAllReduce()
