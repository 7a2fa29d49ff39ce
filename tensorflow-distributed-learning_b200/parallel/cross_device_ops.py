"""Gradient aggregation back ends (the reference's `CollectiveCommunication` choice).

README.md:17,21,23 of the reference: MultiWorkerMirroredStrategy offers RING (ring all-reduce
over the cluster's RPC ports), NCCL, and AUTO (runtime picks by hardware / topology / tensor
size); MirroredStrategy defaults to NcclAllReduce. Here:

  RING  -> TcpRingAllReduce : C++ ring all-reduce over the TF_CONFIG TCP ports (CPU workers)
  NCCL  -> P2PFusedAllReduce: hand-written sm_100a kernels over NVLink peer memory that fuse the
           reduction with the 1/N scale and the SGD/momentum update (ops/allreduce.py) - this is
           the "NCCL slot", but no NCCL call runs on it
  AUTO  -> P2PFused when every worker drives one CUDA device of the SAME host, else RING; inside
           P2PFused the push / push2 / two-shot / multimem variant is picked per message size and group size.
  (single-process MirroredStrategy over several GPUs: ops/allreduce.py:MirroredPeerAllReduce, same kernels)

Every implementation reduces the *flat gradient arena* in place (values.ParamArena.grads), so
there is no packing step, and every implementation produces bit-identical results on all replicas.
"""
from __future__ import annotations

import enum
from typing import List, Optional, Sequence

import numpy as np
import torch


class CollectiveCommunication(enum.Enum):
    """tf.distribute.experimental.CollectiveCommunication (tf_dist_example.py:12)."""

    AUTO = "AUTO"
    RING = "RING"
    NCCL = "NCCL"


CommunicationImplementation = CollectiveCommunication  # newer TF name


class CrossDeviceOps:
    """Sum gradient arenas over all replicas in sync (local replicas x workers)."""

    name = "base"

    def allreduce_grads(self, grads: Sequence[torch.Tensor]) -> None:
        """In place: afterwards every tensor in `grads` (one per local replica) holds the sum over
        all replicas of all workers."""
        raise NotImplementedError

    def allreduce_scalars(self, vals: np.ndarray) -> np.ndarray:
        """Sum a small float64 vector over workers (metric aggregation, K16)."""
        return vals

    def broadcast_params(self, params: Sequence[torch.Tensor]) -> None:
        """Make every replica's parameter arena equal to the chief's replica 0 (K15)."""
        for p in params[1:]:
            p.copy_(params[0])

    def barrier(self) -> None:
        pass

    def close(self) -> None:
        pass


def _sum_local(grads: Sequence[torch.Tensor]) -> torch.Tensor:
    """Deterministic local reduction: replica order, accumulated into replica 0's buffer."""
    acc = grads[0]
    for g in grads[1:]:
        acc.add_(g.to(acc.device, non_blocking=True))
    return acc


class LocalAllReduce(CrossDeviceOps):
    """Single process, any number of local replicas, no peers."""

    name = "local"

    def allreduce_grads(self, grads: Sequence[torch.Tensor]) -> None:
        if len(grads) == 1:
            return
        acc = _sum_local(grads)
        for g in grads[1:]:
            g.copy_(acc, non_blocking=True)


class NcclAllReduce:
    """tf.distribute.NcclAllReduce - MirroredStrategy's default `cross_device_ops` (README.md:17 of the reference). A
    request, not an implementation: the strategy answers it with this framework's NVLink peer-memory kernels
    (ops/allreduce.py) when its devices are CUDA GPUs with peer access, else with the local sum. `num_packs` is accepted
    for compatibility; the gradients already live in one flat arena."""

    def __init__(self, num_packs: int = 1) -> None:
        self.num_packs = num_packs


class ReductionToOneDevice(LocalAllReduce):
    """tf.distribute.ReductionToOneDevice: sum on replica 0's device, copy the result back to the others."""

    name = "reduction_to_one_device"

    def __init__(self, reduce_to_device=None, accumulation_fn=None) -> None:  # arguments accepted for compatibility
        super().__init__()


class HierarchicalCopyAllReduce(LocalAllReduce):
    """tf.distribute.HierarchicalCopyAllReduce. On NVSwitch machines there is no hierarchy to exploit: the same local
    sum-and-copy as ReductionToOneDevice."""

    name = "hierarchical_copy"

    def __init__(self, num_packs: int = 1) -> None:
        super().__init__()


class CommunicationOptions:
    """tf.distribute.experimental.CommunicationOptions: `implementation` chooses RING / NCCL / AUTO for
    MultiWorkerMirroredStrategy(communication_options=...); bytes_per_pack and timeout_seconds are accepted
    (buckets are sized by this framework; the collective timeout is the strategy's `timeout_s`)."""

    def __init__(self, bytes_per_pack: int = 0, timeout_seconds: Optional[float] = None,
                 implementation: "CollectiveCommunication" = CollectiveCommunication.AUTO) -> None:
        if isinstance(implementation, str):
            implementation = CollectiveCommunication[implementation.upper()]
        self.bytes_per_pack, self.timeout_seconds, self.implementation = bytes_per_pack, timeout_seconds, implementation


class TcpRingAllReduce(CrossDeviceOps):
    """RING: local sum, then the C++ TCP ring across workers, then fan the result back out."""

    name = "ring"

    def __init__(self, group) -> None:
        self.group = group
        self._stage: Optional[torch.Tensor] = None

    def allreduce_grads(self, grads: Sequence[torch.Tensor]) -> None:
        acc = _sum_local(grads) if len(grads) > 1 else grads[0]
        if self.group.world > 1:
            if acc.device.type == "cpu":
                assert acc.is_contiguous() and acc.dtype == torch.float32
                self.group.allreduce_ptr_f32(acc.data_ptr(), acc.numel())
            else:  # GPU worker on the RING path: stage through pinned host memory
                if self._stage is None or self._stage.numel() != acc.numel():
                    self._stage = torch.empty(acc.numel(), dtype=torch.float32, pin_memory=True)
                self._stage.copy_(acc, non_blocking=False)
                self.group.allreduce_ptr_f32(self._stage.data_ptr(), acc.numel())
                acc.copy_(self._stage, non_blocking=False)
        for g in grads[1:]:
            g.copy_(acc, non_blocking=True)

    def allreduce_scalars(self, vals: np.ndarray) -> np.ndarray:
        v = np.ascontiguousarray(vals, dtype=np.float64).copy()
        if self.group.world > 1:
            self.group.allreduce_(v)
        return v

    def broadcast_params(self, params: Sequence[torch.Tensor]) -> None:
        if self.group.world > 1:
            host = params[0].detach().cpu().contiguous().numpy()
            self.group.broadcast_array(host)
            params[0].copy_(torch.from_numpy(host))
        for p in params[1:]:
            p.copy_(params[0])

    def barrier(self) -> None:
        self.group.barrier()

    def close(self) -> None:
        self.group.close()


class TorchDistAllReduce(CrossDeviceOps):
    """Stand-in baseline: torch.distributed (NCCL on GPU, gloo on CPU). This is the comparison
    arm of the benchmarks ("a path that only calls NCCL is the baseline, not the product")."""

    name = "torch.distributed"

    def __init__(self) -> None:
        import torch.distributed as dist

        if not dist.is_initialized():
            raise RuntimeError("TorchDistAllReduce needs an initialised torch.distributed process group")
        self.dist = dist

    def allreduce_grads(self, grads: Sequence[torch.Tensor]) -> None:
        acc = _sum_local(grads) if len(grads) > 1 else grads[0]
        self.dist.all_reduce(acc)
        for g in grads[1:]:
            g.copy_(acc, non_blocking=True)

    def allreduce_scalars(self, vals: np.ndarray) -> np.ndarray:
        t = torch.from_numpy(np.ascontiguousarray(vals, dtype=np.float64).copy())
        dev = "cuda" if self.dist.get_backend() == "nccl" else "cpu"
        t = t.to(dev)
        self.dist.all_reduce(t)
        return t.cpu().numpy()

    def broadcast_params(self, params: Sequence[torch.Tensor]) -> None:
        self.dist.broadcast(params[0], src=0)
        for p in params[1:]:
            p.copy_(params[0])

    def barrier(self) -> None:
        self.dist.barrier()
