"""tf.distribute look-alike - ref: /root/reference/tf_dist_example.py:12-13,36,56; README.md:13-29."""
from .cross_device_ops import (CollectiveCommunication, CommunicationImplementation, CommunicationOptions,
                               CrossDeviceOps, HierarchicalCopyAllReduce, LocalAllReduce, NcclAllReduce,
                               ReductionToOneDevice, TcpRingAllReduce, TorchDistAllReduce)
from .strategy import (DistributedDataset, MirroredStrategy, MultiWorkerMirroredStrategy, OneDeviceStrategy, ReduceOp,
                       Strategy, get_replica_context, get_strategy, has_strategy)
from .values import MirroredVariable, ParamArena, PerReplica, SyncOnReadVariable, VarSpec


class experimental:  # tf.distribute.experimental namespace look-alike (tf_dist_example.py:12)
    MultiWorkerMirroredStrategy = MultiWorkerMirroredStrategy
    CollectiveCommunication = CollectiveCommunication
    CommunicationImplementation = CollectiveCommunication
    CommunicationOptions = CommunicationOptions


__all__ = ["CollectiveCommunication", "CommunicationImplementation", "CommunicationOptions", "CrossDeviceOps",
           "HierarchicalCopyAllReduce", "NcclAllReduce", "ReductionToOneDevice", "LocalAllReduce",
           "TcpRingAllReduce", "TorchDistAllReduce", "DistributedDataset", "MirroredStrategy",
           "MultiWorkerMirroredStrategy", "OneDeviceStrategy", "ReduceOp", "Strategy", "get_replica_context",
           "get_strategy", "has_strategy", "MirroredVariable", "ParamArena", "PerReplica", "SyncOnReadVariable",
           "VarSpec", "experimental"]
