"""tf.data-style pipeline - ref: /root/reference/tf_dist_example.py:20-37."""
from .dataset import (AUTOTUNE, AutoShardPolicy, Dataset, DistributeOptions, Options, auto_shard, rebatch_for_replicas,
                      set_global_seed, find_batch_node)
from . import datasets


class experimental:  # tf.data.experimental namespace look-alike
    AutoShardPolicy = AutoShardPolicy
    AUTOTUNE = AUTOTUNE


__all__ = ["AUTOTUNE", "AutoShardPolicy", "Dataset", "DistributeOptions", "Options", "auto_shard", "rebatch_for_replicas",
           "set_global_seed", "find_batch_node", "datasets", "experimental"]
