"""Cluster configuration (TF_CONFIG) - ref: /root/reference/tf_dist_example.py:6-10, README.md:32-61."""
from .tf_config import (ClusterSpec, TFConfigClusterResolver, TFConfigError, make_tf_config, KNOWN_JOBS,
                        TRAINING_JOBS)

__all__ = ["ClusterSpec", "TFConfigClusterResolver", "TFConfigError", "make_tf_config", "KNOWN_JOBS",
           "TRAINING_JOBS"]
