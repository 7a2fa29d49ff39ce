"""TF_CONFIG cluster resolver.

The reference configures its whole cluster through one environment variable
(/root/reference/tf_dist_example.py:6-10, README.md:32-61):

    TF_CONFIG = {"cluster": {"worker": ["host:port", ...], "chief": [...], "ps": [...], "evaluator": [...]},
                 "task": {"type": "worker", "index": 1}}

Rules reproduced here (README.md:51,59,82):
  * `cluster` is identical on every node, only `task` differs; `index` is zero-based and must
    address an entry of cluster[type];
  * with no `chief` job, worker 0 is the chief (it alone writes checkpoints / TensorBoard);
  * TF_CONFIG has to be set before the strategy is constructed - the resolver reads the
    environment at construction time, or takes an explicit dict/JSON override;
  * without TF_CONFIG the process is a single-worker cluster (README.md:34).
`ps` and `evaluator` jobs are parsed but are not training replicas of the mirrored strategies.
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass, field
from typing import Dict, List, Mapping, Optional, Tuple, Union

KNOWN_JOBS = ("chief", "worker", "ps", "evaluator")
TRAINING_JOBS = ("chief", "worker")


class TFConfigError(ValueError):
    """TF_CONFIG is malformed or inconsistent."""


def _parse_addr(addr: str) -> Tuple[str, int]:
    if not isinstance(addr, str) or ":" not in addr:
        raise TFConfigError(f"cluster address {addr!r} is not of the form host:port")
    host, _, port = addr.rpartition(":")
    try:
        p = int(port)
    except ValueError as e:
        raise TFConfigError(f"cluster address {addr!r} has a non-integer port") from e
    if not host or not (0 < p < 65536):
        raise TFConfigError(f"cluster address {addr!r} has an empty host or a port out of range")
    return host, p


@dataclass(frozen=True)
class ClusterSpec:
    """Job name -> ordered list of `host:port` task addresses."""

    jobs: Mapping[str, Tuple[str, ...]] = field(default_factory=dict)

    @staticmethod
    def from_dict(d: Mapping[str, object]) -> "ClusterSpec":
        if not isinstance(d, Mapping):
            raise TFConfigError("TF_CONFIG['cluster'] must be an object mapping job names to address lists")
        jobs: Dict[str, Tuple[str, ...]] = {}
        for job, tasks in d.items():
            if job not in KNOWN_JOBS:
                raise TFConfigError(f"unknown job {job!r}; expected one of {KNOWN_JOBS}")
            if isinstance(tasks, Mapping):  # sparse form {index: addr}
                idx = sorted(int(k) for k in tasks)
                if idx != list(range(len(idx))):
                    raise TFConfigError(f"job {job!r}: sparse task indices must be 0..n-1")
                tasks = [tasks[str(i)] if str(i) in tasks else tasks[i] for i in idx]
            if not isinstance(tasks, (list, tuple)):
                raise TFConfigError(f"job {job!r} must list its task addresses")
            for a in tasks:
                _parse_addr(a)
            if len(set(tasks)) != len(tasks):
                raise TFConfigError(f"job {job!r} lists an address twice")
            jobs[job] = tuple(tasks)
        if len(jobs.get("chief", ())) > 1:
            raise TFConfigError("there can be at most one chief task")
        return ClusterSpec(jobs)

    def as_dict(self) -> Dict[str, List[str]]:
        return {k: list(v) for k, v in self.jobs.items()}

    def num_tasks(self, job: str) -> int:
        return len(self.jobs.get(job, ()))

    def task_address(self, job: str, index: int) -> str:
        tasks = self.jobs.get(job)
        if tasks is None:
            raise TFConfigError(f"job {job!r} is not in the cluster")
        if not 0 <= index < len(tasks):
            raise TFConfigError(f"task index {index} out of range for job {job!r} ({len(tasks)} tasks)")
        return tasks[index]

    def training_tasks(self) -> List[Tuple[str, int, str]]:
        """(job, index, address) of every training replica host, chief first - this order
        defines the collective rank of each worker (`id_in_cluster`)."""
        out = []
        for job in TRAINING_JOBS:
            for i, a in enumerate(self.jobs.get(job, ())):
                out.append((job, i, a))
        return out


@dataclass(frozen=True)
class TFConfigClusterResolver:
    """Resolved view of TF_CONFIG for this process."""

    cluster_spec: ClusterSpec
    task_type: str
    task_id: int
    rpc_layer: str = "tcp"
    environment: str = ""

    # ------------------------------------------------------------------ construction
    @staticmethod
    def from_env(override: Union[None, str, Mapping[str, object]] = None,
                 env: Optional[Mapping[str, str]] = None) -> "TFConfigClusterResolver":
        """Resolve from an explicit override (dict or JSON string) or, failing that, from the
        TF_CONFIG environment variable. The reference script overwrites the variable in-process
        with a hard-coded index (tf_dist_example.py:6-10); passing `override` is the explicit
        form of the same thing and leaves the environment untouched."""
        env = os.environ if env is None else env
        raw = override if override is not None else env.get("TF_CONFIG", "")
        if isinstance(raw, str):
            if not raw.strip():
                return TFConfigClusterResolver.single_worker()
            try:
                cfg = json.loads(raw)
            except json.JSONDecodeError as e:
                raise TFConfigError(f"TF_CONFIG is not valid JSON: {e}") from e
        else:
            cfg = raw
        if not isinstance(cfg, Mapping):
            raise TFConfigError("TF_CONFIG must be a JSON object")
        cluster = cfg.get("cluster")
        task = cfg.get("task") or {}
        if not cluster:
            return TFConfigClusterResolver.single_worker()
        spec = ClusterSpec.from_dict(cluster)
        if not isinstance(task, Mapping) or "type" not in task:
            raise TFConfigError("TF_CONFIG['task'] needs 'type' and 'index' when a cluster is given")
        ttype = task["type"]
        try:
            tid = int(task.get("index", 0))
        except (TypeError, ValueError) as e:
            raise TFConfigError("TF_CONFIG['task']['index'] must be an integer") from e
        if ttype not in KNOWN_JOBS:
            raise TFConfigError(f"task type {ttype!r} is not one of {KNOWN_JOBS}")
        if ttype == "evaluator" and "evaluator" not in spec.jobs:
            # TF allows an evaluator outside the cluster spec; it is not a training replica anyway
            if tid != 0:
                raise TFConfigError("an evaluator outside the cluster spec must have index 0")
        else:
            spec.task_address(ttype, tid)  # range check
        return TFConfigClusterResolver(spec, ttype, tid, cfg.get("rpc_layer", "tcp"),
                                       cfg.get("environment", ""))

    @staticmethod
    def single_worker() -> "TFConfigClusterResolver":
        return TFConfigClusterResolver(ClusterSpec({}), "worker", 0)

    # ------------------------------------------------------------------ derived facts
    @property
    def is_single_worker(self) -> bool:
        return self.num_workers <= 1

    @property
    def num_workers(self) -> int:
        """Number of training hosts (chief + workers); 1 when there is no cluster."""
        return max(1, len(self.cluster_spec.training_tasks()))

    @property
    def is_training_task(self) -> bool:
        return self.task_type in TRAINING_JOBS

    @property
    def is_chief(self) -> bool:
        """chief job if present, else worker 0 (README.md:51)."""
        if not self.cluster_spec.jobs:
            return True
        if self.task_type == "chief":
            return True
        if self.task_type == "worker" and self.task_id == 0 and self.cluster_spec.num_tasks("chief") == 0:
            return True
        return False

    @property
    def id_in_cluster(self) -> int:
        """Collective rank of this worker: chief (if any) is 0, then workers in list order."""
        if not self.cluster_spec.jobs:
            return 0
        if not self.is_training_task:
            raise TFConfigError(f"task type {self.task_type!r} is not a training replica of a mirrored strategy")
        for rank, (job, idx, _) in enumerate(self.cluster_spec.training_tasks()):
            if job == self.task_type and idx == self.task_id:
                return rank
        raise TFConfigError("this task is not listed in the cluster")

    @property
    def address(self) -> Optional[str]:
        if not self.cluster_spec.jobs or self.task_type not in self.cluster_spec.jobs:
            return None
        return self.cluster_spec.task_address(self.task_type, self.task_id)

    @property
    def worker_addresses(self) -> List[str]:
        """Addresses of all training hosts in rank order."""
        return [a for _, _, a in self.cluster_spec.training_tasks()]

    @property
    def leader_address(self) -> Optional[str]:
        """The collective leader is the chief's task (collective_leader in TF)."""
        w = self.worker_addresses
        return w[0] if w else None

    def master(self) -> str:
        return self.leader_address or ""

    def to_tf_config(self) -> str:
        return json.dumps({"cluster": self.cluster_spec.as_dict(),
                           "task": {"type": self.task_type, "index": self.task_id}})


def make_tf_config(workers: List[str], task_type: str, index: int, chief: Optional[str] = None) -> str:
    """Helper for launching several cluster nodes on one machine (README.md:61)."""
    cluster: Dict[str, List[str]] = {"worker": list(workers)}
    if chief:
        cluster["chief"] = [chief]
    return json.dumps({"cluster": cluster, "task": {"type": task_type, "index": index}})
