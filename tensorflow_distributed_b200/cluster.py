"""Cluster spec, role/task validation, chief election and worker device choice.

Replaces ``tf.train.ClusterSpec`` / the device-string logic of the reference
(``/root/reference/mnist_python_m.py:137-172``).  A host entry is either a
classic ``host:port`` pair (CPU / gloo transport, one OS process per task) or
``gpu:<ordinal>`` which maps the task onto one B200 of the local box (the ps
role is then a GPU that owns parameter shards in HBM).
ps and worker tasks may name the same GPU (colocated shards).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

from .config import Flags


@dataclass(frozen=True)
class TaskAddress:
    job: str
    index: int
    raw: str

    @property
    def is_gpu(self) -> bool:
        return self.raw.startswith("gpu:")

    @property
    def gpu_ordinal(self) -> Optional[int]:
        return int(self.raw.split(":", 1)[1]) if self.is_gpu else None

    @property
    def host_port(self) -> Tuple[str, int]:
        host, port = self.raw.rsplit(":", 1)
        return host, int(port)


class ClusterSpec:
    """``{"ps": [...], "worker": [...]}`` (m:146-154)."""

    def __init__(self, jobs: Dict[str, List[str]]):
        for job, hosts in jobs.items():
            if not hosts or any(h.strip() == "" for h in hosts):
                raise ValueError("job %r has an empty host entry" % job)
        self._jobs = {j: [h.strip() for h in hosts] for j, hosts in jobs.items()}

    @classmethod
    def from_flags(cls, flags: Flags) -> "ClusterSpec":
        return cls({"ps": flags.ps_hosts.split(","), "worker": flags.worker_hosts.split(",")})

    @property
    def jobs(self) -> List[str]:
        return list(self._jobs)

    def num_tasks(self, job: str) -> int:
        return len(self._jobs[job])

    def job_tasks(self, job: str) -> List[str]:
        return list(self._jobs[job])

    def task_address(self, job: str, index: int) -> TaskAddress:
        hosts = self._jobs[job]
        if not 0 <= index < len(hosts):
            raise ValueError("task_index %d out of range for job %r (%d tasks)" % (index, job, len(hosts)))
        return TaskAddress(job, index, hosts[index])

    def as_dict(self) -> Dict[str, List[str]]:
        return {j: list(h) for j, h in self._jobs.items()}

    @property
    def num_workers(self) -> int:  # m:150
        return self.num_tasks("worker")

    @property
    def num_ps(self) -> int:
        return self.num_tasks("ps")

    @property
    def all_gpu(self) -> bool:
        return all(h.startswith("gpu:") for hs in self._jobs.values() for h in hs)

    # ---- rank mapping for torch.distributed bootstrap: ps tasks first, then workers ----
    @property
    def world_size(self) -> int:
        return self.num_ps + self.num_workers

    def rank_of(self, job: str, index: int) -> int:
        self.task_address(job, index)
        return index if job == "ps" else self.num_ps + index

    def task_of_rank(self, rank: int) -> Tuple[str, int]:
        return ("ps", rank) if rank < self.num_ps else ("worker", rank - self.num_ps)

    def master_endpoint(self) -> Tuple[str, int]:
        """Rendezvous endpoint: ps task 0's host:port (the only address every task knows)."""
        a = self.task_address("ps", 0)
        if a.is_gpu:
            import os
            return "127.0.0.1", int(os.environ.get("TFD_MASTER_PORT", 29500 + (a.gpu_ordinal or 0)))
        return a.host_port


def validate_role(flags: Flags) -> None:
    """m:137-140, plus the cases the reference lets fall through."""
    if flags.job_name is None or flags.job_name == "":
        raise ValueError("Must specify an explicit `job_name`")
    if flags.task_index is None or flags.task_index == "":
        raise ValueError("Must specify an explicit `task_index`")
    if flags.job_name not in ("ps", "worker"):
        raise ValueError("job_name must be 'ps' or 'worker', got %r" % flags.job_name)
    if flags.task_index < 0:
        raise ValueError("task_index must be >= 0")
    if flags.num_gpus < 0:
        # the reference leaves worker_device undefined here (m:164-172)
        raise ValueError("num_gpus must be >= 0")
    if flags.existing_servers and flags.job_name == "ps":
        # the reference would fall through and build a worker graph
        raise ValueError("existing_servers=True is a worker-side option; a ps task is the server")


def is_chief(flags: Flags) -> bool:
    """Static chief election: task 0 (m:163)."""
    return flags.task_index == 0


def worker_device(flags: Flags, cluster: Optional[ClusterSpec] = None) -> str:
    """Device string of this worker (m:164-172).

    ``gpu:<n>`` entries in worker_hosts win; otherwise ``num_gpus>0`` maps
    task i to GPU ``i % num_gpus`` and ``num_gpus==0`` to the CPU.
    """
    if cluster is not None and flags.job_name == "worker":
        a = cluster.task_address("worker", flags.task_index)
        if a.is_gpu:
            return "/job:worker/task:%d/gpu:%d" % (flags.task_index, a.gpu_ordinal)
    if flags.num_gpus > 0:
        return "/job:worker/task:%d/gpu:%d" % (flags.task_index, flags.task_index % flags.num_gpus)
    return "/job:worker/task:%d/cpu:0" % flags.task_index


def torch_device_of(device_string: str) -> str:
    tail = device_string.rsplit("/", 1)[1]
    kind, idx = tail.split(":")
    return "cuda:%s" % idx if kind == "gpu" else "cpu"


def effective_replicas_to_aggregate(flags: Flags, num_workers: int) -> int:
    """Documented intent of the flag: default = num_workers (m:62-65)."""
    r = flags.replicas_to_aggregate
    if r is None or r <= 0:
        return num_workers
    return r


def device_filters(flags: Flags) -> List[str]:
    """A worker only sees the ps job and itself (m:258)."""
    return ["/job:ps", "/job:worker/task:%d" % flags.task_index]


def filter_allows(filters: List[str], job: str, task_index: int) -> bool:
    """tf.ConfigProto(device_filters=...) semantics: a remote task is visible when one filter is a prefix of its name
    ('/job:ps' admits every ps task, '/job:worker/task:1' only that worker)."""
    name = "/job:%s/task:%d" % (job, task_index)
    return any(name == f or name.startswith(f + "/") for f in filters)
