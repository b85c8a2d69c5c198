"""Cluster spec -> GPU ranks of one box (north-star subsystem (1), SURVEY.md §5 last row).

The reference addresses a task by a Service DNS name ``<jobName(i)>:<TfPort>``
(pkg/trainer/training.go:121, replicas.go:157-174).  On one 8xB200 box those strings
become rank labels: training replicas are the master task then the worker tasks in index
order, replica ``r`` computes on GPU ``r % n_gpus``; PS task ``k`` is owned by GPU
``k % n_gpus``.  The same rule is implemented behind the C ABI (``tfg_replica_gpu`` /
``tfg_ps_gpu``); this module is the host-side view the trainer backend uses, and the tests
hold the two against each other.
"""
from __future__ import annotations

from typing import Dict, List, Tuple


class RankMap:
    def __init__(self, cluster: Dict[str, List[str]], n_gpus: int):
        if n_gpus < 1:
            raise ValueError("n_gpus must be >= 1")
        self.cluster = {k: list(v) for k, v in cluster.items()}
        self.n_gpus = n_gpus
        self.n_master = len(self.cluster.get("master", []))
        self.n_worker = len(self.cluster.get("worker", []))
        self.n_ps = len(self.cluster.get("ps", []))
        self.n_replicas = max(1, self.n_master + self.n_worker)

    def replicas(self) -> List[Tuple[str, int]]:
        """Training replicas in replica order: (job, task index)."""
        out = [("master", i) for i in range(self.n_master)]
        out += [("worker", i) for i in range(self.n_worker)]
        return out or [("master", 0)]

    def replica_index(self, job: str, task: int) -> int:
        if job == "master":
            return task
        if job == "worker":
            return self.n_master + task
        raise ValueError("job %r is not a training replica" % job)

    def gpu_of(self, job: str, task: int) -> int:
        n = len(self.cluster.get(job, []))
        if task < 0 or task >= max(n, 1):
            raise ValueError("task %d out of range for job %r" % (task, job))
        if job == "ps":
            return task % self.n_gpus
        if job in ("master", "worker"):
            return self.replica_index(job, task) % self.n_gpus
        return task % self.n_gpus

    def local_replicas(self, gpu_rank: int) -> List[int]:
        return [r for r in range(self.n_replicas) if r % self.n_gpus == gpu_rank]

    def tasks(self) -> List[Tuple[str, int]]:
        """Every task of the cluster, jobs sorted by name (Go map marshalling order)."""
        return [(j, i) for j in sorted(self.cluster) for i in range(len(self.cluster[j]))]


def round_robin_placement(op_types: List[str], n_ps: int,
                          ps_ops=("Variable", "VariableV2", "VarHandleOp")) -> List[int]:
    """[TF1.3] ``tf.train.replica_device_setter`` restated (call site
    examples/tf_sample/tf_sample/tf_smoke.py:116-118): the k-th PS op created goes to
    ``/job:ps/task:(k mod n_ps)``; other ops (and everything when there is no ps job) stay
    on the worker device, reported as -1."""
    out, nxt = [], 0
    for t in op_types:
        if n_ps > 0 and t in ps_ops:
            out.append(nxt)
            nxt = (nxt + 1) % n_ps
        else:
            out.append(-1)
    return out
