"""GPU-rank replica backend: north-star subsystem (1), SURVEY.md §8f row 1.

Where ``TFReplicaSet.Create`` (pkg/trainer/replicas.go:124-271) creates one Service and
one batch/v1 Job per task index, this backend records the launch and, on ``run``, executes
the replicas of one box: every (type, index) is mapped to a GPU rank by
:class:`k8s_b200.cluster.RankMap`, its TF_CONFIG is placed in the environment exactly as
the container would see it, and its exit code feeds the reference's completion contract
(pkg/trainer/training.go:203-238).
"""
from __future__ import annotations

import json
import os
import traceback
from typing import Any, Callable, Dict, Optional

from . import spec as S
from .cluster import RankMap
from .lib import TfgError
from .trainer import ReplicaBackend, ReplicaLaunch


class GpuRankBackend(ReplicaBackend):
    def __init__(self, n_gpus: int = 1, workload: Optional[Callable[..., Any]] = None):
        self.n_gpus = n_gpus
        self.launches: Dict[str, ReplicaLaunch] = {}
        self.exits: Dict[str, int] = {}
        self.results: Dict[str, Any] = {}
        self.workload = workload

    # -- ReplicaBackend -------------------------------------------------------------------
    def create_replica(self, launch: ReplicaLaunch) -> None:
        self.launches.setdefault(launch.name, launch)   # AlreadyExists is tolerated

    def replica_exit(self, name: str):
        if name not in self.launches:
            return (False, None, "")
        return (True, self.exits.get(name), "")

    def delete_replica(self, name: str) -> None:
        self.launches.pop(name, None)
        self.exits.pop(name, None)

    # -- execution ------------------------------------------------------------------------
    def rank_of(self, launch: ReplicaLaunch) -> int:
        cluster = json.loads(launch.tf_config)["cluster"]
        return RankMap(cluster, self.n_gpus).gpu_of(launch.replica_type.lower(), launch.index)

    def run(self) -> None:
        """Run every recorded replica that has not exited.  PS and WORKER replicas of the
        sample workload only serve (tf_smoke.py:126-132), so they stay Running like the
        reference's pods; the MASTER's exit decides the job."""
        from . import tf_smoke
        for name, launch in self.launches.items():
            if name in self.exits:
                continue
            rtype = launch.replica_type
            if rtype != S.MASTER:
                continue
            env_before = os.environ.get("TF_CONFIG")
            os.environ["TF_CONFIG"] = launch.tf_config
            try:
                fn = self.workload or tf_smoke.main
                self.results[name] = fn(n_gpus=self.n_gpus, gpu_rank=self.rank_of(launch))
                self.exits[name] = 0
            except TfgError as e:
                self.results[name] = e
                self.exits[name] = e.code          # 1..127 permanent, >=128 retryable
            except Exception:                      # a Python replica dying = exit code 1
                self.results[name] = traceback.format_exc()
                self.exits[name] = 1
            finally:
                if env_before is None:
                    os.environ.pop("TF_CONFIG", None)
                else:
                    os.environ["TF_CONFIG"] = env_before
