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
import signal
import subprocess
import sys
import tempfile
import traceback
from typing import Any, Callable, Dict, List, Optional

from . import spec as S
from .cluster import RankMap
from .lib import TfgError
from .trainer import ReplicaBackend, ReplicaLaunch, is_retryable_termination_state as is_retryable_exit_code


class GpuRankBackend(ReplicaBackend):
    def __init__(self, n_gpus: int = 1, workload: Optional[Callable[..., Any]] = None, max_restarts: int = 2):
        self.n_gpus = n_gpus
        self.launches: Dict[str, ReplicaLaunch] = {}
        self.exits: Dict[str, int] = {}
        self.results: Dict[str, Any] = {}
        self.restarts: Dict[str, int] = {}
        self.max_restarts = max_restarts
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
                # a retryable exit (>= 128, not OOMKilled) leaves the replica Running in the reference because
                # Kubernetes restarts the container (restartPolicy: OnFailure, training.go:203-238): run it again
                if is_retryable_exit_code(self.exits[name]) and self.restarts.get(name, 0) < self.max_restarts:
                    self.restarts[name] = self.restarts.get(name, 0) + 1
                    del self.exits[name]
                else:
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


class ProcessBackend(ReplicaBackend):
    """One OS process per (type, index), as the reference starts one pod per task index
    (pkg/trainer/replicas.go:152-268): the launcher that lets an unchanged TfJob document drive the
    synchronous PS step on the GPU ranks of this box.

    * MASTER / WORKER run ``python -m k8s_b200.tf_sync_ps <workload_args>`` (or ``tf_smoke`` for the
      sample workload) with the trainer's TF_CONFIG in the environment;
    * a default PS keeps the reference's command line (replicas.go:205-208), with the interpreter and
      the script path replaced by ``<python> -m k8s_b200.grpc_tensorflow_server``;
    * ``replica_exit`` reports the process's exit code, so the trainer's completion contract
      (training.go:163-238) applies unchanged; a retryable code restarts the process up to
      ``max_restarts`` times (restartPolicy: OnFailure);
    * the processes meet through a Unix-socket rendezvous (k8s_b200/rendezvous.py), not torchrun.
    """

    def __init__(self, n_gpus: int = 1, workload: str = "syncps", workload_args: Optional[List[str]] = None,
                 log_dir: Optional[str] = None, max_restarts: int = 1, extra_env: Optional[Dict[str, str]] = None,
                 workload_cmd: Optional[List[str]] = None):
        from . import rendezvous
        self.n_gpus = n_gpus
        self.workload = workload
        self.workload_args = list(workload_args or [])
        self.workload_cmd = list(workload_cmd) if workload_cmd else None   # replaces the training replicas' program
        self.launches: Dict[str, ReplicaLaunch] = {}
        self.procs: Dict[str, subprocess.Popen] = {}
        self.restarts: Dict[str, int] = {}
        self.max_restarts = max_restarts
        self.extra_env = dict(extra_env or {})
        self.tmp = tempfile.mkdtemp(prefix="tfgraft-")
        self.log_dir = log_dir or self.tmp
        os.makedirs(self.log_dir, exist_ok=True)
        self.server = rendezvous.Server(os.path.join(self.tmp, "rdv.sock"))

    # -- ReplicaBackend -------------------------------------------------------------------
    def create_replica(self, launch: ReplicaLaunch) -> None:
        self.launches.setdefault(launch.name, launch)   # AlreadyExists is tolerated
        if launch.name not in self.procs:
            self._spawn(launch)

    def command(self, launch: ReplicaLaunch) -> List[str]:
        root_pkg = __name__.rsplit(".", 1)[0]
        if launch.command:                               # default PS: keep the reference's argv
            args = [a for a in launch.command[2:]]
            return [sys.executable, "-m", root_pkg + ".grpc_tensorflow_server"] + args + \
                   ["--n_gpus", str(self.n_gpus), "--join"]
        if self.workload_cmd:
            return self.workload_cmd + self.workload_args
        mod = ".tf_sync_ps" if self.workload == "syncps" else ".tf_smoke"
        return [sys.executable, "-m", root_pkg + mod] + self.workload_args

    def _spawn(self, launch: ReplicaLaunch) -> None:
        from . import rendezvous
        env = dict(os.environ)
        env.update(self.extra_env)
        env["TF_CONFIG"] = launch.tf_config
        env["TFGRAFT_N_GPUS"] = str(self.n_gpus)
        env[rendezvous.ENV] = self.server.path
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        env["PYTHONPATH"] = root + os.pathsep + env.get("PYTHONPATH", "")
        log = open(os.path.join(self.log_dir, launch.name + ".log"), "ab")
        self.procs[launch.name] = subprocess.Popen(self.command(launch), env=env, stdout=log, stderr=subprocess.STDOUT,
                                                   cwd=root, start_new_session=True)
        log.close()

    def replica_exit(self, name: str):
        if name not in self.launches:
            return (False, None, "")
        code = self.procs[name].poll()
        if code is None:
            return (True, None, "")
        if code < 0:                                    # killed by a signal: the shell convention 128 + signo
            code = 128 - code
        if is_retryable_exit_code(code) and self.restarts.get(name, 0) < self.max_restarts:
            self.restarts[name] = self.restarts.get(name, 0) + 1
            self._spawn(self.launches[name])
            return (True, None, "")
        return (True, code, "")

    def delete_replica(self, name: str) -> None:
        p = self.procs.pop(name, None)
        self.launches.pop(name, None)
        if p is not None and p.poll() is None:
            try:
                os.killpg(p.pid, signal.SIGTERM)        # the exact process group this backend started
                p.wait(timeout=10)
            except (ProcessLookupError, subprocess.TimeoutExpired):
                try:
                    os.killpg(p.pid, signal.SIGKILL)
                except ProcessLookupError:
                    pass

    def log(self, name: str) -> str:
        try:
            return open(os.path.join(self.log_dir, name + ".log"), errors="replace").read()
        except OSError:
            return ""

    def finish(self, exit_code: int = 0) -> None:
        """The job is over: waiting replicas (hosted workers, default PS) return."""
        self.server.finish(exit_code)

    def close(self) -> None:
        self.finish()
        for name in list(self.procs):
            p = self.procs[name]
            try:
                p.wait(timeout=5)
            except subprocess.TimeoutExpired:
                pass
            self.delete_replica(name)
        self.server.close()
