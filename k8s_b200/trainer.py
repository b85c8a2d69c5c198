"""Per-job trainer: cluster spec, TF_CONFIG, replica launch and status.

Host-side mirror of the reference's ``pkg/trainer`` for the hot path (SURVEY.md §8
rows A5-A7, A10).  What changes against the reference is *where a replica runs*:
``TFReplicaSet.create`` hands each (type, index) to a :class:`ReplicaBackend` instead of
creating a Service + batch/v1 Job through a kube clientset.  Everything a replica is
told -- its name, labels, TF_CONFIG, the default-PS command line -- is byte-identical.

Reference:
  * ClusterSpec / TaskSpec / TfConfig .... pkg/trainer/training.go:66-74,114-128; replicas.go:45-53
  * setup / reconcile / GetStatus ........ pkg/trainer/training.go:163-199,245-301,350-409
  * isRetryableTerminationState .......... pkg/trainer/training.go:203-238
  * NewTFReplicaSet / Labels / Create .... pkg/trainer/replicas.go:55-99,124-271
  * transformClusterSpecForDefaultPS ..... pkg/trainer/replicas.go:102-122
  * replicaStatusFromPodList / GetStatus . pkg/trainer/replicas.go:359-492
  * jobName .............................. pkg/trainer/replicas.go:494-500
  * RandString ........................... pkg/util/util.go:27-54
"""
from __future__ import annotations

import copy
import json
import random
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional

from . import spec as S

RECONCILE_INTERVAL_S = 8.0  # pkg/trainer/training.go:23
LETTER_BYTES = "0123456789abcdefghijklmnopqrstuvwxyz"  # pkg/util/util.go:27


def rand_string(n: int, rng: Optional[random.Random] = None) -> str:
    """pkg/util/util.go:38-54: ``n`` characters of ``[0-9a-z]``.

    The reference draws 6-bit indices from a time-seeded source and rejects values
    >= 36; the distribution (uniform over the alphabet) is what is kept, the bit
    stream is not reproducible by design."""
    rng = rng or random
    return "".join(rng.choice(LETTER_BYTES) for _ in range(n))


def go_json(value: Any) -> str:
    """``encoding/json.Marshal``: no spaces, map keys sorted, ``<>&`` escaped.

    Struct field order is the caller's job (pass dicts built in declaration order;
    ``ClusterSpec`` maps are sorted here)."""
    if isinstance(value, ClusterSpec):
        value = dict(sorted(value.items()))
    out = json.dumps(value, separators=(",", ":"), ensure_ascii=False)
    return (out.replace("<", "\\u003c").replace(">", "\\u003e").replace("&", "\\u0026")
            .replace("\u2028", "\\u2028").replace("\u2029", "\\u2029"))


class ClusterSpec(dict):
    """pkg/trainer/training.go:66-69: job name -> list of ``host:port``."""


def tf_config_json(cluster: Dict[str, List[str]], task_type: str, index: int) -> str:
    """pkg/trainer/replicas.go:189-203: ``json.Marshal(TfConfig{...})``.

    Field order cluster, task, environment (struct declaration order, replicas.go:45-53);
    cluster keys sorted (Go marshals maps in key order); a nil/empty cluster map with no
    entries marshals as ``{}``."""
    doc = {"cluster": dict(sorted(cluster.items())),
           "task": {"type": task_type, "index": int(index)},
           "environment": "cloud"}
    return go_json(doc)


def transform_cluster_spec_for_default_ps(cluster: Dict[str, List[str]]) -> str:
    """pkg/trainer/replicas.go:102-122: ``job|h:p;h:p,job|...`` with keys sorted."""
    return ",".join(k + "|" + ";".join(cluster[k]) for k in sorted(cluster))


def is_retryable_termination_state(exit_code: int, reason: str = "") -> bool:
    """pkg/trainer/training.go:203-238."""
    if reason == "OOMKilled":
        return False
    if 0 <= exit_code <= 127:
        return False
    return True


def replica_status_from_pod_list(pods: List[Dict[str, Any]], name: str) -> str:
    """pkg/trainer/replicas.go:359-412 over pods in Kubernetes JSON spelling.

    ``pods``: ``[{"status": {"startTime": <comparable>, "containerStatuses": [{"name",
    "state": {"running"|"waiting"|"terminated": {...}}, "lastState": {...}}]}}]``.
    Newest pod by ``startTime`` wins; a previous termination wins over the current
    state; exit 0 -> Succeeded; retryable -> Running; otherwise Failed."""
    latest = None
    for p in pods:
        if latest is None:
            latest = p
            continue
        lt = (latest.get("status") or {}).get("startTime")
        pt = (p.get("status") or {}).get("startTime")
        if lt is not None and pt is not None and lt < pt:
            latest = p
    if latest is None:
        return S.REPLICA_STATE_RUNNING
    tf_state: Dict[str, Any] = {}
    for cs in (latest.get("status") or {}).get("containerStatuses") or []:
        if cs.get("name") != name:
            continue
        tf_state = cs.get("state") or {}
        last = cs.get("lastState") or cs.get("lastTerminationState") or {}
        if last.get("terminated") is not None:
            tf_state = last
    if tf_state.get("running") is not None or tf_state.get("waiting") is not None:
        return S.REPLICA_STATE_RUNNING
    term = tf_state.get("terminated")
    if term is not None:
        code = int(term.get("exitCode", 0))
        if code == 0:
            return S.REPLICA_STATE_SUCCEEDED
        if is_retryable_termination_state(code, term.get("reason", "") or ""):
            return S.REPLICA_STATE_RUNNING
        return S.REPLICA_STATE_FAILED
    return S.REPLICA_STATE_UNKNOWN


def replica_state_from_exit(exit_code: Optional[int], reason: str = "") -> str:
    """The same classification for a replica that is a local GPU rank rather than a pod:
    ``None`` = still running (SURVEY.md §8b contract B3)."""
    if exit_code is None:
        return S.REPLICA_STATE_RUNNING
    pod = {"status": {"containerStatuses": [{
        "name": S.TENSORFLOW,
        "state": {"terminated": {"exitCode": int(exit_code), "reason": reason}}}]}}
    return replica_status_from_pod_list([pod], S.TENSORFLOW)


# ---------------------------------------------------------------------------------------
# Backends: the seam that replaces ``kubernetes.Interface`` (SURVEY.md §8b "Go seam").
# ---------------------------------------------------------------------------------------
@dataclass
class ReplicaLaunch:
    """Everything ``TFReplicaSet.Create`` decides for one task index."""
    name: str
    labels: Dict[str, str]
    replica_type: str
    index: int
    port: int
    tf_config: str
    command: Optional[List[str]]
    template: Dict[str, Any]
    owner: Dict[str, Any]


class ReplicaBackend:
    """Where replicas run.  Must be idempotent: reconcile re-creates every 8 s
    (pkg/trainer/training.go:363-367) and ``AlreadyExists`` is tolerated
    (replicas.go:180-186,261-268)."""

    def create_replica(self, launch: ReplicaLaunch) -> None:
        raise NotImplementedError

    def replica_exit(self, name: str):
        """-> (known: bool, exit_code: Optional[int], reason: str)."""
        raise NotImplementedError

    def delete_replica(self, name: str) -> None:
        raise NotImplementedError


class RecordingBackend(ReplicaBackend):
    """In-memory stand-in for ``kubernetes/fake.NewSimpleClientset()``:
    records what would have been created so the trainer logic is testable on CPU."""

    def __init__(self):
        self.services: Dict[str, Dict[str, Any]] = {}
        self.jobs: Dict[str, Dict[str, Any]] = {}
        self.exits: Dict[str, Any] = {}

    def create_replica(self, launch: ReplicaLaunch) -> None:
        if launch.name not in self.services:
            self.services[launch.name] = {
                "metadata": {"name": launch.name, "labels": dict(launch.labels),
                             "ownerReferences": [launch.owner]},
                "spec": {"selector": dict(launch.labels),
                         "ports": [{"name": "tf-port", "port": launch.port}]}}
        if launch.name not in self.jobs:
            tmpl = copy.deepcopy(launch.template)
            tmpl.setdefault("metadata", {}).setdefault("labels", {}).update(launch.labels)
            for c in tmpl.get("spec", {}).get("containers", []):
                if c.get("name") != S.TENSORFLOW:
                    continue
                if launch.command is not None and launch.replica_type == S.PS:
                    pass  # command already placed on containers[0] by the caller
                c.setdefault("env", [])
                if c["env"] is None:
                    c["env"] = []
                c["env"].append({"name": "TF_CONFIG", "value": launch.tf_config})
            self.jobs[launch.name] = {
                "metadata": {"name": launch.name, "labels": dict(launch.labels),
                             "ownerReferences": [launch.owner]},
                "spec": {"completions": 1, "parallelism": 1, "template": tmpl}}

    def replica_exit(self, name: str):
        if name not in self.jobs:
            return (False, None, "")
        e = self.exits.get(name)
        if e is None:
            return (True, None, "")
        if isinstance(e, tuple):
            return (True, e[0], e[1])
        return (True, int(e), "")

    def delete_replica(self, name: str) -> None:
        self.services.pop(name, None)
        self.jobs.pop(name, None)


# ---------------------------------------------------------------------------------------
class TFReplicaSet:
    """pkg/trainer/replicas.go:28-34."""

    def __init__(self, backend: ReplicaBackend, tf_replica_spec: S.TfReplicaSpec,
                 job: "TrainingJob"):
        # pkg/trainer/replicas.go:55-88
        if tf_replica_spec.tfReplicaType == S.MASTER and tf_replica_spec.replicas != 1:
            raise S.SpecError("The MASTER must have Replicas = 1")
        if tf_replica_spec.tfPort is None:
            raise S.SpecError("tfReplicaSpec.TfPort can't be nil.")
        if tf_replica_spec.template is None and tf_replica_spec.tfReplicaType != S.PS:
            raise S.SpecError("tfReplicaSpec.Template can't be nil for replica type %s."
                              % tf_replica_spec.tfReplicaType)
        if tf_replica_spec.tfReplicaType not in S.VALID_REPLICA_TYPES:
            raise S.SpecError("tfReplicaSpec.TfReplicaType is %s but must be one of %s"
                              % (tf_replica_spec.tfReplicaType,
                                 S._go_slice(S.VALID_REPLICA_TYPES)))
        self.backend = backend
        self.job = job
        self.spec = tf_replica_spec

    def labels(self) -> Dict[str, str]:
        """pkg/trainer/replicas.go:91-99."""
        return {"tensorflow.org": "",
                "job_type": self.spec.tfReplicaType,
                "runtime_id": self.job.job.spec.RuntimeId,
                "tf_job_name": self.job.job.metadata.name}

    def job_name(self, index: int) -> str:
        """pkg/trainer/replicas.go:494-500: ``%.40s-type-runtimeid-index``."""
        return "%s-%s-%s-%d" % (self.job.job.metadata.name[:40],
                                self.spec.tfReplicaType.lower(),
                                self.job.job.spec.RuntimeId, index)

    def default_ps_config_map_name(self) -> str:
        """pkg/trainer/replicas.go:502-504."""
        return "cm-ps-%s" % self.job.job.spec.RuntimeId

    def create(self, config: S.ControllerConfig) -> None:
        """pkg/trainer/replicas.go:124-271 with the clientset replaced by the backend."""
        if self.spec.IsDefaultPS:
            # :140-149 -- the ConfigMap volume carrying grpc_tensorflow_server.py; here the
            # default PS is k8s_b200.grpc_tensorflow_server, so only the volume entry is kept.
            vols = self.spec.template["spec"].setdefault("volumes", [])
            ent = {"name": "ps-config-volume",
                   "configMap": {"name": self.default_ps_config_map_name()}}
            if ent not in vols:  # idempotent, unlike the reference
                vols.append(ent)
        for index in range(int(self.spec.replicas)):
            task_labels = self.labels()
            task_labels["task_index"] = "%d" % index
            cluster = self.job.cluster_spec()
            cfg = tf_config_json(cluster, self.spec.tfReplicaType.lower(), index)
            command = None
            if self.spec.IsDefaultPS:
                cs = transform_cluster_spec_for_default_ps(cluster)
                command = ["python", "/ps-server/grpc_tensorflow_server.py", "--cluster_spec",
                           cs, "--job_name", "ps", "--task_id", "%d" % index]
                self.spec.template["spec"]["containers"][0]["command"] = command
            self.backend.create_replica(ReplicaLaunch(
                name=self.job_name(index), labels=task_labels,
                replica_type=self.spec.tfReplicaType, index=index, port=int(self.spec.tfPort),
                tf_config=cfg, command=command, template=copy.deepcopy(self.spec.template),
                owner=self.job.job.as_owner()))

    def delete(self) -> None:
        """pkg/trainer/replicas.go:299-356."""
        for index in range(int(self.spec.replicas)):
            self.backend.delete_replica(self.job_name(index))

    def get_status(self) -> S.TfReplicaStatus:
        """pkg/trainer/replicas.go:415-492."""
        status = S.TfReplicaStatus(tf_replica_type=self.spec.tfReplicaType,
                                   state=S.REPLICA_STATE_UNKNOWN, ReplicasStates={})

        def increment(state: str) -> None:
            status.ReplicasStates[state] = status.ReplicasStates.get(state, 0) + 1

        for index in range(int(self.spec.replicas)):
            known, code, reason = self.backend.replica_exit(self.job_name(index))
            if not known:
                increment(S.REPLICA_STATE_UNKNOWN)
                continue
            increment(replica_state_from_exit(code, reason))
        if S.REPLICA_STATE_FAILED in status.ReplicasStates:
            status.state = S.REPLICA_STATE_FAILED
            return status
        if S.REPLICA_STATE_RUNNING in status.ReplicasStates:
            status.state = S.REPLICA_STATE_RUNNING
            return status
        if status.ReplicasStates.get(S.REPLICA_STATE_SUCCEEDED, 0) == int(self.spec.replicas):
            status.state = S.REPLICA_STATE_SUCCEEDED
        return status


class TrainingJob:
    """pkg/trainer/training.go:41-64 without the goroutine: ``reconcile`` is called by the
    runner (once per poll) on one host thread, as the reference calls it serially."""

    def __init__(self, backend: ReplicaBackend, job: S.TfJob):
        # initJob, pkg/trainer/training.go:76-89
        self.backend = backend
        self.job = job
        self.replicas: List[TFReplicaSet] = []
        self.tensorboard = None
        self.status = job.status.copy()

    def cluster_spec(self) -> ClusterSpec:
        """pkg/trainer/training.go:114-128."""
        cs = ClusterSpec()
        for p in self.replicas:
            cs[p.spec.tfReplicaType.lower()] = [
                "%s:%d" % (p.job_name(i), int(p.spec.tfPort)) for i in range(int(p.spec.replicas))]
        return cs

    def setup(self, config: S.ControllerConfig, rng: Optional[random.Random] = None) -> None:
        """pkg/trainer/training.go:245-301: defaults -> validate -> replica sets ->
        tensorboard -> accelerators -> RuntimeId; any error lands in status.reason."""
        try:
            if self.status.phase != S.PHASE_NONE:
                return
            try:
                self.job.spec.set_defaults()
            except S.SpecError as e:
                raise S.SpecError("there was a problem setting defaults for job spec: %s" % e)
            try:
                self.job.spec.validate()
            except S.SpecError as e:
                raise S.SpecError("invalid job spec: %s" % e)
            for t in self.job.spec.replicaSpecs:
                self.replicas.append(TFReplicaSet(self.backend, t, self))
            tb = self.job.spec.tensorboard
            if tb is not None:
                # NewTBReplicaSet, pkg/trainer/tensorboard.go:29-32; TensorBoard itself is
                # accepted and ignored (out of the hot path).
                if tb.logDir == "":
                    raise S.SpecError("tbReplicaSpec.LogDir must be specified")
                self.tensorboard = tb
            try:
                self.job.spec.configure_accelerators(config.Accelerators)
            except S.SpecError as e:
                raise S.SpecError("ConfigureAccelerators(...) error; %s" % e)
            if self.job.spec.RuntimeId == "":
                self.job.spec.RuntimeId = rand_string(4, rng)
        except S.SpecError as e:
            self.status.reason = str(e)
            self.status.phase = S.PHASE_FAILED
            self.status.state = S.STATE_FAILED
        else:
            self.status.phase = S.PHASE_CREATING
            self.status.state = S.STATE_RUNNING

    def create_resources(self, config: S.ControllerConfig) -> None:
        """pkg/trainer/training.go:131-145."""
        for r in self.replicas:
            r.create(config)

    def delete_resources(self) -> None:
        """pkg/trainer/training.go:148-161."""
        for r in self.replicas:
            r.delete()

    def get_status(self):
        """pkg/trainer/training.go:163-199 -> (state, [TfReplicaStatus])."""
        state = S.STATE_UNKNOWN
        statuses: List[S.TfReplicaStatus] = []
        set_states: Dict[str, str] = {}
        for r in self.replicas:
            rs = r.get_status()
            set_states[r.spec.tfReplicaType] = rs.state
            statuses.append(rs)
            if rs.state == S.REPLICA_STATE_FAILED:
                state = S.STATE_FAILED
        if set_states.get(S.MASTER) == S.REPLICA_STATE_SUCCEEDED:
            return S.STATE_SUCCEEDED, statuses
        if set_states.get(S.MASTER) == S.REPLICA_STATE_FAILED:
            return S.STATE_FAILED, statuses
        return S.STATE_RUNNING, statuses

    def update_tpr_status(self) -> None:
        """pkg/trainer/training.go:331-347: persist ``status`` onto the job object (the
        CRD PUT of the reference becomes an in-memory assignment)."""
        self.job.status = self.status.copy()

    def reconcile(self, config: S.ControllerConfig) -> None:
        """pkg/trainer/training.go:350-409."""
        if self.status.phase == S.PHASE_NONE:
            self.setup(config)
            self.update_tpr_status()
        if self.job.status.phase in (S.PHASE_CREATING, S.PHASE_RUNNING):
            self.create_resources(config)
            state, statuses = self.get_status()
            self.status.replicaStatuses = statuses
            if state == S.STATE_FAILED:
                self.status.phase = S.PHASE_DONE
                self.status.state = S.STATE_FAILED
            elif state == S.STATE_SUCCEEDED:
                self.status.phase = S.PHASE_DONE
                self.status.state = S.STATE_SUCCEEDED
        self.update_tpr_status()
        if self.job.status.phase == S.PHASE_CLEANUP:
            self.delete_resources()

    def delete(self) -> None:
        """The ``eventDeleteJob`` branch of run(), pkg/trainer/training.go:431-450."""
        if self.job.status.phase != S.PHASE_CLEANUP:
            self.status.phase = S.PHASE_CLEANUP
        self.delete_resources()
        self.update_tpr_status()
