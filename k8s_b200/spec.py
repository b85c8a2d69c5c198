"""TfJob v1alpha1 schema, defaults, validation and accelerator injection.

Host-side mirror of the reference's ``pkg/spec`` for the hot path (SURVEY.md §8 rows
A1-A4).  Same field names (in their JSON spelling), same defaulting order, same
error strings -- the error strings are API because they surface in
``status.reason`` (pkg/trainer/training.go:293-296).

Reference:
  * types + JSON tags ........ pkg/spec/tf_job.go:33-123
  * Validate ................. pkg/spec/tf_job.go:126-176
  * ConfigureAccelerators .... pkg/spec/tf_job.go:179-233
  * SetDefaults .............. pkg/spec/tf_job.go:236-273
  * default PS template ...... pkg/spec/tf_job.go:282-301
  * status enums ............. pkg/spec/tf_job.go:303-383
  * ControllerConfig ......... pkg/spec/controller.go:3-29

Pod templates stay plain dicts in the Kubernetes JSON spelling (``spec.containers[]
.name`` ...): this package never talks to an apiserver, it only has to read the same
YAML and hand the same TF_CONFIG to each replica.
"""
from __future__ import annotations

import copy
import json
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional

# pkg/spec/tf_job.go:15-27
CRD_KIND = "TfJob"
CRD_KIND_PLURAL = "tfjobs"
CRD_GROUP = "tensorflow.org"
CRD_VERSION = "v1alpha1"
APP_LABEL = "tensorflow-job"
TF_PORT = 2222
REPLICAS = 1

# pkg/spec/tf_job.go:76-88
MASTER = "MASTER"
PS = "PS"
WORKER = "WORKER"
VALID_REPLICA_TYPES = [MASTER, PS, WORKER]
TENSORFLOW = "tensorflow"
DEFAULT_TF_IMAGE = "tensorflow/tensorflow:1.3.0"

# pkg/spec/tf_job.go:303-312
PHASE_NONE = ""
PHASE_CREATING = "Creating"
PHASE_RUNNING = "Running"
PHASE_CLEANUP = "CleanUp"
PHASE_FAILED = "Failed"
PHASE_DONE = "Done"

# pkg/spec/tf_job.go:338-345
STATE_UNKNOWN = "Unknown"
STATE_RUNNING = "Running"
STATE_SUCCEEDED = "Succeeded"
STATE_FAILED = "Failed"

# pkg/spec/tf_job.go:366-374
REPLICA_STATE_UNKNOWN = "Unknown"
REPLICA_STATE_STARTING = "Starting"
REPLICA_STATE_RUNNING = "Running"
REPLICA_STATE_FAILED = "Failed"
REPLICA_STATE_SUCCEEDED = "Succeeded"


class SpecError(ValueError):
    """A Go ``error`` returned by pkg/spec; ``str(e)`` is the reference's message."""


def crd_name() -> str:
    """pkg/spec/tf_job.go:29-31."""
    return "%s.%s" % (CRD_KIND_PLURAL, CRD_GROUP)


def pformat(value: Any) -> str:
    """pkg/util/util.go:13-24 -- ``json.MarshalIndent(value, "", "  ")``."""
    if isinstance(value, str):
        return value
    if hasattr(value, "to_dict"):
        value = value.to_dict()
    return json.dumps(value, indent=2, sort_keys=False)


@dataclass
class ChiefSpec:
    """pkg/spec/tf_job.go:120-123."""
    replicaName: str = ""
    replicaIndex: int = 0

    def to_dict(self):
        return {"replicaName": self.replicaName, "replicaIndex": self.replicaIndex}


@dataclass
class TerminationPolicySpec:
    """pkg/spec/tf_job.go:115-118."""
    chief: Optional[ChiefSpec] = None

    def to_dict(self):
        d = {}
        if self.chief is not None:
            d["chief"] = self.chief.to_dict()
        return d


@dataclass
class TensorBoardSpec:
    """pkg/spec/tf_job.go:107-113.  Parsed and validated (LogDir), never launched:
    the TensorBoard Deployment is out of the hot path."""
    logDir: str = ""
    volumes: Optional[list] = None
    volumeMounts: Optional[list] = None
    serviceType: str = ""

    def to_dict(self):
        return {"logDir": self.logDir, "volumes": self.volumes,
                "volumeMounts": self.volumeMounts, "serviceType": self.serviceType}


@dataclass
class TfReplicaSpec:
    """pkg/spec/tf_job.go:92-105.  ``None`` stands for Go's nil pointer."""
    replicas: Optional[int] = None
    template: Optional[Dict[str, Any]] = None
    tfPort: Optional[int] = None
    tfReplicaType: str = ""
    IsDefaultPS: bool = False

    def to_dict(self):
        d: Dict[str, Any] = {}
        if self.replicas is not None:
            d["replicas"] = self.replicas
        if self.template is not None:
            d["template"] = self.template
        if self.tfPort is not None:
            d["tfPort"] = self.tfPort
        d["tfReplicaType"] = self.tfReplicaType
        d["IsDefaultPS"] = self.IsDefaultPS
        return d

    def containers(self) -> List[Dict[str, Any]]:
        """``r.Template.Spec.Containers``; like the reference this dereferences the
        template unguarded (pkg/spec/tf_job.go:157)."""
        if self.template is None:
            raise SpecError("nil Template dereferenced")  # Go would panic here
        return (self.template.get("spec") or {}).get("containers") or []


@dataclass
class TfJobSpec:
    """pkg/spec/tf_job.go:55-71."""
    RuntimeId: str = ""
    tensorboard: Optional[TensorBoardSpec] = None
    replicaSpecs: List[TfReplicaSpec] = field(default_factory=list)
    tfImage: str = ""
    terminationPolicy: Optional[TerminationPolicySpec] = None

    # -- pkg/spec/tf_job.go:236-273 -------------------------------------------------
    def set_defaults(self) -> None:
        if self.tfImage == "":
            self.tfImage = DEFAULT_TF_IMAGE
        for r in self.replicaSpecs:
            if r.template is None and r.tfReplicaType != PS:
                raise SpecError("ReplicaType: %s, Replica is missing Template; %s"
                                % (r.tfReplicaType, pformat(r)))
            if r.tfPort is None:
                r.tfPort = TF_PORT
            if r.tfReplicaType == "":
                r.tfReplicaType = MASTER
            if r.replicas is None:
                r.replicas = REPLICAS
            if r.template is None and r.tfReplicaType == PS:
                _set_default_ps_pod_template_spec(r, self.tfImage)
        if self.terminationPolicy is None:
            self.terminationPolicy = TerminationPolicySpec(
                chief=ChiefSpec(replicaName="MASTER", replicaIndex=0))

    # -- pkg/spec/tf_job.go:126-176 -------------------------------------------------
    def validate(self) -> None:
        for r in self.replicaSpecs:
            found = False
            if r.template is None and r.tfReplicaType != PS:
                raise SpecError("Replica is missing Template; %s" % pformat(r))
            if r.tfReplicaType == MASTER and r.replicas != 1:
                raise SpecError("The MASTER must have Replicas = 1")
            if r.tfPort is None:
                raise SpecError("tfReplicaSpec.TfPort can't be nil.")
            if r.tfReplicaType not in VALID_REPLICA_TYPES:
                raise SpecError("tfReplicaSpec.TfReplicaType is %s but must be one of %s"
                                % (r.tfReplicaType, _go_slice(VALID_REPLICA_TYPES)))
            for c in r.containers():
                if c.get("name") == TENSORFLOW:
                    found = True
                    break
            if not found:
                raise SpecError("Replica type %s is missing a container named %s"
                                % (r.tfReplicaType, TENSORFLOW))
        if self.terminationPolicy is not None:
            if self.terminationPolicy.chief is None:
                raise SpecError("invalid termination policy, Chief cannot be nil")
            ch = self.terminationPolicy.chief
            if ch.replicaName != "MASTER" or ch.replicaIndex != 0:
                raise SpecError("invalid termination policy, Chief should have "
                                "replicaName=MASTER and index=0")

    # -- pkg/spec/tf_job.go:179-233 -------------------------------------------------
    def configure_accelerators(self, accelerators: Dict[str, "AcceleratorConfig"]) -> None:
        accelerators = accelerators or {}
        for r in self.replicaSpecs:
            if r.template is None:
                raise SpecError("Replica is missing Template; %s" % pformat(r))
            pod_spec = r.template.setdefault("spec", {})
            for c in pod_spec.get("containers") or []:
                if c.get("name") != TENSORFLOW:
                    continue
                # Go iterates a map here (order unspecified); sorted keeps it stable.
                attached: Dict[str, AcceleratorConfig] = {}
                res = c.get("resources") or {}
                for lst in (res.get("limits") or {}, res.get("requests") or {}):
                    for name in lst:
                        if name in accelerators:
                            attached[name] = accelerators[name]
                for name in sorted(attached):
                    cfg = attached[name]
                    for v in cfg.Volumes:
                        pod_spec.setdefault("volumes", [])
                        if pod_spec["volumes"] is None:
                            pod_spec["volumes"] = []
                        pod_spec["volumes"].append(
                            {"name": v.Name, "hostPath": {"path": v.HostPath}})
                        if c.get("volumeMounts") is None:
                            c["volumeMounts"] = []
                        c["volumeMounts"].append({"name": v.Name, "mountPath": v.MountPath})
                    for e in cfg.EnvVars:
                        if c.get("env") is None:
                            c["env"] = []
                        c["env"].append({"name": e.Name, "value": e.Value})
                break

    def to_dict(self):
        d: Dict[str, Any] = {
            "RuntimeId": self.RuntimeId,
            "tensorboard": None if self.tensorboard is None else self.tensorboard.to_dict(),
            "replicaSpecs": [r.to_dict() for r in self.replicaSpecs],
        }
        if self.tfImage:
            d["tfImage"] = self.tfImage
        if self.terminationPolicy is not None:
            d["terminationPolicy"] = self.terminationPolicy.to_dict()
        return d


def _go_slice(items) -> str:
    """``fmt.Sprintf("%v", []T{...})`` -> ``[a b c]``."""
    return "[" + " ".join(str(i) for i in items) + "]"


def _set_default_ps_pod_template_spec(r: TfReplicaSpec, tf_image: str) -> None:
    """pkg/spec/tf_job.go:282-301."""
    r.IsDefaultPS = True
    r.template = {
        "spec": {
            "containers": [{
                "image": tf_image,
                "name": "tensorflow",
                "volumeMounts": [{"name": "ps-config-volume", "mountPath": "/ps-server"}],
            }],
            "restartPolicy": "OnFailure",
        }
    }


@dataclass
class TfReplicaStatus:
    """pkg/spec/tf_job.go:376-383.  ``ReplicasStates`` has no JSON tag in the
    reference, hence the capitalised key (README.md:443-446)."""
    tf_replica_type: str = ""
    state: str = REPLICA_STATE_UNKNOWN
    ReplicasStates: Dict[str, int] = field(default_factory=dict)

    def to_dict(self):
        return {"tf_replica_type": self.tf_replica_type, "state": self.state,
                "ReplicasStates": dict(sorted(self.ReplicasStates.items()))}


@dataclass
class TfJobStatus:
    """pkg/spec/tf_job.go:347-364."""
    phase: str = PHASE_NONE
    reason: str = ""
    controlPaused: bool = False
    conditions: Optional[list] = None
    state: str = ""
    replicaStatuses: Optional[List[TfReplicaStatus]] = None

    def to_dict(self):
        return {
            "phase": self.phase, "reason": self.reason,
            "controlPaused": self.controlPaused, "conditions": self.conditions,
            "state": self.state,
            "replicaStatuses": None if self.replicaStatuses is None
            else [r.to_dict() for r in self.replicaStatuses],
        }

    def copy(self) -> "TfJobStatus":
        return copy.deepcopy(self)

    def is_failed(self) -> bool:
        return self.state == STATE_FAILED


@dataclass
class ObjectMeta:
    name: str = ""
    namespace: str = ""
    uid: str = ""
    extra: Dict[str, Any] = field(default_factory=dict)


@dataclass
class TfJob:
    """pkg/spec/tf_job.go:33-38."""
    apiVersion: str = ""
    kind: str = ""
    metadata: ObjectMeta = field(default_factory=ObjectMeta)
    spec: TfJobSpec = field(default_factory=TfJobSpec)
    status: TfJobStatus = field(default_factory=TfJobStatus)

    def as_owner(self) -> Dict[str, Any]:
        """pkg/spec/tf_job.go:40-52."""
        return {"apiVersion": self.apiVersion, "kind": self.kind,
                "name": self.metadata.name, "uid": self.metadata.uid,
                "controller": True, "blockOwnerDeletion": True}

    def to_dict(self):
        md = dict(self.metadata.extra)
        md["name"] = self.metadata.name
        if self.metadata.namespace:
            md["namespace"] = self.metadata.namespace
        if self.metadata.uid:
            md["uid"] = self.metadata.uid
        return {"apiVersion": self.apiVersion, "kind": self.kind, "metadata": md,
                "spec": self.spec.to_dict(), "status": self.status.to_dict()}


# ---- pkg/spec/controller.go:3-29 ------------------------------------------------------
@dataclass
class AcceleratorVolume:
    Name: str = ""
    HostPath: str = ""
    MountPath: str = ""


@dataclass
class EnvironmentVariableConfig:
    Name: str = ""
    Value: str = ""


@dataclass
class AcceleratorConfig:
    Volumes: List[AcceleratorVolume] = field(default_factory=list)
    EnvVars: List[EnvironmentVariableConfig] = field(default_factory=list)


@dataclass
class ControllerConfig:
    Accelerators: Dict[str, AcceleratorConfig] = field(default_factory=dict)
    GrpcServerFilePath: str = ""

    @staticmethod
    def from_dict(d: Optional[Dict[str, Any]]) -> "ControllerConfig":
        """The operator's ``-controller_config_file`` YAML (cmd/tf_operator/main.go:68-84;
        sample README.md:122-143).  Go's yaml->json unmarshal matches keys
        case-insensitively, so both ``accelerators`` and ``Accelerators`` load."""
        d = d or {}
        low = {k.lower(): v for k, v in d.items()}
        cfg = ControllerConfig(GrpcServerFilePath=low.get("grpcserverfilepath", "") or "")
        for name, a in (low.get("accelerators") or {}).items():
            al = {k.lower(): v for k, v in (a or {}).items()}
            ac = AcceleratorConfig()
            for v in al.get("volumes") or []:
                vl = {k.lower(): x for k, x in v.items()}
                ac.Volumes.append(AcceleratorVolume(vl.get("name", ""), vl.get("hostpath", ""),
                                                    vl.get("mountpath", "")))
            for e in al.get("envvars") or []:
                el = {k.lower(): x for k, x in e.items()}
                ac.EnvVars.append(EnvironmentVariableConfig(el.get("name", ""),
                                                            str(el.get("value", ""))))
            cfg.Accelerators[name] = ac
        return cfg


# ---- document loading (B1 contract) --------------------------------------
def tf_job_from_dict(doc: Dict[str, Any]) -> TfJob:
    """``json.Unmarshal`` into ``TfJob`` (pkg/spec/tf_job_list.go:28-37).

    Only the keys the Go struct tags name are honoured: the README's
    ``replica_specs`` spelling deserialises to an empty replica list there too
    (SURVEY.md §2 row 17)."""
    doc = doc or {}
    md = dict(doc.get("metadata") or {})
    meta = ObjectMeta(name=str(md.pop("name", "") or ""),
                      namespace=str(md.pop("namespace", "") or ""),
                      uid=str(md.pop("uid", "") or ""), extra=md)
    s = doc.get("spec") or {}
    spec = TfJobSpec(RuntimeId=str(s.get("RuntimeId", "") or ""),
                     tfImage=str(s.get("tfImage", "") or ""))
    tb = s.get("tensorboard")
    if tb is not None:
        spec.tensorboard = TensorBoardSpec(
            logDir=str(tb.get("logDir", "") or ""), volumes=tb.get("volumes"),
            volumeMounts=tb.get("volumeMounts"), serviceType=str(tb.get("serviceType", "") or ""))
    for r in s.get("replicaSpecs") or []:
        r = r or {}
        spec.replicaSpecs.append(TfReplicaSpec(
            replicas=None if r.get("replicas") is None else int(r["replicas"]),
            template=copy.deepcopy(r.get("template")),
            tfPort=None if r.get("tfPort") is None else int(r["tfPort"]),
            tfReplicaType=str(r.get("tfReplicaType", "") or ""),
            IsDefaultPS=bool(r.get("IsDefaultPS", False))))
    tp = s.get("terminationPolicy")
    if tp is not None:
        ch = tp.get("chief")
        spec.terminationPolicy = TerminationPolicySpec(
            chief=None if ch is None else ChiefSpec(str(ch.get("replicaName", "")),
                                                    int(ch.get("replicaIndex", 0))))
    st = doc.get("status") or {}
    status = TfJobStatus(phase=st.get("phase", "") or "", reason=st.get("reason", "") or "",
                         controlPaused=bool(st.get("controlPaused", False)),
                         conditions=st.get("conditions"), state=st.get("state", "") or "")
    return TfJob(apiVersion=str(doc.get("apiVersion", "") or ""),
                 kind=str(doc.get("kind", "") or ""), metadata=meta, spec=spec, status=status)


def load_tf_job(path: str) -> TfJob:
    """Read a TfJob YAML/JSON document (what ``kubectl create -f`` would post)."""
    import yaml
    with open(path) as f:
        return tf_job_from_dict(yaml.safe_load(f))


def gpu_limit(r: TfReplicaSpec, names=("nvidia.com/gpu", "alpha.kubernetes.io/nvidia-gpu",
                                      "nvidia-gpu")) -> int:
    """GPU count a replica's ``tensorflow`` container asks for in limits/requests -- the
    signal ConfigureAccelerators keys on (pkg/spec/tf_job.go:189-199), read here by the
    GPU-rank backend."""
    if r.template is None:
        return 0
    n = 0
    for c in (r.template.get("spec") or {}).get("containers") or []:
        if c.get("name") != TENSORFLOW:
            continue
        res = c.get("resources") or {}
        for lst in (res.get("limits") or {}, res.get("requests") or {}):
            for k, v in lst.items():
                if k in names:
                    n = max(n, int(str(v)))
    return n
