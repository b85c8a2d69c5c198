"""pkg/trainer/*_test.go transcribed (golden strings and tables), plus the
e2e acceptance semantics of test/e2e/main.go:129-151 against the GPU-rank backend."""
import json
import os
import random

import pytest

from k8s_b200 import spec as S
from k8s_b200 import trainer as T
from k8s_b200.backend import GpuRankBackend
from k8s_b200.cluster import RankMap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def tmpl(resources=None):
    c = {"name": "tensorflow"}
    if resources:
        c["resources"] = resources
    return {"spec": {"containers": [c]}}


# TestIsRetryableTerminationState, pkg/trainer/training_test.go:17-73
@pytest.mark.parametrize("code,reason,expected", [
    (0, "", False), (0, "", False), (1, "", False), (1, "", False), (244, "", True),
    (244, "OOMKilled", False),
    (127, "", False), (128, "", True), (137, "", True), (143, "", True), (-1, "", True),
])
def test_is_retryable_termination_state(code, reason, expected):
    assert T.is_retryable_termination_state(code, reason) is expected


# TestClusterSpec, pkg/trainer/training_test.go:75-172
def test_cluster_spec_golden():
    job = S.TfJob(metadata=S.ObjectMeta(name="myjob"), spec=S.TfJobSpec(RuntimeId="runtime", replicaSpecs=[
        S.TfReplicaSpec(replicas=2, tfPort=22, template=tmpl(), tfReplicaType=S.PS),
        S.TfReplicaSpec(replicas=1, tfPort=42, template=tmpl(), tfReplicaType=S.MASTER),
        S.TfReplicaSpec(replicas=3, tfPort=40, template=tmpl(), tfReplicaType=S.WORKER)]))
    tj = T.TrainingJob(T.RecordingBackend(), job)
    tj.setup(S.ControllerConfig())
    assert dict(tj.cluster_spec()) == {
        "ps": ["myjob-ps-runtime-0:22", "myjob-ps-runtime-1:22"],
        "master": ["myjob-master-runtime-0:42"],
        "worker": ["myjob-worker-runtime-0:40", "myjob-worker-runtime-1:40", "myjob-worker-runtime-2:40"]}


# TestJobSetup, pkg/trainer/training_test.go:174-327
CFG = S.ControllerConfig(Accelerators={"nvidia-gpu": S.AcceleratorConfig(
    Volumes=[S.AcceleratorVolume("cuda-lib", "/home/cuda", "/usr/local/cuda")])})


@pytest.mark.parametrize("resources,tensorboard,mounts,phase,state,reason", [
    (None, None, 0, S.PHASE_CREATING, S.STATE_RUNNING, ""),
    ({"requests": {"nvidia-gpu": "1"}}, None, 1, S.PHASE_CREATING, S.STATE_RUNNING, ""),
    ({"requests": {"nvidia-gpu": "1"}}, S.TensorBoardSpec(), 0, S.PHASE_FAILED, S.STATE_FAILED,
     "tbReplicaSpec.LogDir must be specified"),
])
def test_job_setup(resources, tensorboard, mounts, phase, state, reason):
    job = S.TfJob(spec=S.TfJobSpec(tensorboard=tensorboard, replicaSpecs=[
        S.TfReplicaSpec(replicas=2, tfPort=10, template=tmpl(resources), tfReplicaType=S.PS)]))
    tj = T.TrainingJob(T.RecordingBackend(), job)
    tj.setup(CFG)
    assert (tj.status.phase, tj.status.state, tj.status.reason) == (phase, state, reason)
    if state != S.STATE_FAILED:
        rid = job.spec.RuntimeId
        assert len(rid) == 4 and all(ch in T.LETTER_BYTES for ch in rid)
    pod = job.spec.replicaSpecs[0].template["spec"]
    assert len(pod.get("volumes") or []) == mounts
    assert len(pod["containers"][0].get("volumeMounts") or []) == mounts


def test_setup_wraps_validation_error_in_reason():
    job = S.TfJob(spec=S.TfJobSpec(replicaSpecs=[
        S.TfReplicaSpec(replicas=2, tfPort=1, template=tmpl(), tfReplicaType=S.MASTER)]))
    tj = T.TrainingJob(T.RecordingBackend(), job)
    tj.setup(S.ControllerConfig())
    assert tj.status.reason == "invalid job spec: The MASTER must have Replicas = 1"
    assert tj.status.phase == S.PHASE_FAILED


# TestTFReplicaSet, pkg/trainer/replicas_test.go:22-182
def test_tf_replica_set_create():
    be = T.RecordingBackend()
    job = S.TfJob(metadata=S.ObjectMeta(name="some-job", uid="some-uid"),
                  spec=S.TfJobSpec(RuntimeId="some-runtime", replicaSpecs=[
                      S.TfReplicaSpec(replicas=2, tfPort=10, template=tmpl(), tfReplicaType=S.PS)]))
    tj = T.TrainingJob(be, job)
    rs = T.TFReplicaSet(be, job.spec.replicaSpecs[0], tj)
    rs.create(S.ControllerConfig())
    rs.create(S.ControllerConfig())          # reconcile calls Create again: AlreadyExists tolerated
    owner = {"apiVersion": "", "kind": "", "name": "some-job", "uid": "some-uid",
             "controller": True, "blockOwnerDeletion": True}
    assert len(be.services) == 2 and len(be.jobs) == 2
    for index in range(2):
        labels = {"tensorflow.org": "", "task_index": str(index), "job_type": "PS",
                  "runtime_id": "some-runtime", "tf_job_name": "some-job"}
        name = "some-job-ps-some-runtime-%d" % index
        s, j = be.services[name], be.jobs[name]
        assert s["metadata"]["labels"] == labels and j["metadata"]["labels"] == labels
        assert s["metadata"]["ownerReferences"] == [owner] and j["metadata"]["ownerReferences"] == [owner]
        assert s["spec"]["ports"] == [{"name": "tf-port", "port": 10}]
        cs = j["spec"]["template"]["spec"]["containers"]
        assert len(cs) == 1 and len(cs[0]["env"]) == 1 and cs[0]["env"][0]["name"] == "TF_CONFIG"
        # the replica set was never added to job.Replicas, hence the empty cluster (:161-168)
        assert json.loads(cs[0]["env"][0]["value"]) == {
            "cluster": {}, "task": {"type": "ps", "index": index}, "environment": "cloud"}
        assert cs[0]["env"][0]["value"] == \
            '{"cluster":{},"task":{"type":"ps","index":%d},"environment":"cloud"}' % index
        assert j["spec"]["template"]["metadata"]["labels"] == labels
    rs.delete()
    assert not be.jobs and not be.services


# TestTFReplicaSetStatusFromPodList, pkg/trainer/replicas_test.go:184-340
def _pod(statuses, start=None):
    return {"status": {"containerStatuses": statuses, "startTime": start}}


POD_CASES = [
    ([_pod([{"name": "master", "state": {"running": {}}}])], S.REPLICA_STATE_RUNNING),
    ([_pod([{"name": "master", "state": {"terminated": {"exitCode": 0}}}])], S.REPLICA_STATE_SUCCEEDED),
    ([_pod([{"name": "other", "state": {"running": {}}},
            {"name": "master", "state": {"terminated": {"exitCode": 0}}}])], S.REPLICA_STATE_SUCCEEDED),
    ([_pod([{"name": "master", "state": {"running": {}},
             "lastState": {"terminated": {"exitCode": 100, "message": "some reason"}}}])],
     S.REPLICA_STATE_FAILED),
    ([_pod([{"name": "master", "state": {"running": {}}}], start="2016-11-30T00:00:00Z"),
      _pod([{"name": "master", "state": {"terminated": {"exitCode": 100, "message": "some reason"}}}],
           start="2017-11-30T00:00:00Z")], S.REPLICA_STATE_FAILED),
]


@pytest.mark.parametrize("pods,expected", POD_CASES)
def test_replica_status_from_pod_list(pods, expected):
    assert T.replica_status_from_pod_list(pods, "master") == expected


def test_replica_status_extra_cases():
    assert T.replica_status_from_pod_list([], "tensorflow") == S.REPLICA_STATE_RUNNING      # :372-374
    assert T.replica_status_from_pod_list([_pod([])], "tensorflow") == S.REPLICA_STATE_UNKNOWN  # :411
    p = [_pod([{"name": "tensorflow", "state": {"terminated": {"exitCode": 137}}}])]
    assert T.replica_status_from_pod_list(p, "tensorflow") == S.REPLICA_STATE_RUNNING       # retryable
    p = [_pod([{"name": "tensorflow", "state": {"terminated": {"exitCode": 137, "reason": "OOMKilled"}}}])]
    assert T.replica_status_from_pod_list(p, "tensorflow") == S.REPLICA_STATE_FAILED


# TestTransformClusterSpecForDefaultPS, pkg/trainer/replicas_test.go:342-356
def test_transform_cluster_spec_for_default_ps():
    cs = {"master": ["master-0:2222"], "worker": ["worker-0:2222", "worker-1:2222"],
          "ps": ["localhost:2222", "ps-1:2222"]}
    assert T.transform_cluster_spec_for_default_ps(cs) == \
        "master|master-0:2222,ps|localhost:2222;ps-1:2222,worker|worker-0:2222;worker-1:2222"


def test_tf_config_json_is_go_marshal_compatible():
    cs = {"worker": ["w-0:1"], "master": ["m<&>:2"], "ps": []}
    s = T.tf_config_json(cs, "worker", 3)
    assert s == ('{"cluster":{"master":["m\\u003c\\u0026\\u003e:2"],"ps":[],"worker":["w-0:1"]},'
                 '"task":{"type":"worker","index":3},"environment":"cloud"}')


def test_job_name_truncates_to_40_chars():
    long = "x" * 50
    job = S.TfJob(metadata=S.ObjectMeta(name=long), spec=S.TfJobSpec(RuntimeId="abcd", replicaSpecs=[
        S.TfReplicaSpec(replicas=1, tfPort=1, template=tmpl(), tfReplicaType=S.WORKER)]))
    rs = T.TFReplicaSet(T.RecordingBackend(), job.spec.replicaSpecs[0], T.TrainingJob(T.RecordingBackend(), job))
    assert rs.job_name(7) == "x" * 40 + "-worker-abcd-7"
    assert rs.default_ps_config_map_name() == "cm-ps-abcd"


def test_rand_string_alphabet():
    r = random.Random(1)
    for _ in range(50):
        s = T.rand_string(4, r)
        assert len(s) == 4 and set(s) <= set("0123456789abcdefghijklmnopqrstuvwxyz")


# reconcile + status: training.go:163-199,350-409 with exit codes fed by the fake backend
def _example(name):
    return S.load_tf_job(os.path.join(ROOT, "examples", name))


def test_reconcile_state_machine_with_recording_backend():
    be = T.RecordingBackend()
    tj = T.TrainingJob(be, _example("tf_job.yaml"))
    tj.reconcile(S.ControllerConfig())
    assert (tj.job.status.phase, tj.job.status.state) == (S.PHASE_CREATING, S.STATE_RUNNING)
    rid = tj.job.spec.RuntimeId
    tj.reconcile(S.ControllerConfig())                       # creates resources, all Running
    assert sorted(be.jobs) == sorted(["example-job-master-%s-0" % rid, "example-job-worker-%s-0" % rid,
                                      "example-job-ps-%s-0" % rid, "example-job-ps-%s-1" % rid])
    ps_cmd = be.jobs["example-job-ps-%s-1" % rid]["spec"]["template"]["spec"]["containers"][0]["command"]
    assert ps_cmd[:4] == ["python", "/ps-server/grpc_tensorflow_server.py", "--cluster_spec",
                          "master|example-job-master-%s-0:2222,ps|example-job-ps-%s-0:2222;"
                          "example-job-ps-%s-1:2222,worker|example-job-worker-%s-0:2222" % ((rid,) * 4)]
    assert ps_cmd[4:] == ["--job_name", "ps", "--task_id", "1"]
    assert tj.job.status.state == S.STATE_RUNNING
    st = {r.tf_replica_type: r for r in tj.job.status.replicaStatuses}
    assert st["PS"].ReplicasStates == {"Running": 2}
    be.exits["example-job-worker-%s-0" % rid] = 137           # retryable: stays Running
    tj.reconcile(S.ControllerConfig())
    assert tj.job.status.state == S.STATE_RUNNING
    be.exits["example-job-master-%s-0" % rid] = 0             # MASTER exit 0 decides the job
    tj.reconcile(S.ControllerConfig())
    assert (tj.job.status.phase, tj.job.status.state) == (S.PHASE_DONE, S.STATE_SUCCEEDED)
    tj.delete()
    assert tj.job.status.phase == S.PHASE_CLEANUP and not be.jobs


def test_master_failure_and_worker_failure():
    # training.go:181-198: a failed WORKER marks state Failed, but the unconditional
    # `state = StateRunning` at :197 overrides it unless the MASTER has decided
    for who, code, want_phase, want in [("master", 1, S.PHASE_DONE, S.STATE_FAILED),
                                        ("worker", 2, S.PHASE_CREATING, S.STATE_RUNNING)]:
        be = T.RecordingBackend()
        tj = T.TrainingJob(be, _example("tf_job.yaml"))
        tj.reconcile(S.ControllerConfig())
        tj.reconcile(S.ControllerConfig())
        be.exits["example-job-%s-%s-0" % (who, tj.job.spec.RuntimeId)] = code
        tj.reconcile(S.ControllerConfig())
        assert (tj.job.status.phase, tj.job.status.state) == (want_phase, want)


def test_gpu_rank_backend_maps_replicas_and_classifies_exit_codes():
    from k8s_b200.lib import TfgError
    seen = {}

    def workload(n_gpus, gpu_rank):
        cfg = json.loads(os.environ["TF_CONFIG"])
        seen["cfg"], seen["rank"], seen["n"] = cfg, gpu_rank, n_gpus
        return "ok"

    be = GpuRankBackend(n_gpus=2, workload=workload)
    tj = T.TrainingJob(be, _example("tf_job.yaml"))
    for _ in range(3):
        tj.reconcile(S.ControllerConfig())
        be.run()
    tj.reconcile(S.ControllerConfig())
    assert (tj.job.status.phase, tj.job.status.state) == (S.PHASE_DONE, S.STATE_SUCCEEDED)
    assert seen["cfg"]["task"] == {"type": "master", "index": 0} and seen["rank"] == 0
    rm = RankMap(seen["cfg"]["cluster"], 2)
    assert [rm.gpu_of(j, i) for j, i in rm.tasks()] == [0, 0, 1, 1]   # master0, ps0, ps1, worker0

    def retryable(n_gpus, gpu_rank):
        raise TfgError(130, "CUDA error")

    be = GpuRankBackend(n_gpus=1, workload=retryable)
    tj = T.TrainingJob(be, _example("tf_job_gpu.yaml"))
    for _ in range(3):
        tj.reconcile(S.ControllerConfig())
        be.run()
    assert tj.job.status.state == S.STATE_RUNNING      # >=128 is retried, never Failed

    def permanent(n_gpus, gpu_rank):
        raise TfgError(3, "bad TF_CONFIG")

    be = GpuRankBackend(n_gpus=1, workload=permanent)
    tj = T.TrainingJob(be, _example("tf_job_gpu.yaml"))
    for _ in range(3):
        tj.reconcile(S.ControllerConfig())
        be.run()
    assert (tj.job.status.phase, tj.job.status.state) == (S.PHASE_DONE, S.STATE_FAILED)
