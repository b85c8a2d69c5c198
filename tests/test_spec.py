"""pkg/spec/tf_job_test.go transcribed: the Go tables are the golden vectors of the job
surface."""
import copy

import pytest

from k8s_b200 import spec as S


def _ps_spec(resources=None):
    c = {"name": "tensorflow"}
    if resources is not None:
        c["resources"] = resources
    return S.TfJobSpec(replicaSpecs=[S.TfReplicaSpec(
        replicas=2, tfPort=10, template={"spec": {"containers": [c]}}, tfReplicaType=S.PS)])


CUDA_CFG = {"nvidia-gpu": S.AcceleratorConfig(
    Volumes=[S.AcceleratorVolume(Name="cuda-lib", HostPath="/home/cuda", MountPath="/usr/local/cuda")])}


# TestAddAccelertor, pkg/spec/tf_job_test.go:13-233
@pytest.mark.parametrize("resources,expect_mount", [
    ({"requests": {"nvidia-gpu": "1"}}, True),    # case 1: requests are looked at (:22-96)
    ({"limits": {"nvidia-gpu": "1"}}, True),      # case 2: limits are looked at (:98-172)
    (None, False),                                # case 3: no GPUs (:174-222)
])
def test_add_accelerator(resources, expect_mount):
    s = _ps_spec(resources)
    before = copy.deepcopy(s)
    s.configure_accelerators(CUDA_CFG)
    pod = s.replicaSpecs[0].template["spec"]
    if expect_mount:
        assert pod["volumes"] == [{"name": "cuda-lib", "hostPath": {"path": "/home/cuda"}}]
        assert pod["containers"][0]["volumeMounts"] == [{"name": "cuda-lib", "mountPath": "/usr/local/cuda"}]
        assert pod["containers"][0]["resources"] == resources
    else:
        assert s == before


def test_add_accelerator_env_vars_and_missing_template():
    s = _ps_spec({"limits": {"nvidia.com/gpu": 1}})
    cfg = {"nvidia.com/gpu": S.AcceleratorConfig(EnvVars=[S.EnvironmentVariableConfig("LD_LIBRARY_PATH", "/x")])}
    s.configure_accelerators(cfg)
    assert s.replicaSpecs[0].template["spec"]["containers"][0]["env"] == [{"name": "LD_LIBRARY_PATH", "value": "/x"}]
    bad = S.TfJobSpec(replicaSpecs=[S.TfReplicaSpec(tfReplicaType=S.PS)])
    with pytest.raises(S.SpecError, match="Replica is missing Template"):
        bad.configure_accelerators(cfg)


# TestSetDefaults, pkg/spec/tf_job_test.go:235-339
def test_set_defaults_empty_replica():
    s = S.TfJobSpec(replicaSpecs=[S.TfReplicaSpec(template={"spec": {"containers": [{"name": "tensorflow"}]}})],
                    tfImage="tensorflow/tensorflow:1.3.0")
    s.set_defaults()
    r = s.replicaSpecs[0]
    assert (r.replicas, r.tfPort, r.tfReplicaType, r.IsDefaultPS) == (1, 2222, S.MASTER, False)
    assert r.template == {"spec": {"containers": [{"name": "tensorflow"}]}}
    assert s.tfImage == "tensorflow/tensorflow:1.3.0"
    assert s.terminationPolicy == S.TerminationPolicySpec(chief=S.ChiefSpec("MASTER", 0))


def test_set_defaults_default_ps():
    s = S.TfJobSpec(replicaSpecs=[S.TfReplicaSpec(tfReplicaType=S.PS)], tfImage="tensorflow/tensorflow:1.3.0")
    s.set_defaults()
    r = s.replicaSpecs[0]
    assert (r.replicas, r.tfPort, r.tfReplicaType, r.IsDefaultPS) == (1, 2222, S.PS, True)
    assert r.template == {"spec": {
        "containers": [{"image": "tensorflow/tensorflow:1.3.0", "name": "tensorflow",
                        "volumeMounts": [{"name": "ps-config-volume", "mountPath": "/ps-server"}]}],
        "restartPolicy": "OnFailure"}}
    assert s.terminationPolicy == S.TerminationPolicySpec(chief=S.ChiefSpec("MASTER", 0))


def test_set_defaults_image_and_missing_template():
    s = S.TfJobSpec(replicaSpecs=[S.TfReplicaSpec(tfReplicaType=S.WORKER)])
    with pytest.raises(S.SpecError, match="ReplicaType: WORKER, Replica is missing Template"):
        s.set_defaults()
    assert s.tfImage == S.DEFAULT_TF_IMAGE == "tensorflow/tensorflow:1.3.0"


# Validate, pkg/spec/tf_job.go:126-176 (error strings are API)
def _valid():
    return S.TfJobSpec(replicaSpecs=[S.TfReplicaSpec(
        replicas=1, tfPort=2222, tfReplicaType=S.MASTER,
        template={"spec": {"containers": [{"name": "tensorflow"}]}})])


def test_validate_ok():
    _valid().validate()


@pytest.mark.parametrize("mutate,msg", [
    (lambda s: setattr(s.replicaSpecs[0], "replicas", 2), "The MASTER must have Replicas = 1"),
    (lambda s: setattr(s.replicaSpecs[0], "tfPort", None), "tfReplicaSpec.TfPort can't be nil."),
    (lambda s: setattr(s.replicaSpecs[0], "tfReplicaType", "CHIEF"),
     r"tfReplicaSpec.TfReplicaType is CHIEF but must be one of \[MASTER PS WORKER\]"),
    (lambda s: s.replicaSpecs[0].template["spec"]["containers"][0].update(name="tf"),
     "Replica type MASTER is missing a container named tensorflow"),
    (lambda s: setattr(s, "terminationPolicy", S.TerminationPolicySpec(chief=None)),
     "invalid termination policy, Chief cannot be nil"),
    (lambda s: setattr(s, "terminationPolicy", S.TerminationPolicySpec(chief=S.ChiefSpec("WORKER", 0))),
     "invalid termination policy, Chief should have replicaName=MASTER and index=0"),
    (lambda s: setattr(s, "terminationPolicy", S.TerminationPolicySpec(chief=S.ChiefSpec("MASTER", 1))),
     "invalid termination policy, Chief should have replicaName=MASTER and index=0"),
])
def test_validate_errors(mutate, msg):
    s = _valid()
    mutate(s)
    with pytest.raises(S.SpecError, match=msg):
        s.validate()


def test_validate_missing_template_non_ps():
    s = S.TfJobSpec(replicaSpecs=[S.TfReplicaSpec(replicas=1, tfPort=1, tfReplicaType=S.WORKER)])
    with pytest.raises(S.SpecError, match="Replica is missing Template"):
        s.validate()


# the acceptance inputs, examples/*.yaml, unchanged
def test_example_yamls_load_and_default(examples_dir="examples"):
    import os
    root = os.path.join(os.path.dirname(os.path.dirname(__file__)), examples_dir)
    j = S.load_tf_job(os.path.join(root, "tf_job.yaml"))
    assert (j.apiVersion, j.kind, j.metadata.name) == ("tensorflow.org/v1alpha1", "TfJob", "example-job")
    assert [(r.tfReplicaType, r.replicas, r.tfPort) for r in j.spec.replicaSpecs] == \
        [("MASTER", 1, None), ("WORKER", 1, None), ("PS", 2, None)]
    assert j.spec.replicaSpecs[2].template is None
    j.spec.set_defaults()
    j.spec.validate()
    assert [(r.tfReplicaType, r.replicas, r.tfPort, r.IsDefaultPS) for r in j.spec.replicaSpecs] == \
        [("MASTER", 1, 2222, False), ("WORKER", 1, 2222, False), ("PS", 2, 2222, True)]

    g = S.load_tf_job(os.path.join(root, "tf_job_gpu.yaml"))   # omits replicas and tfPort
    g.spec.set_defaults()
    g.spec.validate()
    r = g.spec.replicaSpecs[0]
    assert (r.tfReplicaType, r.replicas, r.tfPort) == ("MASTER", 1, 2222)
    assert S.gpu_limit(r) == 1

    d = S.load_tf_job(os.path.join(root, "tf_job_defaults.yaml"))  # omits the type as well
    d.spec.set_defaults()
    d.spec.validate()
    assert (d.spec.replicaSpecs[0].tfReplicaType, d.spec.replicaSpecs[0].replicas) == ("MASTER", 1)


def test_readme_replica_specs_spelling_is_not_honoured():
    # only the JSON tag `replicaSpecs` deserialises
    j = S.tf_job_from_dict({"spec": {"replica_specs": [{"tfReplicaType": "MASTER"}]}})
    assert j.spec.replicaSpecs == []


def test_controller_config_from_yaml_dict():
    # README.md:122-143 / tf-job-operator-chart/templates/config.yaml:9-22
    cfg = S.ControllerConfig.from_dict({
        "grpcServerFilePath": "/opt/mlkube/grpc_tensorflow_server/grpc_tensorflow_server.py",
        "accelerators": {"alpha.kubernetes.io/nvidia-gpu": {
            "volumes": [{"name": "nvidia-libraries", "mountPath": "/usr/local/nvidia/lib64",
                         "hostPath": "/home/kubernetes/bin/nvidia/lib"}],
            "envVars": [{"name": "LD_LIBRARY_PATH", "value": "/usr/local/nvidia/lib64"}]}}})
    a = cfg.Accelerators["alpha.kubernetes.io/nvidia-gpu"]
    assert a.Volumes[0] == S.AcceleratorVolume("nvidia-libraries", "/home/kubernetes/bin/nvidia/lib",
                                               "/usr/local/nvidia/lib64")
    assert a.EnvVars[0] == S.EnvironmentVariableConfig("LD_LIBRARY_PATH", "/usr/local/nvidia/lib64")
    assert cfg.GrpcServerFilePath.endswith("grpc_tensorflow_server.py")


def test_crd_name_and_status_json_keys():
    assert S.crd_name() == "tfjobs.tensorflow.org"
    st = S.TfJobStatus(phase=S.PHASE_DONE, state=S.STATE_SUCCEEDED, replicaStatuses=[
        S.TfReplicaStatus("MASTER", S.REPLICA_STATE_SUCCEEDED, {"Succeeded": 1})])
    d = st.to_dict()   # README.md:437-455
    assert set(d) == {"phase", "reason", "controlPaused", "conditions", "state", "replicaStatuses"}
    assert d["replicaStatuses"][0] == {"tf_replica_type": "MASTER", "state": "Succeeded",
                                       "ReplicasStates": {"Succeeded": 1}}
