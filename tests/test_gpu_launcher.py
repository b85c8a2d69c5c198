"""North-star target: the UNCHANGED examples/tf_job.yaml (1 MASTER + 1 WORKER + 2
default PS) drives the synchronous PS training step -- YAML -> trainer reconcile -> one process per
task index with the trainer's TF_CONFIG (pkg/trainer/replicas.go:189-203,241-255) -> GPU ranks ->
exit codes -> `phase: Done, state: Succeeded` (pkg/trainer/training.go:163-238) -- and the loss and
parameters after three steps match the oracle's sync_step at the fp32 bounds.  On one GPU both
replicas and both PS shards share GPU rank 0; on two GPUs MASTER and WORKER are separate rank-host
processes that map each other's buffers through CUDA IPC and meet at the Unix-socket rendezvous."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from k8s_b200 import spec as S
from k8s_b200.run import run_tf_job_processes
from k8s_b200.syncps import MNIST_VARS, mnist_init, synthetic_batch
from oracle import model_ref as M

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TF_JOB = os.path.join(ROOT, "examples", "tf_job.yaml")
BATCH, STEPS = 16, 3


def oracle_run():
    params = mnist_init(0)
    init = [p.copy() for p in params]
    accum = [np.zeros_like(p) for p in params]
    losses = []
    for step in range(STEPS):
        batches = [synthetic_batch(r, BATCH, step) for r in range(2)]
        losses.append(M.sync_step(params, accum, batches, "momentum", 0.001, 0.9)[0])
    return losses, init, params


def check_rank_file(path, losses_ref, init, params):
    got = np.load(path)
    for s in range(STEPS):
        assert abs(got["losses"][s] - losses_ref[s]) <= 1e-5 * (s + 1) * abs(losses_ref[s]) + 1e-6, (s, got["losses"], losses_ref)
    for v, name in enumerate(MNIST_VARS):
        upd = got["var%d" % v] - init[v].ravel()
        ref = params[v].ravel() - init[v].ravel()
        e = np.linalg.norm(upd - ref) / np.linalg.norm(ref)
        assert e <= 1e-3 * STEPS, "update of %s: rel L2 %.3e" % (name, e)
    assert int(got["global_step"]) == STEPS
    assert list(got["ps_task"]) == [0, 1, 0, 1, 0, 1, 0, 1]      # replica_device_setter round-robin over the 2 PS tasks


@pytest.mark.parametrize("n_gpus", [1, 2])
@pytest.mark.parametrize("layout", ["logical", "balanced"])
def test_unchanged_tf_job_yaml_drives_the_sync_ps_step(lib, tmp_path, n_gpus, layout):
    if lib.device_count_or_zero() < n_gpus:
        pytest.skip("needs %d GPUs" % n_gpus)
    res = str(tmp_path / "res")
    args = ["--steps", str(STEPS), "--batch", str(BATCH), "--math", "tf32x3", "--layout", layout, "--result_dir", res]
    tj, backend = run_tf_job_processes(TF_JOB, n_gpus=n_gpus, workload="syncps", workload_args=args, timeout=240,
                                       log_dir=str(tmp_path / "logs"))
    try:
        logs = "\n".join("---- %s ----\n%s" % (n, backend.log(n)[-1500:]) for n in backend.launches)
        st = tj.job.to_dict()["status"]
        assert (st["phase"], st["state"]) == ("Done", "Succeeded"), logs
        by_type = {r["tf_replica_type"]: r for r in st["replicaStatuses"]}
        assert by_type["MASTER"]["ReplicasStates"] == {"Succeeded": 1}
        assert len(backend.procs) == 4                            # one process per task index
    finally:
        tj.delete()
        backend.close()
    losses_ref, init, params = oracle_run()
    for r in range(n_gpus):                                       # every rank host holds the full, identical parameters
        check_rank_file(os.path.join(res, "rank%d.npz" % r), losses_ref, init, params)
    if n_gpus == 2:
        a, b = np.load(os.path.join(res, "rank0.npz")), np.load(os.path.join(res, "rank1.npz"))
        for v in range(8):
            np.testing.assert_array_equal(a["var%d" % v].view(np.uint32), b["var%d" % v].view(np.uint32))


def test_command_line_prints_the_status_the_operator_would_write(tmp_path):
    r = subprocess.run([sys.executable, "-m", "k8s_b200.run", TF_JOB, "--gpus", "1", "--workload", "syncps",
                        "--steps", "2", "--batch", "8", "--status_file", str(tmp_path / "st.json")],
                       capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    st = json.loads(r.stdout)
    assert st["phase"] == "Done" and st["state"] == "Succeeded"
    doc = json.load(open(tmp_path / "st.json"))
    assert doc["kind"] == "TfJob" and doc["status"]["phase"] == "Done" and doc["metadata"]["name"] == "example-job"
