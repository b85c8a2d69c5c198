"""The local-process launcher (k8s_b200.backend.ProcessBackend + k8s_b200.run.run_tf_job_processes):
an unchanged TfJob document -> one OS process per task index with the trainer's TF_CONFIG -> exit
codes -> TfJob status, and the reference-style poller (py/tf_job_client.py:63-96) over it.  CPU only:
the training replicas' program is replaced by small scripts; the default PS runs the real
k8s_b200.grpc_tensorflow_server with the reference's command line."""
import datetime
import json
import os
import sys
import threading

import pytest

from k8s_b200 import rendezvous
from k8s_b200 import spec as S
from k8s_b200 import tf_job_client
from k8s_b200.run import run_tf_job_processes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TF_JOB = "/root/reference/examples/tf_job.yaml"
if not os.path.exists(TF_JOB):
    TF_JOB = os.path.join(ROOT, "examples", "tf_job.yaml")


def test_rendezvous_all_gather_and_done(tmp_path):
    srv = rendezvous.Server(str(tmp_path / "r.sock"))
    out = {}

    def rank(r):
        c = rendezvous.Client(r, 3, srv.path, timeout=20)
        got = [None] * 3
        c.all_gather_object(got, b"blob-%d" % r)
        c.barrier()
        out[r] = got

    ts = [threading.Thread(target=rank, args=(r,)) for r in range(3)]
    [t.start() for t in ts]
    [t.join(30) for t in ts]
    assert all(out[r] == [b"blob-0", b"blob-1", b"blob-2"] for r in range(3))
    waited = []
    t = threading.Thread(target=lambda: waited.append(rendezvous.Client(0, 1, srv.path).wait_done()))
    t.start()
    srv.finish(7)
    t.join(10)
    assert waited == [7]
    srv.close()


def test_rendezvous_times_out_when_a_rank_is_missing(tmp_path):
    srv = rendezvous.Server(str(tmp_path / "r.sock"))
    c = rendezvous.Client(0, 2, srv.path, timeout=0.5)
    with pytest.raises(TimeoutError):
        c.barrier()
    srv.close()


def script(tmp_path, body):
    p = tmp_path / "replica.py"
    p.write_text(body)
    return [sys.executable, str(p)]


def test_tf_job_yaml_runs_one_process_per_task_and_the_poller_sees_done(tmp_path):
    """examples/tf_job.yaml: 1 MASTER + 1 WORKER + 2 default PS = four processes; the MASTER's exit 0
    makes the job Succeeded (training.go:187-195) while WORKER and PS are still serving."""
    cmd = script(tmp_path, "import json, os, sys, time\n"
                           "t = json.loads(os.environ['TF_CONFIG'])['task']\n"
                           "open(os.path.join(%r, 'seen-%%s-%%d' %% (t['type'], t['index'])), 'w').write(os.environ['TF_CONFIG'])\n"
                           "time.sleep(0 if t['type'] == 'master' else 60)\n" % str(tmp_path))
    status_file = str(tmp_path / "status.json")
    polled = []
    poller_result = {}

    def poll():
        poller_result["job"] = tf_job_client.wait_for_job(
            status_file, "default", "example-job", timeout=datetime.timedelta(seconds=60),
            polling_interval=datetime.timedelta(milliseconds=50), status_callback=lambda j: polled.append(j["status"]["phase"]))

    t = threading.Thread(target=poll)
    t.start()
    tj, backend = run_tf_job_processes(TF_JOB, n_gpus=2, workload_cmd=cmd, timeout=60, interval=0.05,
                                       log_dir=str(tmp_path / "logs"), status_file=status_file)
    t.join(30)
    try:
        st = tj.job.status
        assert (st.phase, st.state) == (S.PHASE_DONE, S.STATE_SUCCEEDED), backend.log(next(iter(backend.launches)))
        assert len(backend.procs) == 4
        rid = tj.job.spec.RuntimeId
        names = sorted(backend.launches)
        assert names == sorted(["example-job-master-%s-0" % rid, "example-job-worker-%s-0" % rid,
                                "example-job-ps-%s-0" % rid, "example-job-ps-%s-1" % rid])
        # every training replica saw exactly the trainer's TF_CONFIG (Go marshalling order: cluster, task, environment)
        cfg = json.loads(open(tmp_path / "seen-master-0").read())
        assert list(cfg) == ["cluster", "task", "environment"] and cfg["environment"] == "cloud"
        assert cfg["cluster"]["ps"] == ["example-job-ps-%s-0:2222" % rid, "example-job-ps-%s-1:2222" % rid]
        # the default PS keeps the reference's argv (replicas.go:205-208) behind the local interpreter
        ps = backend.command(backend.launches["example-job-ps-%s-1" % rid])
        assert ps[3:9] == ["--cluster_spec", ps[4], "--job_name", "ps", "--task_id", "1"]
        assert "ps|example-job-ps-%s-0:2222;example-job-ps-%s-1:2222" % (rid, rid) in ps[4]
        # the poller of py/tf_job_client.py saw the job reach Done with state Succeeded
        assert poller_result["job"]["status"]["phase"] == "Done"
        assert poller_result["job"]["status"]["state"] == "Succeeded"
        assert polled[-1] == "Done" and set(polled) <= {"", "Creating", "Running", "Done"}
        rs = {r["tf_replica_type"]: r for r in poller_result["job"]["status"]["replicaStatuses"]}
        assert rs["MASTER"]["state"] == "Succeeded" and rs["PS"]["ReplicasStates"] == {"Running": 2}
        # CleanUp terminates the serving replicas
        tj.delete()
        assert tj.job.status.phase == S.PHASE_CLEANUP
    finally:
        procs = list(backend.procs.values())
        backend.close()
    assert all(p.poll() is not None for p in procs)


def test_retryable_exit_restarts_the_replica_and_permanent_exit_fails_the_job(tmp_path):
    marker = tmp_path / "ran-once"
    cmd = script(tmp_path, "import json, os, sys, time\n"
                           "t = json.loads(os.environ['TF_CONFIG'])['task']\n"
                           "if t['type'] != 'master': time.sleep(60)\n"
                           "if not os.path.exists(%r):\n"
                           "    open(%r, 'w').close(); sys.exit(130)\n"
                           "sys.exit(0)\n" % (str(marker), str(marker)))
    tj, backend = run_tf_job_processes(TF_JOB, n_gpus=1, workload_cmd=cmd, timeout=60, interval=0.05)
    try:
        assert (tj.job.status.phase, tj.job.status.state) == (S.PHASE_DONE, S.STATE_SUCCEEDED)
        assert sum(backend.restarts.values()) == 1
    finally:
        backend.close()
    cmd = script(tmp_path, "import json, os, sys, time\n"
                           "t = json.loads(os.environ['TF_CONFIG'])['task']\n"
                           "if t['type'] != 'master': time.sleep(60)\n"
                           "sys.exit(3)\n")
    tj, backend = run_tf_job_processes(TF_JOB, n_gpus=1, workload_cmd=cmd, timeout=60, interval=0.05)
    try:
        assert (tj.job.status.phase, tj.job.status.state) == (S.PHASE_DONE, S.STATE_FAILED)
    finally:
        backend.close()


def test_poller_times_out_like_the_reference(tmp_path):
    f = tmp_path / "status.json"
    f.write_text(json.dumps({"metadata": {"name": "j", "namespace": "default"}, "status": {"phase": "Running", "state": "Running"}}))
    with pytest.raises(tf_job_client.TimeoutError, match="Timeout waiting for job j in namespace default to finish."):
        tf_job_client.wait_for_job(str(f), "default", "j", timeout=datetime.timedelta(milliseconds=200),
                                   polling_interval=datetime.timedelta(milliseconds=50))


def test_in_process_backend_reruns_a_replica_after_a_retryable_exit():
    """GpuRankBackend: a TfgError with a retryable code (>= 128) leaves the replica Running
    (training.go:203-238) and the next run() executes it again."""
    from k8s_b200.backend import GpuRankBackend
    from k8s_b200.lib import TfgError
    from k8s_b200.trainer import TrainingJob
    calls = []

    def workload(n_gpus, gpu_rank):
        calls.append(1)
        if len(calls) == 1:
            raise TfgError(130, "transient CUDA error")
        return "ok"

    be = GpuRankBackend(n_gpus=1, workload=workload)
    tj = TrainingJob(be, S.load_tf_job(TF_JOB))
    cfg = S.ControllerConfig()
    for _ in range(4):
        tj.reconcile(cfg)
        if tj.job.status.phase == S.PHASE_DONE:
            break
        be.run()
    assert len(calls) == 2
    assert (tj.job.status.phase, tj.job.status.state) == (S.PHASE_DONE, S.STATE_SUCCEEDED)
