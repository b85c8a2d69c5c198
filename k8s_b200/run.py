"""Run a TfJob document on this box: ``python -m k8s_b200.run examples/tf_job.yaml``.

The one-box stand-in for ``kubectl create -f`` + the operator's reconcile loop
(pkg/trainer/training.go:412-456): load the YAML unchanged, reconcile until the job phase
is Done or Failed, print the TfJob status the operator would have written
(pkg/spec/tf_job.go:347-383).
"""
from __future__ import annotations

import argparse
import json
import sys
import time

from . import spec as S
from .backend import GpuRankBackend, ProcessBackend
from .trainer import TrainingJob


def run_tf_job(path: str, n_gpus: int = 1, config: S.ControllerConfig = None, max_reconciles: int = 4,
               workload=None):
    job = S.load_tf_job(path)
    backend = GpuRankBackend(n_gpus=n_gpus, workload=workload)
    tj = TrainingJob(backend, job)
    config = config or S.ControllerConfig()
    for _ in range(max_reconciles):
        tj.reconcile(config)
        if tj.job.status.phase in (S.PHASE_DONE, S.PHASE_FAILED):
            break
        backend.run()
    return tj, backend


def run_tf_job_processes(path: str, n_gpus: int = 1, workload: str = "syncps", workload_args=None,
                         config: S.ControllerConfig = None, timeout: float = 600.0, interval: float = 0.2,
                         log_dir: str = None, status_file: str = None, on_status=None, workload_cmd=None):
    """The document at `path`, unchanged, driven the way the operator drives it: reconcile (which
    creates one replica PROCESS per task index through ProcessBackend, each with the TF_CONFIG the
    trainer built), poll the replicas' exit codes every `interval` seconds (the reference's 8 s,
    pkg/trainer/training.go:23), until the phase is Done or Failed.  The whole TfJob object is
    written to `status_file` after every reconcile -- what `kubectl get tfjob -o json` returns, and
    what k8s_b200.tf_job_client.wait_for_job polls.  Returns (TrainingJob, backend); the caller
    closes the backend (CleanUp)."""
    job = S.load_tf_job(path)
    backend = ProcessBackend(n_gpus=n_gpus, workload=workload, workload_args=workload_args, log_dir=log_dir,
                             workload_cmd=workload_cmd)
    tj = TrainingJob(backend, job)
    config = config or S.ControllerConfig()
    deadline = time.time() + timeout

    def publish():
        doc = tj.job.to_dict()
        doc.setdefault("metadata", {}).setdefault("namespace", "default")
        if status_file:
            with open(status_file + ".tmp", "w") as f:
                json.dump(doc, f)
            import os
            os.replace(status_file + ".tmp", status_file)
        if on_status:
            on_status(doc)

    while True:
        tj.reconcile(config)
        publish()
        if tj.job.status.phase in (S.PHASE_DONE, S.PHASE_FAILED):
            break
        if time.time() > deadline:
            tj.status.phase, tj.status.state = S.PHASE_FAILED, S.STATE_FAILED
            tj.status.reason = "timeout waiting for the replicas after %.0f s" % timeout
            tj.update_tpr_status()
            publish()
            break
        time.sleep(interval)
    backend.finish(0 if tj.job.status.state == S.STATE_SUCCEEDED else 1)
    return tj, backend


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("tf_job")
    ap.add_argument("--gpus", type=int, default=0, help="GPU ranks of this box (0 = all visible)")
    ap.add_argument("--controller_config_file", default="")
    ap.add_argument("--workload", default="smoke", choices=["smoke", "syncps"],
                    help="smoke = the sample image's multiply, in this process (tf_smoke.py); syncps = the "
                         "synchronous PS training step, one process per replica (k8s_b200.tf_sync_ps)")
    ap.add_argument("--status_file", default="", help="syncps: the TfJob object as JSON, rewritten after every reconcile")
    ap.add_argument("--log_dir", default="", help="syncps: one log per replica process")
    ap.add_argument("--timeout", type=float, default=600.0)
    a, rest = ap.parse_known_args(argv)
    n_gpus = a.gpus
    if n_gpus <= 0:
        from .lib import load
        n_gpus = max(1, load().device_count_or_zero())
    cfg = S.ControllerConfig()
    if a.controller_config_file:
        import yaml
        cfg = S.ControllerConfig.from_dict(yaml.safe_load(open(a.controller_config_file)))
    if a.workload == "syncps":       # remaining arguments go to every training replica (k8s_b200.tf_sync_ps)
        tj, backend = run_tf_job_processes(a.tf_job, n_gpus, "syncps", rest, cfg, a.timeout,
                                           log_dir=a.log_dir or None, status_file=a.status_file or None)
        status = tj.job.to_dict()["status"]
        if tj.job.status.state != S.STATE_SUCCEEDED:
            for name in backend.launches:
                sys.stderr.write("---- %s ----\n%s\n" % (name, backend.log(name)[-2000:]))
        tj.delete()                  # CleanUp: the replica processes that only serve (default PS) are terminated
        backend.close()
        print(json.dumps(status, indent=2))
        return 0 if status["state"] == S.STATE_SUCCEEDED else 1
    if rest:
        ap.error("unrecognized arguments: %s" % " ".join(rest))
    tj, backend = run_tf_job(a.tf_job, n_gpus, cfg)
    print(json.dumps(tj.job.to_dict()["status"], indent=2))
    return 0 if tj.job.status.state == S.STATE_SUCCEEDED else 1


if __name__ == "__main__":
    sys.exit(main())
