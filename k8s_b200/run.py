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

from . import spec as S
from .backend import GpuRankBackend
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


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("tf_job")
    ap.add_argument("--gpus", type=int, default=0, help="GPU ranks of this box (0 = all visible)")
    ap.add_argument("--controller_config_file", default="")
    a = ap.parse_args(argv)
    n_gpus = a.gpus
    if n_gpus <= 0:
        from .lib import load
        n_gpus = max(1, load().device_count_or_zero())
    cfg = S.ControllerConfig()
    if a.controller_config_file:
        import yaml
        cfg = S.ControllerConfig.from_dict(yaml.safe_load(open(a.controller_config_file)))
    tj, backend = run_tf_job(a.tf_job, n_gpus, cfg)
    print(json.dumps(tj.job.to_dict()["status"], indent=2))
    return 0 if tj.job.status.state == S.STATE_SUCCEEDED else 1


if __name__ == "__main__":
    sys.exit(main())
