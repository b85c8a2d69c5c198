"""The training replica program of a synchronous PS job: what a MASTER / WORKER container of a
TfJob runs when the workload is the sync-PS training step instead of the smoke multiply.

Launch contract = the reference's (SURVEY.md par. 8b B2/B3): one OS process per (type, index),
``TF_CONFIG`` in the environment (pkg/trainer/replicas.go:189-203,241-255), the exit code decides the
replica's state (pkg/trainer/training.go:203-238; a TfgError exits with its own code, 1..127
permanent, >= 128 retryable).  Same role dispatch as examples/tf_sample/tf_sample/tf_smoke.py:125-138.

Mapping onto the box (DESIGN.md "Process and rank model"): training replica r computes on GPU
r % n_gpus, PS task k is owned by GPU k % n_gpus.  The lowest replica of a GPU is that GPU's rank
host: its process owns the ``tfg_job`` of the rank, hosts every replica and PS shard mapped to the
GPU, exchanges CUDA-IPC handles with the other rank hosts through the launcher's rendezvous socket
(k8s_b200/rendezvous.py) and runs the steps.  A replica that shares its GPU with a lower one has
nothing to drive -- "The worker just blocks because we let the master assign all ops"
(tf_smoke.py:131) -- and waits for the job to finish.
"""
from __future__ import annotations

import argparse
import json
import logging
import os
import sys

import numpy as np


def parse_args(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--steps", type=int, default=3)
    p.add_argument("--batch", type=int, default=128, help="examples per replica per step")
    p.add_argument("--math", default="tf32x3", choices=["tf32x3", "fp32", "bf16", "tf32"])
    p.add_argument("--opt", default="momentum", choices=["sgd", "momentum", "nesterov"])
    p.add_argument("--lr", type=float, default=0.001)
    p.add_argument("--momentum", type=float, default=0.9)
    p.add_argument("--layout", default="logical", choices=["logical", "balanced"])
    p.add_argument("--n_gpus", type=int, default=0, help="GPU ranks of the box (0 = TFGRAFT_N_GPUS, else all visible)")
    p.add_argument("--result_dir", default="", help="rank hosts drop losses and final parameters here (npz)")
    p.add_argument("--checkpoint_dir", default="", help="resume from / save to this directory (k8s_b200.checkpoint)")
    args, _ = p.parse_known_args(argv)
    return args


def main(argv=None) -> int:
    args = parse_args(argv)
    tf_config_json = os.environ.get("TF_CONFIG", "{}")
    tf_config = json.loads(tf_config_json)
    task = tf_config.get("task", {})
    cluster = tf_config.get("cluster", {})
    job_type = str(task.get("type", "")).lower()
    index = int(task.get("index", 0))
    logging.info("tf_config: %s", tf_config)
    from . import rendezvous
    from .cluster import RankMap
    from .lib import TfgError, load
    n_gpus = args.n_gpus or int(os.environ.get("TFGRAFT_N_GPUS", "0")) or max(1, load().device_count_or_zero())
    rm = RankMap(cluster, n_gpus)
    has_rdv = rendezvous.ENV in os.environ
    if job_type == "ps":
        logging.info("Running PS code.")            # a shard owner is hosted by its GPU's rank host
        return rendezvous.Client(0, 1).wait_done() if has_rdv else 0
    if job_type not in ("master", "worker", ""):
        raise ValueError("invalid job_type %s" % (job_type,))
    replica = rm.replica_index(job_type or "master", index)
    gpu = replica % n_gpus
    host = min(rm.local_replicas(gpu))
    if replica != host:
        logging.info("Running Worker code: replica %d is hosted by replica %d on GPU rank %d.", replica, host, gpu)
        return rendezvous.Client(0, 1).wait_done() if has_rdv else 0
    logging.info("Running %s: rank host of GPU rank %d of %d.", job_type or "master", gpu, n_gpus)

    from .syncps import SyncPSJob, synthetic_batch
    try:
        job = SyncPSJob(tf_config_json, gpu, n_gpus, cuda_device=gpu)
        job.build_mnist(args.batch, math=args.math, layout=args.layout)
        job.optimizer(args.opt, args.lr, args.momentum)
        client = rendezvous.Client(gpu, n_gpus) if n_gpus > 1 else None
        job.connect_peers(client)
        first_step = 0
        if args.checkpoint_dir:
            from . import checkpoint
            if checkpoint.exists(args.checkpoint_dir):
                first_step = checkpoint.restore(job, args.checkpoint_dir)
                if client is not None:
                    client.barrier()       # every rank has overwritten its own buffers before anyone steps
        losses = []
        for step in range(first_step, first_step + args.steps):
            xs, ys = zip(*[synthetic_batch(r, args.batch, step) for r in job.local_replicas])
            losses.append(job.step(list(xs), list(ys)))
            logging.info("step %d loss %.6f", step, losses[-1])
        if args.checkpoint_dir:
            from . import checkpoint
            checkpoint.save(job, args.checkpoint_dir, first_step + args.steps)
        if args.result_dir:
            os.makedirs(args.result_dir, exist_ok=True)
            out = {"losses": np.asarray(losses, dtype=np.float64), "global_step": np.asarray(job.global_step),
                   "ps_task": np.asarray([job.placement(v)["ps_task"] for v in range(8)])}
            for v in range(8):
                out["var%d" % v] = job.get_var(v)
            np.savez(os.path.join(args.result_dir, "rank%d.npz" % gpu), **out)
        if client is not None:
            client.barrier()               # nobody unmaps a peer's buffers while a step may still read them
        job.close()
    except TfgError as e:
        logging.error("%s", e)
        return e.code
    return 0


if __name__ == "__main__":
    logging.getLogger().setLevel(logging.INFO)
    sys.exit(main())
