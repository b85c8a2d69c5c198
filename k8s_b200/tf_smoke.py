"""The sample workload of the TfJob examples, on GPU ranks.

Mirror of ``examples/tf_sample/tf_sample/tf_smoke.py``: read ``TF_CONFIG`` (:88-96),
dispatch on the task type (:125-138) -- ``ps`` and ``worker`` serve, ``master`` (or no type)
drives -- and in ``run`` place ``c = a * b`` with ``a = b = arange(100).reshape(10, 10)``
(int32) on every task of the cluster and fetch each result (:53-60, :74-76).  Where the
reference starts a ``tf.train.Server`` and ships sub-graphs over gRPC, each
``/job:J/task:i`` here is a GPU rank and the multiply is a CUDA kernel behind
``tfg_smoke``.
"""
from __future__ import annotations

import argparse
import json
import logging
import os
import time
from typing import Dict, List

import numpy as np


def parse_args(argv=None):
    """tf_smoke.py:18-31."""
    parser = argparse.ArgumentParser()
    parser.add_argument("--sleep_secs", default=0, type=int,
                        help="Amount of time to sleep at the end")
    args, _ = parser.parse_known_args(argv)
    return args


def run(server, cluster_spec: Dict[str, List[str]]) -> Dict[str, np.ndarray]:
    """tf_smoke.py:34-76.  ``server`` is the SyncPSJob of this process (or None, the
    reference's direct-session branch).  Returns {"/job:J/task:i": 10x10 int32}; the
    reference visits ``cluster_spec.keys()`` in unspecified (py2 dict) order, so results
    are keyed by device, not by position."""
    results: Dict[str, np.ndarray] = {}
    if server is None:
        from .syncps import SyncPSJob
        server = SyncPSJob(json.dumps({"cluster": cluster_spec}), 0, 1)
    out = server.smoke(all_tasks=True)
    t = 0
    for job_name in sorted(cluster_spec.keys()):
        for i in range(len(cluster_spec[job_name])):
            d = "/job:{0}/task:{1}".format(job_name, i)
            results[d] = out[t]
            logging.info("Result: %s", out[t])
            t += 1
    return results


def main(argv=None, n_gpus: int = 1, gpu_rank: int = 0):
    """tf_smoke.py:79-138."""
    tf_config_json = os.environ.get("TF_CONFIG", "{}")
    tf_config = json.loads(tf_config_json)
    logging.info("tf_config: %s", tf_config)
    task = tf_config.get("task", {})
    logging.info("task: %s", task)
    cluster_spec = tf_config.get("cluster", {})
    logging.info("cluster_spec: %s", cluster_spec)

    server = None
    if cluster_spec:
        from .syncps import SyncPSJob
        logging.info("Building server.")
        server = SyncPSJob(tf_config_json, gpu_rank, n_gpus)
        logging.info("Finished building server.")
    else:
        logging.error("Using default device function.")

    job_type = task.get("type", "").lower()
    if job_type == "ps":
        logging.info("Running PS code.")
        return None          # server.join(): a shard owner has nothing to drive
    elif job_type == "worker":
        logging.info("Running Worker code.")
        return None          # "The worker just blocks because we let the master assign all ops"
    elif job_type == "master" or not job_type:
        logging.info("Running master.")
        res = run(server=server, cluster_spec=cluster_spec)
        args = parse_args(argv)
        if args.sleep_secs:
            time.sleep(args.sleep_secs)
        return res
    else:
        raise ValueError("invalid job_type %s" % (job_type,))


if __name__ == "__main__":
    logging.getLogger().setLevel(logging.INFO)
    main()
