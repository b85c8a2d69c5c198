"""Default parameter-server replica process, GPU-rank edition.

Mirror of ``grpc_tensorflow_server/grpc_tensorflow_server.py``: the same three flags
(``--cluster_spec --job_name --task_id``, reference :119-157), the same
``parse_cluster_spec`` grammar and ``ValueError`` messages (reference :46-88).  Where
the reference builds a ``ServerDef(protocol="grpc")`` and blocks in ``Server.join()``
(:93-115), a PS task here is a shard owner on a GPU rank: it has no process of its own
to block in, so ``main`` resolves which GPU owns task ``task_id`` and returns.
"""
from __future__ import annotations

import argparse
import os
import sys
from collections import OrderedDict
from typing import Dict, List


def parse_cluster_spec(cluster_spec: str, cluster: "OrderedDict[str, Dict[int, str]]" = None,
                       verbose: bool = False) -> "OrderedDict[str, Dict[int, str]]":
    """``job|h:p;h:p,job|...`` -> ordered ``{job: {task_index: address}}``.

    Follows grpc_tensorflow_server.py:46-88 check for check: empty string; exactly one
    ``|`` per job; non-empty job name; non-empty task strings.  Like the reference (which
    ``add()``s the job to the protobuf before validating it, :65-68) a job that fails a
    check has already been appended when the error is raised."""
    if cluster is None:
        cluster = OrderedDict()
    job_strings = cluster_spec.split(",")
    if not cluster_spec:
        raise ValueError("Empty cluster_spec string")
    for n, job_string in enumerate(job_strings):
        tasks: Dict[int, str] = {}
        key = "\0pending%d" % n
        cluster[key] = tasks
        if job_string.count("|") != 1:
            raise ValueError("Not exactly one instance of '|' in cluster_spec")
        job_name = job_string.split("|")[0]
        if not job_name:
            raise ValueError("Empty job_name in cluster_spec")
        # rename the placeholder in place, keeping insertion order; repeated job names
        # stay separate entries in the reference's repeated field, here the later wins
        del cluster[key]
        cluster[job_name] = tasks
        if verbose:
            print("Added job named \"%s\"" % job_name)
        job_tasks = job_string.split("|")[1].split(";")
        for i in range(len(job_tasks)):
            if not job_tasks[i]:
                raise ValueError("Empty task string at position %d" % i)
            tasks[i] = job_tasks[i]
            if verbose:
                print("  Added task \"%s\" to job \"%s\"" % (job_tasks[i], job_name))
    return cluster


def cluster_lists(cluster: "OrderedDict[str, Dict[int, str]]") -> Dict[str, List[str]]:
    """``{job: {i: addr}}`` -> ``{job: [addr...]}`` (the TF_CONFIG ``cluster`` form)."""
    return {j: [t[i] for i in sorted(t)] for j, t in cluster.items()}


def build_parser() -> argparse.ArgumentParser:
    """grpc_tensorflow_server.py:119-157 (same names, types and defaults)."""
    parser = argparse.ArgumentParser()
    parser.register("type", "bool", lambda v: v.lower() == "true")
    parser.add_argument("--cluster_spec", type=str, default="")
    parser.add_argument("--job_name", type=str, default="")
    parser.add_argument("--task_id", type=int, default=0)
    parser.add_argument("--gpu_memory_fraction", type=float, default=1.0)
    parser.add_argument("--verbose", type="bool", nargs="?", const=True, default=False)
    parser.add_argument("--n_gpus", type=int, default=0,
                        help="GPUs of this box (0 = ask the driver); not a reference flag")
    parser.add_argument("--join", action="store_true",
                        help="block like server.join() (grpc_tensorflow_server.py:115) until the launcher's "
                             "rendezvous reports the job finished; not a reference flag")
    return parser


def main(argv=None) -> int:
    flags, _ = build_parser().parse_known_args(argv)
    cluster = parse_cluster_spec(flags.cluster_spec, verbose=flags.verbose)
    if not flags.job_name:
        raise ValueError("Empty job_name")
    if flags.task_id < 0:
        raise ValueError("Invalid task_id: %d" % flags.task_id)
    from .cluster import RankMap
    n_gpus = flags.n_gpus
    if n_gpus <= 0:
        from .lib import load
        n_gpus = max(1, load().device_count_or_zero())
    rm = RankMap(cluster_lists(cluster), n_gpus)
    gpu = rm.gpu_of(flags.job_name, flags.task_id)
    print("/job:%s/task:%d -> shard owner on GPU rank %d of %d"
          % (flags.job_name, flags.task_id, gpu, n_gpus), flush=True)
    if flags.join:
        # server.join(): the shard itself lives in the tfg_job of GPU rank `gpu`'s host process; this
        # process stays up, like the reference's PS pod, until the job is over
        from . import rendezvous
        if rendezvous.ENV in os.environ:
            return rendezvous.Client(0, 1).wait_done()
        import time
        while True:
            time.sleep(3600)
    return 0


if __name__ == "__main__":
    sys.exit(main())
