"""Checkpoint of the sharded PS variables (SURVEY.md par. 8f row 3): one flat file per shard plus
a placement table.

The reference has no checkpoint format (its operator only persists the TfJob status,
pkg/controller/controller.go:172-201, and skips setup when Phase != None,
pkg/trainer/training.go:252-257), so this one is designed for the layout in HBM:

  <dir>/placement.json          written by GPU rank 0: format version, global_step, layout, n_gpus,
                                length of the flat space, and the variable table (name, shape,
                                trainable, ps_task, owner gpu, offset, count) in creation order
  <dir>/shard-<r>-of-<n>.bin    written by GPU rank r: for each flat-space range whose update the
                                rank applies, a 24-byte record header (int64 offset, int64 count,
                                int64 reserved) followed by `count` fp32 parameters and `count`
                                fp32 optimizer-slot values (momentum accumulator)

Every element of the flat space is owned by exactly one rank (tfg_owned_ranges), so the shard files
of one save partition it.  Restore does not need the saver's sharding: a rank reads ALL shard
files, fills its replicated parameter buffer, and its accumulator for the whole space (only the
ranges it owns are ever used) -- a job saved on 8 GPUs restores on 1, or with the other layout.
Round trip is bit-exact (tests/test_gpu_checkpoint.py).
"""
from __future__ import annotations

import glob
import json
import os
import struct
from typing import List

import numpy as np

from . import lib as L

FORMAT = "tfgraft-ckpt-1"


def exists(path: str) -> bool:
    return os.path.exists(os.path.join(path, "placement.json"))


def save(job, path: str, step: int = None) -> None:
    """Called by every GPU rank of the job (no collective: each rank writes only its own file)."""
    os.makedirs(path, exist_ok=True)
    name = os.path.join(path, "shard-%d-of-%d.bin" % (job.gpu_rank, job.n_gpus))
    with open(name + ".tmp", "wb") as f:
        for off, cnt in job.owned_ranges():
            f.write(struct.pack("<qqq", off, cnt, 0))
            f.write(job.flat_read(L.BUF_PARAMS, off, cnt).tobytes())
            f.write(job.flat_read(L.BUF_ACCUM, off, cnt).tobytes())
    os.replace(name + ".tmp", name)
    if job.gpu_rank == 0:
        table = []
        for v in range(job.n_vars):
            vname, shape, trainable = job.var_info(v)
            pl = job.placement(v)
            table.append({"name": vname, "shape": list(shape), "trainable": trainable, "ps_task": pl["ps_task"],
                          "gpu": pl["gpu"], "offset": pl["offset"], "count": pl["count"]})
        doc = {"format": FORMAT, "global_step": int(job.global_step if step is None else step),
               "n_gpus": job.n_gpus, "flat_count": job.flat_count, "variables": table}
        with open(os.path.join(path, "placement.json.tmp"), "w") as f:
            json.dump(doc, f, indent=1)
        os.replace(os.path.join(path, "placement.json.tmp"), os.path.join(path, "placement.json"))


def read_records(path: str):
    """-> (placement dict, [(offset, params, accum)]) of a saved checkpoint (numpy only)."""
    doc = json.load(open(os.path.join(path, "placement.json")))
    if doc.get("format") != FORMAT:
        raise ValueError("%s is not a %s checkpoint" % (path, FORMAT))
    files = sorted(glob.glob(os.path.join(path, "shard-*-of-%d.bin" % doc["n_gpus"])))
    if len(files) != doc["n_gpus"]:
        raise ValueError("checkpoint %s holds %d of %d shard files" % (path, len(files), doc["n_gpus"]))
    recs = []
    for name in files:
        with open(name, "rb") as f:
            while True:
                h = f.read(24)
                if not h:
                    break
                off, cnt, _ = struct.unpack("<qqq", h)
                p = np.frombuffer(f.read(4 * cnt), dtype=np.float32)
                a = np.frombuffer(f.read(4 * cnt), dtype=np.float32)
                if p.size != cnt or a.size != cnt:
                    raise ValueError("truncated shard file %s" % name)
                recs.append((off, p, a))
    covered = sum(r[1].size for r in recs)
    if covered != doc["flat_count"]:
        raise ValueError("shard files cover %d of %d elements" % (covered, doc["flat_count"]))
    return doc, recs


def restore(job, path: str) -> int:
    """Load a checkpoint into a built, finalized job of the same model (any n_gpus / layout);
    returns the saved global_step (also set on the job)."""
    doc, recs = read_records(path)
    names: List[str] = [job.var_info(v)[0] for v in range(job.n_vars)]
    if [t["name"] for t in doc["variables"]] != names or doc["flat_count"] != job.flat_count:
        raise ValueError("checkpoint %s was written by a different model" % path)
    for off, p, a in recs:
        job.flat_write(L.BUF_PARAMS, off, p)
        job.flat_write(L.BUF_ACCUM, off, a)
    job.global_step = doc["global_step"]
    return int(doc["global_step"])
