"""Checkpoint of the sharded variables (k8s_b200/checkpoint.py, SURVEY.md par. 8f row 3): save after two
steps, restore into a fresh job -- also one with another layout or another number of (emulated) GPU
ranks -- and continue: parameters, momentum slots and global_step must be bit-identical to the
uninterrupted run."""
import json

import numpy as np
import pytest

from k8s_b200 import checkpoint
from k8s_b200 import lib as L
from k8s_b200.syncps import SyncPSJob, synthetic_batch

pytestmark = pytest.mark.gpu
BATCH = 8


def cfg():
    return json.dumps({"cluster": {"master": ["m-0:2222"], "worker": ["w-0:2222"], "ps": ["p-0:2222", "p-1:2222"]},
                       "task": {"type": "master", "index": 0}, "environment": "cloud"})


def build(layout="logical", math="tf32x3"):
    j = SyncPSJob(cfg(), 0, 1, cuda_device=0)
    j.build_mnist(BATCH, math=math, layout=layout)
    j.optimizer("momentum", 0.001, 0.9)
    return j


def steps(j, first, n):
    out = []
    for s in range(first, first + n):
        b = [synthetic_batch(r, BATCH, s) for r in range(2)]
        out.append(j.step([x for x, _ in b], [y for _, y in b]))
    return out


def state(j):
    return [j.get_var(v).view(np.uint32).copy() for v in range(8)] + [j.get_accum(v).view(np.uint32).copy() for v in range(8)]


@pytest.mark.parametrize("restore_layout", ["logical", "balanced"])
@pytest.mark.parametrize("math", ["tf32x3", "bf16"])
def test_round_trip_is_bit_exact(tmp_path, restore_layout, math):
    a = build("logical", math)
    steps(a, 0, 2)
    checkpoint.save(a, str(tmp_path))
    tail_a = steps(a, 2, 2)
    want = state(a)
    doc, recs = checkpoint.read_records(str(tmp_path))
    assert doc["global_step"] == 2 and doc["n_gpus"] == 1 and sum(r[1].size for r in recs) == a.flat_count
    assert [v["ps_task"] for v in doc["variables"][:8]] == [0, 1, 0, 1, 0, 1, 0, 1]
    a.close()
    b = build(restore_layout, math)
    assert checkpoint.restore(b, str(tmp_path)) == 2 and b.global_step == 2
    tail_b = steps(b, 2, 2)           # the operand-precision shadow of the weights must have been refreshed too
    assert tail_a == tail_b
    for x, y in zip(want, state(b)):
        np.testing.assert_array_equal(x, y)
    assert b.global_step == 4
    b.close()


def test_two_rank_save_partitions_the_flat_space_and_restores_on_one_rank(tmp_path):
    """Two (emulated, same-device) GPU ranks each write their own shard file; the files partition the
    flat space; a one-rank job restores from them and continues bit-identically."""
    two = [SyncPSJob(cfg(), r, 2, cuda_device=0) for r in range(2)]
    for j in two:
        j.build_mnist(BATCH, math="fp32", layout="balanced")
        j.optimizer("momentum", 0.001, 0.9)
    two[0].attach_local_peer(1, two[1])
    two[1].attach_local_peer(0, two[0])
    torch = pytest.importorskip("torch")
    for s in range(2):
        for r, j in enumerate(two):
            x, y = synthetic_batch(r, BATCH, s)
            dx, dy = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
            j.forward_backward(0, dx.data_ptr(), dy.data_ptr())
            j.sync()
        for j in two:
            j.push_apply_pull(with_barriers=False)
        for j in two:
            j.sync()
    for j in two:
        checkpoint.save(j, str(tmp_path), step=2)
    doc, recs = checkpoint.read_records(str(tmp_path))
    covered = np.zeros(doc["flat_count"], dtype=np.int32)
    for off, p, _ in recs:
        covered[off:off + p.size] += 1
    assert doc["n_gpus"] == 2 and (covered == 1).all()
    one = SyncPSJob(cfg(), 0, 1, cuda_device=0)
    one.build_mnist(BATCH, math="fp32", layout="logical")
    one.optimizer("momentum", 0.001, 0.9)
    assert checkpoint.restore(one, str(tmp_path)) == 2
    for v in range(8):
        np.testing.assert_array_equal(one.get_var(v).view(np.uint32), two[0].get_var(v).view(np.uint32))
        own = two[0].placement(v)
        # each accumulator element lives on the rank that owns it
        acc = np.where(_owned_mask(two[0], own), two[0].get_accum(v), two[1].get_accum(v))
        np.testing.assert_array_equal(one.get_accum(v).view(np.uint32), acc.view(np.uint32))
    for j in two + [one]:
        j.close()


def _owned_mask(job, pl):
    m = np.zeros(pl["count"], dtype=bool)
    for off, cnt in job.owned_ranges():
        lo, hi = max(off, pl["offset"]), min(off + cnt, pl["offset"] + pl["count"])
        if lo < hi:
            m[lo - pl["offset"]:hi - pl["offset"]] = True
    return m


def test_restore_rejects_a_checkpoint_of_another_model_and_a_truncated_file(tmp_path):
    a = build()
    checkpoint.save(a, str(tmp_path))
    f = tmp_path / "shard-0-of-1.bin"
    data = f.read_bytes()
    f.write_bytes(data[:-8])
    with pytest.raises(ValueError):
        checkpoint.restore(a, str(tmp_path))
    f.write_bytes(data)
    assert checkpoint.restore(a, str(tmp_path)) == 0
    assert a.lib.flat_read(a.h, L.BUF_LOSS, 0, 1, None) != 0       # only params / accum are addressable
    a.close()
