"""World-size-2 test of the host-side multi-rank logic on CPU (gloo): every rank derives
its TF_CONFIG through the trainer, hosts the replicas RankMap assigns it, agrees with its
peer on variable ownership, and exchanges peer blobs in rank order."""
import json
import os
import socket
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import bench
        from k8s_b200.cluster import RankMap
        from k8s_b200.syncps import MNIST_SHAPES, MNIST_VARS, SyncPSJob
        cfg = bench.tf_config_for_rank(world, rank)
        doc = json.loads(cfg)
        assert doc["task"] == ({"type": "master", "index": 0} if rank == 0 else {"type": "worker", "index": rank - 1})
        job = SyncPSJob(cfg, rank, world)              # host-only: no CUDA call is made
        rm = RankMap(doc["cluster"], world)
        assert job.local_replicas == rm.local_replicas(rank) == [rank]
        owners = []
        for name, shape in zip(MNIST_VARS, MNIST_SHAPES):
            v, task = job.register(name, shape)
            owners.append((task, job.lib.ps_gpu(job.h, task)))
        mine = {"rank": rank, "cluster": doc["cluster"], "owners": owners, "blob": bytes([rank]) * 8}
        got = [None] * world
        dist.all_gather_object(got, mine)
        assert [g["rank"] for g in got] == list(range(world))
        assert all(g["cluster"] == doc["cluster"] for g in got)       # same cluster spec on every rank
        assert all(g["owners"] == owners for g in got)                # ranks agree on who owns what
        owned = [i for i, (_, gpu) in enumerate(owners) if gpu == rank]
        all_owned = [None] * world
        dist.all_gather_object(all_owned, owned)
        assert sorted(sum(all_owned, [])) == list(range(8))           # every variable has exactly one owner
        q.put((rank, "ok", owned))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, "fail: %s\n%s" % (e, traceback.format_exc()), None))
    finally:
        dist.destroy_process_group()


def test_two_ranks_agree_on_cluster_and_ownership():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert [r[1] for r in res] == ["ok", "ok"], res
    assert res[0][2] == [0, 2, 4, 6] and res[1][2] == [1, 3, 5, 7]    # round robin over 2 PS = 2 GPUs


def test_bench_reference_arm_line():
    import bench
    a = bench.parse_args(["--impl", "reference", "--steps", "1", "--warmup", "1", "--batch", "8"])
    line = bench.run_reference(a)
    assert line["impl"] == "reference" and line["unit"] == "examples/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0
