"""Multi-GPU parity of the sparse embedding path, one process per GPU (run under torchrun): every
rank looks up and updates a table sharded over all ranks through CUDA-IPC peer mappings and the
stream-ordered barrier; the looked-up rows and the final shards must be bit-identical to the C oracle
run on the concatenated batch.  Also times lookup and update at BASELINE config 5's sizes
(--bench): GB/s against the algorithmic bytes 8 + 8 D (gather) and 8 + 12 D (scatter-sub) per id."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    import bench
    from k8s_b200 import lib as L
    from k8s_b200.sparse import ShardedEmbedding
    from k8s_b200.syncps import SyncPSJob
    from oracle import oracle as O

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    do_bench = "--bench" in sys.argv
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stream = torch.cuda.Stream()
    job = SyncPSJob(bench.tf_config_for_rank(world, rank), rank, world, cuda_device=local, stream=stream.cuda_stream)
    job.register("dummy", (32,))
    job.finalize("balanced")
    job.connect_peers(dist)
    ok = True
    for strategy in ("mod", "div"):
        vocab, dim, batch = 50021, 128, 4096
        strat = {"mod": L.PART_MOD, "div": L.PART_DIV}[strategy]
        with torch.cuda.stream(stream):
            emb = ShardedEmbedding(job, vocab, dim, batch, strategy, n_shards=world, stream=stream)
        rng = np.random.default_rng(7)
        tables = {}
        for k in range(world):                         # every rank draws the same tables; owners upload theirs
            rows = emb.lib.sparse_shard_rows(vocab, world, k, strat)
            tables[k] = rng.standard_normal((max(rows, 1), dim)).astype(np.float32)
        with torch.cuda.stream(stream):
            for k in emb.owned:
                emb.set_shard(k, tables[k])
        stream.synchronize()
        emb.connect(dist)
        for step in range(2):
            ids = [np.minimum(np.random.default_rng([step, r]).zipf(1.05, size=batch) - 1, vocab - 1).astype(np.int64)
                   for r in range(world)]
            grads = [np.random.default_rng([step, r, 1]).standard_normal((batch, dim)).astype(np.float32) for r in range(world)]
            with torch.cuda.stream(stream):
                emb.stage_ids(ids[rank])
                job.barrier()                          # tables from the previous update are complete everywhere
                got = emb.lookup().cpu().numpy()
            want = np.full((batch, dim), np.nan, dtype=np.float32)
            for k in range(world):
                O.sparse_gather(tables[k], vocab, world, k, strat, ids[rank], want)
            good = np.array_equal(got.view(np.uint32), want.view(np.uint32))
            with torch.cuda.stream(stream):
                emb.apply_gradients(grads[rank], 0.05)
            stream.synchronize()
            all_ids, all_grads = np.concatenate(ids), np.concatenate(grads)
            for k in range(world):
                O.sparse_scatter_sub(tables[k], vocab, world, k, strat, all_ids, all_grads, 0.05)
            for k in emb.owned:
                good = good and np.array_equal(emb.get_shard(k).view(np.uint32), tables[k].view(np.uint32))
            flags = [None] * world
            dist.all_gather_object(flags, bool(good))
            ok = ok and all(flags)
            if rank == 0:
                print("sparse %s step %d: lookup and shards bit-exact on all ranks = %s" % (strategy, step, all(flags)), flush=True)
        dist.barrier()
        emb.close()
    if do_bench:
        peak = bench.measured_peaks()["hbm_gbs"]
        for vocab, dim, batch in [(1 << 20, 128, 1 << 16), (1 << 23, 128, 1 << 20), (1 << 23, 64, 1 << 20)]:
            with torch.cuda.stream(stream):
                emb = ShardedEmbedding(job, vocab, dim, batch, "mod", n_shards=world, stream=stream, chunk_long_runs=True)
                for k in emb.owned:
                    emb.shards[k].normal_()
            emb.connect(dist)
            for name in ("uniform", "zipf"):
                rng = np.random.default_rng([rank, 5])
                ids = rng.integers(0, vocab, size=batch) if name == "uniform" else np.minimum(rng.zipf(1.05, size=batch) - 1, vocab - 1)
                with torch.cuda.stream(stream):
                    emb.stage_ids(ids.astype(np.int64))
                    emb.rows.normal_()
                    job.barrier()

                def timed(fn, reps=8, warm=2):
                    ts = []
                    for i in range(warm + reps):
                        dist.barrier()
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        with torch.cuda.stream(stream):
                            e0.record(stream)
                            fn()
                            e1.record(stream)
                        stream.synchronize()
                        if i >= warm:
                            ts.append(e0.elapsed_time(e1))
                    t = torch.tensor([float(np.median(ts))], device="cuda", dtype=torch.float64)
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                    return float(t.item())

                def update():
                    job.barrier()
                    emb.apply_owned(0.01)
                    job.barrier()
                ms_g, ms_s = timed(emb.lookup), timed(update)
                if rank == 0:
                    n_all = batch * world
                    print(json.dumps({"what": "sparse_exchange", "n_gpus": world, "vocab": vocab, "dim": dim,
                                      "ids_per_rank": batch, "dist": name,
                                      "gather_ms": ms_g, "gather_gbs_per_gpu": batch * (8 + 8 * dim) / ms_g / 1e6,
                                      "gather_frac_hbm": batch * (8 + 8 * dim) / ms_g / 1e6 / peak,
                                      "scatter_ms": ms_s, "scatter_gbs_aggregate": n_all * (8 + 12 * dim) / ms_s / 1e6,
                                      "scatter_gbs_per_gpu": n_all * (8 + 12 * dim) / ms_s / 1e6 / world}), flush=True)
            dist.barrier()
            emb.close()
            del emb
            torch.cuda.empty_cache()
    job.close()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("MGPU_SPARSE_CHECK %s world=%d" % ("PASS" if ok else "FAIL", world), flush=True)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
