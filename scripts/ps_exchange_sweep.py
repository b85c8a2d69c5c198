"""BASELINE config 5 across GPUs: the dense PS exchange (gradient push + mean + optimizer apply +
parameter pull) on N ranks, one process per GPU (run under torchrun on an N-GPU box):

    torchrun --nproc-per-node N scripts/ps_exchange_sweep.py [--quick] [--out gpurun_out/x.jsonl]

For P in {1M .. 1B} fp32 parameters in S in {1,2,4,8,16} logical PS shards (S variables of P/S
elements, round-robin placed on the PS tasks, PS task k owned by GPU k % N) and for the balanced
physical layout, two arms on the same buffers:

  peer : ONE launch per rank of the fused peer-memory kernel (tfg_push_apply_pull, both cross-rank
         synchronisation points inside the kernel);
  nccl : ncclReduceScatter(sum) of the flat gradient -> tfg_apply_owned (the same apply arithmetic on
         the owned range) -> ncclAllGather of the parameters (torch.distributed on the job's stream):
         north-star subsystems (4) and (5) as literally described.  Balanced layout only (NCCL's
         collectives split the buffer evenly).

Time = CUDA events on the job's stream around each step, median of 10 after 3 warm-up steps, MAX over
ranks.  NVLink bytes per direction of the busiest GPU: an owner reads its range from the N-1 other
replicas and writes the new values to the N-1 other ranks, 4 B x owned x (N-1) each way.  `link_frac`
is against the pool's measured peer-copy bandwidth (770 GB/s per direction per GPU, B200_PROFILING.md;
900 nominal); `hbm_frac` (N = 1, or when HBM is the tighter bound) against MEASURED_PEAKS.json.
Both arms leave bit-identical parameters on every rank (checked once per configuration at the
smallest size: a sum of N fp32 values in rank order vs NCCL's order may differ in the last bit, so
the cross-arm check is 1e-6 relative; within an arm all ranks must agree bit for bit).
"""
import ctypes as C
import json
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
from k8s_b200 import lib as L  # noqa: E402
from k8s_b200.syncps import SyncPSJob  # noqa: E402

LINK_GBS = 770.0


def view(ptr, n):
    class _V:
        __cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 3, "strides": None}
    return torch.as_tensor(_V(), device="cuda")


def main():
    quick = "--quick" in sys.argv
    out_path = sys.argv[sys.argv.index("--out") + 1] if "--out" in sys.argv else None
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stream = torch.cuda.Stream()
    peaks = bench.measured_peaks()
    rows = []

    def timed(fn, reps=10, warm=3):
        ts = []
        for i in range(warm + reps):
            if world > 1:
                dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            fn()
            e1.record(stream)
            stream.synchronize()
            if i >= warm:
                ts.append(e0.elapsed_time(e1))
        t = torch.tensor([float(np.median(ts))], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    sizes = [1 << 20, 1 << 24, 1 << 28] if quick else [1 << 20, 1 << 22, 1 << 24, 1 << 26, 1 << 28, 1 << 30]
    shard_counts = [1, 8, 16] if quick else [1, 2, 4, 8, 16]
    for P in sizes:
        configs = [("logical", S) for S in shard_counts] + [("balanced", max(world, 1))]
        for layout, S in configs:
            for opt in (("momentum",) if quick else ("sgd", "momentum")):
                cluster = {"master": ["m:1"], "ps": ["p%d:1" % i for i in range(S)]}
                if world > 1:
                    cluster["worker"] = ["w%d:1" % i for i in range(world - 1)]
                job = SyncPSJob(json.dumps({"cluster": cluster}), rank, world, cuda_device=local, stream=stream.cuda_stream)
                per = (P // S) // (32 * world) * (32 * world)            # every variable splits evenly and stays 128-B aligned
                for i in range(S):
                    job.register("shard%d" % i, (per,))
                job.finalize(layout)
                job.optimizer(opt, 0.01, 0.9)
                n = job.flat_count
                gptr, _ = job.buffer(L.BUF_GRADS, 0)
                pptr, _ = job.buffer(L.BUF_PARAMS, 0)
                aptr, _ = job.buffer(L.BUF_ACCUM, 0)
                grads, params, accum = view(gptr, n), view(pptr, n), view(aptr, n)
                with torch.cuda.stream(stream):
                    gen = torch.Generator(device="cuda").manual_seed(1000 + rank)
                    grads.normal_(generator=gen)
                    params.fill_(1.0)
                stream.synchronize()
                job.connect_peers(dist if world > 1 else None)
                owned = sum(c for _, c in job.owned_ranges())
                t_owned = torch.tensor([owned], device="cuda", dtype=torch.int64)
                if world > 1:
                    dist.all_reduce(t_owned, op=dist.ReduceOp.MAX)
                max_owned = int(t_owned.item())
                arms = [("peer", lambda: job.push_apply_pull(True))]
                if world > 1 and layout == "balanced":
                    g0 = grads.clone()
                    part = n // world
                    g_mine, p_mine = grads[rank * part:(rank + 1) * part], params[rank * part:(rank + 1) * part]

                    def nccl_step():
                        with torch.cuda.stream(stream):
                            dist.reduce_scatter_tensor(g_mine, grads)          # in place: the sum lands in the owner's slice
                            job.apply_owned(world)
                            dist.all_gather_into_tensor(params, p_mine)
                    arms.append(("nccl", nccl_step))
                crcs = {}
                for arm, fn in arms:
                    with torch.cuda.stream(stream):                           # same start for both arms
                        params.fill_(1.0)
                        accum.zero_()
                        if arm == "nccl":
                            grads.copy_(g0)
                    stream.synchronize()
                    if P <= 1 << 22:      # correctness at the small sizes: one step, all ranks identical
                        if world > 1:
                            dist.barrier()
                        fn()
                        stream.synchronize()
                        host = params.cpu().numpy()
                        crc = zlib.crc32(host.tobytes())
                        all_crc = [crc]
                        if world > 1:
                            all_crc = [None] * world
                            dist.all_gather_object(all_crc, crc)
                        crcs[arm] = (all(c == all_crc[0] for c in all_crc), host)
                        with torch.cuda.stream(stream):
                            params.fill_(1.0)
                            accum.zero_()
                            if arm == "nccl":
                                grads.copy_(g0)
                        stream.synchronize()
                    ms = timed(fn)    # nccl arm: the in-place reduce-scatter re-sums its own output; values drift, timing does not
                    bpp = 12 if opt == "sgd" else 20
                    link_bytes = 4.0 * max_owned * (world - 1) + 4.0 * (n - max_owned)
                    row = {"what": "ps_exchange", "n_gpus": world, "arm": arm, "params": n, "shards": S, "layout": layout,
                           "optimizer": opt, "ms": ms, "owned_max": max_owned,
                           "hbm_gbs": bpp * max_owned / (ms * 1e-3) / 1e9 if world == 1 else (4.0 * world + bpp - 4) * max_owned / (ms * 1e-3) / 1e9,
                           "link_gbs_per_dir": link_bytes / (ms * 1e-3) / 1e9 if world > 1 else None,
                           "link_frac": link_bytes / (ms * 1e-3) / 1e9 / LINK_GBS if world > 1 else None,
                           "hbm_frac": (bpp * max_owned / (ms * 1e-3) / 1e9) / peaks["hbm_gbs"] if world == 1 else None}
                    if arm in crcs:
                        row["ranks_bit_identical"] = crcs[arm][0]
                    if arm == "nccl" and "peer" in crcs and "nccl" in crcs:
                        a, b = crcs["peer"][1], crcs["nccl"][1]
                        row["max_rel_diff_vs_peer"] = float(np.abs(a - b).max() / np.abs(a).max())
                    rows.append(row)
                    if rank == 0:
                        print(json.dumps(row), flush=True)
                job.close()
                del grads, params, accum
                torch.cuda.empty_cache()
    if rank == 0 and out_path:
        with open(out_path, "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
