"""Multi-GPU parity check, one process per GPU (run under torchrun on an N-GPU box):
N replicas + N PS shards of the MNIST CNN for 3 synchronous steps through the peer-memory
push/apply/pull kernel; every rank's parameters must be bit-identical across ranks and match
the oracle's step (fp32 and tf32x3 math: relative L2 1e-3 per step on the accumulated update)."""
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
    from k8s_b200.syncps import MNIST_VARS, SyncPSJob, mnist_init, synthetic_batch
    from oracle import model_ref as M

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    layout = sys.argv[1] if len(sys.argv) > 1 else "balanced"
    math = sys.argv[2] if len(sys.argv) > 2 else "tf32x3"
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    batch = 16
    job = SyncPSJob(bench.tf_config_for_rank(world, rank), rank, world, cuda_device=local)
    job.build_mnist(batch, math=math, layout=layout)
    job.optimizer("momentum", 0.001, 0.9)
    job.connect_peers(dist)
    params = mnist_init(0)
    init = [p.copy() for p in params]
    accum = [np.zeros_like(p) for p in params]
    ok = True
    for step in range(3):
        batches = [synthetic_batch(r, batch, step) for r in range(world)]
        loss = job.step([batches[rank][0]], [batches[rank][1]])
        loss_ref, _ = M.sync_step(params, accum, batches, "momentum", 0.001, 0.9, math=math)
        mine = np.concatenate([job.get_var(v) for v in range(8)])
        gathered = [None] * world
        dist.all_gather_object(gathered, mine.tobytes())
        same = all(g == gathered[0] for g in gathered)
        ref = np.concatenate([p.ravel() for p in params])
        ini = np.concatenate([p.ravel() for p in init])
        upd_err = np.linalg.norm((mine - ini) - (ref - ini)) / np.linalg.norm(ref - ini)
        tol = 1e-3 if math in ("fp32", "tf32x3") else 5e-2
        good = same and upd_err <= tol * (step + 1) and abs(loss - loss_ref) <= 5e-3 * abs(loss_ref)
        ok = ok and good
        if rank == 0:
            print("step %d loss %.6f (oracle %.6f) ranks_bit_identical=%s update_rel_l2=%.3e -> %s"
                  % (step, loss, loss_ref, same, upd_err, "ok" if good else "FAIL"), flush=True)
    job.close()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("MGPU_CHECK %s world=%d layout=%s math=%s" % ("PASS" if ok else "FAIL", world, layout, math), flush=True)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
