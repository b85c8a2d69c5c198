"""A small run of every hand-written kernel family, for compute-sanitizer (one tool per call):

    compute-sanitizer --tool memcheck  python scripts/sanitize_case.py
    compute-sanitizer --tool racecheck python scripts/sanitize_case.py nontma

One tf32x3 step and one fp32 step of the sample model (batch 8, two replicas + two PS shards on one
rank, eager launches), the dense exchange on two emulated ranks, sparse gather / scatter-sub and the
smoke multiply.  `nontma` leaves out the kernels that synchronise through mbarriers and the async
proxy (tcgen05 / TMA), whose shared-memory traffic racecheck cannot follow."""
import json
import os
import sys

import numpy as np

os.environ["TFGRAFT_GRAPH"] = "0"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch

    from k8s_b200 import lib as L
    from k8s_b200.sparse import ShardedEmbedding
    from k8s_b200.syncps import SyncPSJob, synthetic_batch
    nontma = "nontma" in sys.argv
    lib = L.load()
    cfg = json.dumps({"cluster": {"master": ["m:1"], "worker": ["w:1"], "ps": ["p0:1", "p1:1"]}})
    for math in (["fp32"] if nontma else ["tf32x3", "fp32", "bf16"]):
        j = SyncPSJob(cfg, 0, 1, cuda_device=0)
        j.build_mnist(8, math=math)
        j.optimizer("momentum", 0.001, 0.9)
        b = [synthetic_batch(r, 8) for r in range(2)]
        print(math, "loss", j.step([x for x, _ in b], [y for _, y in b]), flush=True)
        j.close()
    # dense exchange, two emulated ranks (no cross-rank waiting inside the kernels)
    jobs = [SyncPSJob(cfg, r, 2, cuda_device=0) for r in range(2)]
    for j in jobs:
        j.register("w", (70001,))
        j.finalize("balanced")
        j.optimizer("nesterov", 0.05, 0.9)
        j.set_var(0, np.ones(70001, np.float32))
        j.set_grad(0, 0, np.full(70001, 0.5, np.float32))
    jobs[0].attach_local_peer(1, jobs[1])
    jobs[1].attach_local_peer(0, jobs[0])
    if not nontma:     # multi-rank jobs stream through the bulk-copy kernel
        for j in jobs:
            j.push_apply_pull(with_barriers=False)
        for j in jobs:
            j.sync()
        print("dense exchange", float(jobs[1].get_var(0)[0]), flush=True)
    # sparse path, two emulated ranks
    embs = [ShardedEmbedding(j, 1000, 33, 512, "div", n_shards=3) for j in jobs]
    rng = np.random.default_rng(0)
    for e in embs:
        e.attach_local([o for o in embs if o is not e])
        e.stage_ids(np.minimum(rng.zipf(1.05, 512) - 1, 999))
        e.stage_rows(rng.standard_normal((512, 33)).astype(np.float32))
    torch.cuda.synchronize()
    for e in embs:
        e.lookup()
        e.apply_owned(0.01)
    torch.cuda.synchronize()
    print("sparse ok", flush=True)
    print("smoke", int(jobs[0].smoke(all_tasks=True).sum()), flush=True)
    n = 100003
    v, g = torch.ones(n, device="cuda"), torch.ones(n, device="cuda")
    lib.check(lib.dense_apply(L.OPT_SGD, v.data_ptr(), None, L.ptr_array([g]), 1, 1.0, n, 0.01, 0.0, None, 0, 0))
    torch.cuda.synchronize()
    for e in embs:
        e.close()
    for j in jobs:
        j.close()
    print("SANITIZE_CASE DONE", flush=True)


if __name__ == "__main__":
    main()
