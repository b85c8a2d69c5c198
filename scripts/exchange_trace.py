"""Where the in-kernel exchange of a multi-GPU step spends its time (development aid; run under torchrun).

Needs the library built with the trace stamps: make -C k8s_b200/csrc clean all EXTRA=-DTFG_TRACE (rebuild
without EXTRA afterwards).  Runs the sample model's step from device buffers (graph replays, batch 128 per
replica), then prints for every rank the %globaltimer stamps of the LAST exchange launch as phase lengths.
"""
import ctypes as C
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
    from k8s_b200.syncps import SyncPSJob, synthetic_batch

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stream = torch.cuda.Stream()
    job = SyncPSJob(bench.tf_config_for_rank(world, rank), rank, world, cuda_device=local, stream=stream.cuda_stream)
    job.build_mnist(128, math="tf32x3", layout="balanced")
    job.optimizer("momentum", 0.001, 0.9)
    job.connect_peers(dist)
    x, y = synthetic_batch(rank, 128)
    dx, dy = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    torch.cuda.synchronize()
    dist.barrier()
    for _ in range(40):
        job.step_dev([dx.data_ptr()], [dy.data_ptr()])
    job.sync()
    raw = C.CDLL(L.LIB_PATH)
    raw.tfg_debug_xchg_trace.argtypes = [C.c_void_p]
    buf = np.zeros(8, dtype=np.uint64)
    assert raw.tfg_debug_xchg_trace(buf.ctypes.data) == 0
    t = buf.astype(np.int64)
    line = ("rank %d: ready wait %.1f us | CTA 0 chunks %.1f | CTA 0 flush %.1f | last CTA through at %.1f | "
            "pulled wait %.1f | total %.1f us" % (rank, (t[1] - t[0]) / 1e3, (t[2] - t[1]) / 1e3, (t[3] - t[2]) / 1e3,
                                                   (t[4] - t[0]) / 1e3, (t[5] - t[4]) / 1e3, (t[5] - t[0]) / 1e3))
    lines = [None] * world
    dist.all_gather_object(lines, line)
    if rank == 0:
        print("\n".join(lines), flush=True)
    job.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
