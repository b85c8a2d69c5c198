"""BASELINE config 5: PS variable-shard sweep on one GPU (run on the GPU box).

  dense : P in {1M .. 1B} fp32 parameters in S in {1,2,4,8,16} logical PS shards (S variables
          of P/S elements, round-robin placed), one fused push+apply+pull launch per step.
          GB/s = algorithmic bytes (12 B/param SGD, 20 B/param momentum) / CUDA-event time.
  sparse: table V x D fp32, B int64 ids (uniform and Zipf(1.05)); gather 8+8D B/id,
          scatter-sub 8+12D B/id.
Prints one JSON object per line; `frac` is against MEASURED_PEAKS.json's HBM copy bandwidth.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench  # noqa: E402
from k8s_b200 import lib as L  # noqa: E402
from k8s_b200.syncps import SyncPSJob  # noqa: E402


def timed(fn, stream, reps=10, warm=3):
    ts = []
    for i in range(warm + reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        fn()
        e1.record(stream)
        stream.synchronize()
        if i >= warm:
            ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def main():
    quick = "--quick" in sys.argv
    sparse_only = "--sparse-only" in sys.argv      # skip the dense sweep
    one = "--one" in sys.argv or "--one-zipf" in sys.argv   # sparse: V = 1 M, D = 128, 1 M ids of one distribution (for ncu)
    lib = L.load()
    peak = bench.measured_peaks()["hbm_gbs"]
    stream = torch.cuda.Stream()
    out = []
    # ---- dense ---------------------------------------------------------------------------
    sizes = [1 << 20, 1 << 22, 1 << 24, 1 << 26, 1 << 28] + ([] if quick else [1 << 30])
    for P in ([] if sparse_only or one else sizes):
        for S in ([1, 8] if quick else [1, 2, 4, 8, 16]):
            for opt in ("sgd", "momentum"):
                cfg = json.dumps({"cluster": {"master": ["m:1"], "ps": ["p%d:1" % i for i in range(S)]}})
                job = SyncPSJob(cfg, 0, 1, cuda_device=0, stream=stream.cuda_stream)
                for i in range(S):
                    job.register("shard%d" % i, (P // S,))
                job.finalize("logical")
                job.optimizer(opt, 0.01, 0.9)
                gptr, gbytes = job.buffer(L.BUF_GRADS, 0)
                with torch.cuda.stream(stream):   # N(0,1) gradients straight into the job's buffer
                    g = torch.randn(P, device="cuda")
                    torch.cuda.synchronize()
                C = __import__("ctypes")
                C.CDLL("libcudart.so.12").cudaMemcpy(C.c_void_p(gptr), C.c_void_p(g.data_ptr()),
                                                     C.c_size_t(4 * P), C.c_int(3))
                del g
                ms = timed(lambda: job.push_apply_pull(False), stream)
                bpp = 12 if opt == "sgd" else 20
                gbs = bpp * P / (ms * 1e-3) / 1e9
                out.append({"op": "dense_update", "params": P, "ps_shards": S, "optimizer": opt, "ms": ms,
                            "bytes_per_param": bpp, "GB/s": gbs, "frac_of_hbm_peak": gbs / peak})
                print(json.dumps(out[-1]), flush=True)
                job.close()
                torch.cuda.empty_cache()
    # ---- sparse --------------------------------------------------------------------------
    rng = np.random.default_rng(0)
    for V, D in ([(1 << 20, 128)] if quick or one else [(1 << 20, 64), (1 << 20, 128), (1 << 23, 64), (1 << 23, 128)]):
        table = torch.randn(V, D, device="cuda")
        for B in ([1 << 20] if one else [1 << 16] if quick else [1 << 16, 1 << 20]):
            for dist in (("zipf",) if "--one-zipf" in sys.argv else ("uniform",) if one else ("uniform", "zipf")):
                ids_np = (rng.integers(0, V, size=B) if dist == "uniform"
                          else np.minimum(rng.zipf(1.05, size=B) - 1, V - 1)).astype(np.int64)
                ids = torch.from_numpy(ids_np).cuda()
                rows = torch.randn(B, D, device="cuda")
                outb = torch.empty(B, D, device="cuda")
                scratch = torch.empty(lib.sparse_scratch_bytes_fast(B, D), dtype=torch.uint8, device="cuda")
                s = stream.cuda_stream
                ms_g = timed(lambda: lib.check(lib.sparse_gather(table.data_ptr(), V, D, 1, 0, L.PART_MOD, ids.data_ptr(),
                                                                 B, outb.data_ptr(), None, s)), stream)
                ms_s = timed(lambda: lib.check(lib.sparse_scatter_sub(table.data_ptr(), V, D, 1, 0, L.PART_MOD,
                                                                      ids.data_ptr(), B, rows.data_ptr(), 0.01,
                                                                      scratch.data_ptr(), scratch.numel(), s)), stream)
                uniq = int(np.unique(ids_np).size)
                top = int(np.bincount(ids_np[ids_np < 1 << 24] if False else ids_np).max()) if B <= 1 << 20 else -1
                for op, ms, bpi in (("sparse_gather", ms_g, 8 + 8 * D), ("sparse_scatter_sub", ms_s, 8 + 12 * D)):
                    gbs = bpi * B / (ms * 1e-3) / 1e9
                    out.append({"op": op, "vocab": V, "dim": D, "ids": B, "dist": dist, "unique_ids": uniq,
                                "max_multiplicity": top, "ms": ms, "bytes_per_id": bpi, "GB/s": gbs,
                                "frac_of_hbm_peak": gbs / peak})
                    print(json.dumps(out[-1]), flush=True)
        del table
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
