"""Where a CTA of the conv2 3xTF32 forward / data-gradient kernel spends its time (development aid).

Needs the library built with the trace counters: make -C k8s_b200/csrc clean all EXTRA=-DTFG_TRACE
(rebuild without EXTRA afterwards).  Prints, as medians over the CTAs, the cycles each role waited on
each barrier and when it finished, counted from the CTA's first instruction.
"""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SLOTS = ["producer: wait slab_empty", "producer: wait b_empty", "producer: done at",
         "mma: prologue (entry -> slab wait)", "mma: wait slab_full", "mma: wait tmem_empty", "mma: wait b_full", "mma: done at",
         "epilogue: wait tmem_full", "epilogue: drain (ld + add + arrive)", "epilogue: last chunk full at", "epilogue: done at",
         "cta: done at"]


def main():
    import torch

    from k8s_b200 import lib as L
    from k8s_b200.syncps import SyncPSJob, synthetic_batch
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    cfg = json.dumps({"cluster": {"master": ["m-0:2222"]}, "task": {"type": "master", "index": 0}, "environment": "cloud"})
    j = SyncPSJob(cfg, 0, 1, cuda_device=0)
    j.build_mnist(batch, math="tf32x3")
    x, y = synthetic_batch(0, batch)
    dx, dy = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    for _ in range(5):
        j.forward_backward(0, dx.data_ptr(), dy.data_ptr())
    j.sync()
    raw = C.CDLL(L.LIB_PATH)
    raw.tfg_debug_c2_trace.argtypes = [C.c_int, C.c_void_p]
    for which, name in ((0, "forward"), (1, "data gradient")):
        buf = np.zeros((160, 16), dtype=np.int64)
        rc = raw.tfg_debug_c2_trace(which, buf.ctypes.data)
        assert rc == 0, rc
        live = buf[:min(batch, 148)]
        print("conv2 %s, %d CTAs (cycles: median / max over CTAs)" % (name, len(live)))
        for i, label in enumerate(SLOTS):
            print("   %-40s %8d %8d" % (label, int(np.median(live[:, i])), int(live[:, i].max())))
    j.close()


if __name__ == "__main__":
    main()
