"""Per-tensor error of one forward/backward of the sample model in a given math mode against the
fp32 PyTorch-CPU reference (development aid: prints every tensor instead of stopping at the first)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch

    from k8s_b200.syncps import MNIST_VARS, SyncPSJob, mnist_init, synthetic_batch
    from oracle import model_ref as M
    math = sys.argv[1] if len(sys.argv) > 1 else "tf32x3"
    cfg = json.dumps({"cluster": {"master": ["m-0:2222"]}, "task": {"type": "master", "index": 0}, "environment": "cloud"})
    for batch in [int(b) for b in (sys.argv[2:] or ["1", "8", "128"])]:
        j = SyncPSJob(cfg, 0, 1, cuda_device=0)
        j.build_mnist(batch, math=math)
        x, y = synthetic_batch(0, batch)
        dx, dy = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
        j.forward_backward(0, dx.data_ptr(), dy.data_ptr())
        j.sync()
        loss32, grads32, logits32 = M.loss_and_grads(mnist_init(0), x, y)
        lg = j.logits(0)
        print("batch %d math %s: logits err/scale %.3e  loss %.7f (ref %.7f)" %
              (batch, math, np.abs(lg - logits32).max() / np.abs(logits32).max(), j.loss(0), loss32), flush=True)
        for v, name in enumerate(MNIST_VARS):
            g = j.get_grad(0, v)
            r = grads32[v].ravel()
            print("   grad %-8s max err/scale %.3e  rel L2 %.3e" %
                  (name, np.abs(g - r).max() / np.abs(r).max(), np.linalg.norm(g - r) / np.linalg.norm(r)), flush=True)
        j.close()


if __name__ == "__main__":
    main()
