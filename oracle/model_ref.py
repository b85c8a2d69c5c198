"""ORACLE -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.  PARITY UNPINNED (see oracle.c).

Floating-point reference of the sample model's sync-PS step in plain PyTorch on the CPU
(fp32 for like-for-like, fp64 for error bounds).  The model is build-defined -- the
reference's tf_sample image holds only the smoke multiply (SURVEY.md par. 0.2) -- with the
canonical mnist_deep shapes; the step semantics are TF 1.3's between-graph synchronous
replication (SURVEY.md par. 3.5): every replica computes gradients of its mean batch loss,
the accumulator takes their mean, the optimizer applies it once.
"""
from __future__ import annotations

from typing import List, Sequence

import numpy as np
import torch
import torch.nn.functional as F

from . import oracle as O


def forward(params: Sequence[torch.Tensor], x: torch.Tensor) -> torch.Tensor:
    """x [B,28,28,1] NHWC -> logits [B,10].  Filters are HWIO, the flatten is NHWC order."""
    w1, b1, w2, b2, w3, b3, w4, b4 = params
    h = x.permute(0, 3, 1, 2)
    h = F.conv2d(h, w1.permute(3, 2, 0, 1), b1, padding=2)
    h = F.max_pool2d(F.relu(h), 2)
    h = F.conv2d(h, w2.permute(3, 2, 0, 1), b2, padding=2)
    h = F.max_pool2d(F.relu(h), 2)
    h = h.permute(0, 2, 3, 1).reshape(x.shape[0], 3136)
    h = F.relu(h @ w3 + b3)
    return h @ w4 + b4


def loss_and_grads(params_np: Sequence[np.ndarray], x: np.ndarray, y: np.ndarray, dtype=torch.float32):
    ps = [torch.tensor(p, dtype=dtype, requires_grad=True) for p in params_np]
    logits = forward(ps, torch.tensor(x, dtype=dtype))
    loss = F.cross_entropy(logits, torch.tensor(y, dtype=torch.long))
    grads = torch.autograd.grad(loss, ps)
    return float(loss.detach()), [g.numpy() for g in grads], logits.detach().numpy()


def sync_step(params_np: List[np.ndarray], accum_np, batches, kind: str = "sgd", lr: float = 0.01,
              mu: float = 0.9, dtype=torch.float32):
    """One synchronous step over ``batches`` = [(x_r, y_r) per replica].  Updates params (and
    accum) in place through the C oracle's apply, returns (mean loss, per-replica grads)."""
    losses, all_grads = [], []
    for x, y in batches:
        l, g, _ = loss_and_grads(params_np, x, y, dtype)
        losses.append(np.float32(l))
        all_grads.append(g)
    n = len(batches)
    for v in range(len(params_np)):
        gs = [np.ascontiguousarray(all_grads[r][v], dtype=np.float32).ravel() for r in range(n)]
        flat = params_np[v].reshape(-1)
        acc = None if accum_np is None else accum_np[v].reshape(-1)
        O.dense_apply(O.KINDS[kind], flat, acc, gs, float(n), lr, mu)
    s = np.float32(losses[0])
    for l in losses[1:]:
        s = np.float32(s + l)
    return float(np.float32(s / np.float32(n))), all_grads
