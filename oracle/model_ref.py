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


def quantizer(math: str):
    """Operand rounding of the tensor-path contractions: bf16 = round-to-nearest-even to 8
    mantissa bits; tf32 = the low 13 mantissa bits of an fp32 operand are not read."""
    if math == "bf16":
        return lambda t: t.bfloat16().float()
    if math == "tf32":
        return lambda t: (t.contiguous().view(torch.int32) & -8192).view(torch.float32)
    return None


class _QMatmul(torch.autograd.Function):
    """a @ w with both operands (and, in backward, the incoming gradient) rounded to the
    operand precision, products and sums in fp32 -- what a tcgen05 GEMM with fp32
    accumulation computes for fwd, dgrad and wgrad."""

    @staticmethod
    def forward(ctx, a, w, q):
        qa, qw = q(a), q(w)
        ctx.save_for_backward(qa, qw)
        ctx.q = q
        return qa @ qw

    @staticmethod
    def backward(ctx, g):
        qa, qw = ctx.saved_tensors
        qg = ctx.q(g)
        return qg @ qw.t(), qa.t() @ qg, None


def forward_q(params: Sequence[torch.Tensor], x: torch.Tensor, q) -> torch.Tensor:
    """The same network with conv2 and fc1 computed as operand-rounded GEMMs (conv2 through
    an explicit im2col), conv1 and fc2 in fp32 -- the arithmetic of math modes bf16/tf32."""
    w1, b1, w2, b2, w3, b3, w4, b4 = params
    B = x.shape[0]
    h = F.conv2d(x.permute(0, 3, 1, 2), w1.permute(3, 2, 0, 1), b1, padding=2)
    h = F.max_pool2d(F.relu(h), 2)                                   # [B,32,14,14]
    col = F.unfold(h, kernel_size=5, padding=2)                      # [B, 32*25, 196], k = (c, kh, kw)
    col = col.permute(0, 2, 1).reshape(B * 196, 800)
    w2m = w2.permute(2, 0, 1, 3).reshape(800, 64)                    # HWIO -> (c, kh, kw) x O
    y2 = _QMatmul.apply(col, w2m, q) + b2
    h = y2.reshape(B, 14, 14, 64).permute(0, 3, 1, 2)
    h = F.max_pool2d(F.relu(h), 2)
    h = h.permute(0, 2, 3, 1).reshape(B, 3136)
    h = F.relu(_QMatmul.apply(h, w3, q) + b3)
    return h @ w4 + b4


def loss_and_grads(params_np: Sequence[np.ndarray], x: np.ndarray, y: np.ndarray, dtype=torch.float32,
                   math: str = "fp32"):
    ps = [torch.tensor(p, dtype=dtype, requires_grad=True) for p in params_np]
    q = quantizer(math)
    xt = torch.tensor(x, dtype=dtype)
    logits = forward(ps, xt) if q is None else forward_q(ps, xt, q)
    loss = F.cross_entropy(logits, torch.tensor(y, dtype=torch.long))
    grads = torch.autograd.grad(loss, ps)
    return float(loss.detach()), [g.numpy() for g in grads], logits.detach().numpy()


def sync_step(params_np: List[np.ndarray], accum_np, batches, kind: str = "sgd", lr: float = 0.01,
              mu: float = 0.9, dtype=torch.float32, math: str = "fp32"):
    """One synchronous step over ``batches`` = [(x_r, y_r) per replica].  Updates params (and
    accum) in place through the C oracle's apply, returns (mean loss, per-replica grads)."""
    losses, all_grads = [], []
    for x, y in batches:
        l, g, _ = loss_and_grads(params_np, x, y, dtype, math)
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
