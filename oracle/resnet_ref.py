"""ORACLE -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.  PARITY UNPINNED (see oracle.c): the
reference tree holds no model; ResNet-50 is build-defined (v1.5, stride on the 3x3).

PyTorch-CPU reference of the ResNet-50 forward/backward with the arithmetic of the GPU
path restated: activations, operand weights and activation gradients are rounded to bf16
at tensor boundaries (`_Round`), contractions accumulate in fp32, batch norm uses batch
statistics (biased variance, eps 1e-5) in fp32.  `emulate=False` gives plain fp32.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F


class _Round(torch.autograd.Function):
    """bf16 rounding of a tensor on the way forward and of its gradient on the way back."""

    @staticmethod
    def forward(ctx, x):
        return x.bfloat16().float()

    @staticmethod
    def backward(ctx, g):
        return g.bfloat16().float()


def _rnd(x, on):
    return _Round.apply(x) if on else x


def units() -> List[Tuple[str, str, int, int, int, bool, str]]:
    """(name, input, R, stride, kout, relu, residual) in creation order; tensors named like the
    product's ("conv1/y", "pool1", "stage1/block1/c/y")."""
    out = [("conv1", "image", 7, 2, 64, True, "")]
    cur = "pool1"
    for st, (nb, mid) in enumerate(zip([3, 4, 6, 3], [64, 128, 256, 512])):
        for b in range(nb):
            base = "stage%d/block%d" % (st + 1, b + 1)
            stride = 2 if (b == 0 and st > 0) else 1
            sc = cur
            if b == 0:
                out.append((base + "/shortcut", cur, 1, stride, 4 * mid, False, ""))
                sc = base + "/shortcut/y"
            out.append((base + "/a", cur, 1, 1, mid, True, ""))
            out.append((base + "/b", base + "/a/y", 3, stride, mid, True, ""))
            out.append((base + "/c", base + "/b/y", 1, 1, 4 * mid, True, sc))
            cur = base + "/c/y"
    return out


def forward_backward(params: Dict[str, np.ndarray], x: np.ndarray, y: np.ndarray, emulate: bool = True,
                     keep: Sequence[str] = ()):
    """params: {variable name: array} (HWIO filters).  Returns (loss, {name: grad}, logits,
    {tensor name: NHWC activation} for names in `keep`)."""
    P = {k: torch.tensor(v, dtype=torch.float32, requires_grad=True) for k, v in params.items()}
    T: Dict[str, torch.Tensor] = {}
    T["image"] = _rnd(torch.tensor(x, dtype=torch.float32).permute(0, 3, 1, 2), emulate)
    for name, inp, R, stride, kout, relu, res in units():
        w = _rnd(P[name + "/weights"], emulate).permute(3, 2, 0, 1)
        z = _rnd(F.conv2d(T[inp], w, None, stride=stride, padding=R // 2), emulate)
        T[name + "/z"] = z
        mean = z.mean(dim=(0, 2, 3), keepdim=True)
        var = (z * z).mean(dim=(0, 2, 3), keepdim=True) - mean * mean
        invstd = torch.rsqrt(var.clamp_min(0) + 1e-5)
        v = P[name + "/bn/gamma"].view(1, -1, 1, 1) * (z - mean) * invstd + P[name + "/bn/beta"].view(1, -1, 1, 1)
        if res:
            v = v + T[res]
        if relu:
            v = F.relu(v)
        T[name + "/y"] = _rnd(v, emulate)
        if name == "conv1":
            T["pool1"] = F.max_pool2d(T["conv1/y"], 3, 2, 1)
    last = T["stage4/block3/c/y"]
    pooled = _rnd(last.mean(dim=(2, 3)), emulate)
    logits = pooled @ _rnd(P["fc/weights"], emulate) + P["fc/biases"]
    loss = F.cross_entropy(logits, torch.tensor(y, dtype=torch.long))
    names = list(P)
    grads = torch.autograd.grad(loss, [P[n] for n in names], allow_unused=True)
    g = {n: (gr.numpy() if gr is not None else np.zeros_like(params[n])) for n, gr in zip(names, grads)}
    acts = {k: T[k].detach().permute(0, 2, 3, 1).contiguous().numpy() for k in keep}
    return float(loss.detach()), g, logits.detach().numpy(), acts
