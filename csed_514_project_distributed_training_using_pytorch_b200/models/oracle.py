"""Plain-PyTorch statement of the reference network's math: the numerics oracle.

This is the semantics of src/model.py:15-22 written functionally, with the two dropout
masks as explicit inputs so a CUDA kernel and the oracle can be fed the same masks
(SURVEY §4 item 2). It runs on CPU in fp32 or fp64 and is also the CPU execution path
of `models.Net` (BASELINE config 1: the single-process CPU loop).
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch
import torch.nn.functional as F

MEAN, STD = 0.1307, 0.3081  # src/train.py:29-31, src/train_dist.py:18


def normalize_u8(x_u8: torch.Tensor) -> torch.Tensor:
    """torchvision ToTensor + Normalize on a uint8 [B,784] or [B,28,28] batch (T12)."""
    x = x_u8.to(torch.float32).div(255.0).sub(MEAN).div(STD)
    return x.view(-1, 1, 28, 28)


def make_masks(batch: int, generator: Optional[torch.Generator] = None, p: float = 0.5,
               device="cpu", dtype=torch.float32):
    """Dropout2d channel mask [B,20] and element mask [B,50], values in {0, 1/(1-p)} (T9)."""
    keep = 1.0 - p
    m2 = (torch.rand(batch, 20, generator=generator) < keep).to(dtype) / keep
    m3 = (torch.rand(batch, 50, generator=generator) < keep).to(dtype) / keep
    return m2.to(device), m3.to(device)


def net_forward(params: Sequence[torch.Tensor], x: torch.Tensor, m2=None, m3=None):
    """log-probabilities [B,10]. params = (c1w, c1b, c2w, c2b, f1w, f1b, f2w, f2b).
    m2 / m3: None (no dropout), a mask tensor ([B,20] channel mask / [B,50] element mask,
    values 0 or 1/(1-p)), or a callable applied to the activation (e.g. a torch dropout
    module, which keeps torch's own random stream)."""
    c1w, c1b, c2w, c2b, f1w, f1b, f2w, f2b = params
    x = x.view(-1, 1, 28, 28)
    h = F.relu(F.max_pool2d(F.conv2d(x, c1w, c1b), 2))
    h = F.conv2d(h, c2w, c2b)
    if callable(m2):
        h = m2(h)
    elif m2 is not None:
        h = h * m2.view(-1, 20, 1, 1)
    h = F.relu(F.max_pool2d(h, 2))
    h = h.reshape(-1, 320)
    h = F.relu(F.linear(h, f1w, f1b))
    if callable(m3):
        h = m3(h)
    elif m3 is not None:
        h = h * m3
    h = F.linear(h, f2w, f2b)
    return F.log_softmax(h, dim=1)


def net_loss(params, x, target, m2=None, m3=None, reduction="mean"):
    return F.nll_loss(net_forward(params, x, m2, m3), target, reduction=reduction)


def sgd_momentum_step(params, grads, bufs, lr: float, momentum: float):
    """torch.optim.SGD(momentum, dampening=0) with zero-initialised buffers (T6)."""
    with torch.no_grad():
        for p, g, b in zip(params, grads, bufs):
            b.mul_(momentum).add_(g)
            p.add_(b, alpha=-lr)
