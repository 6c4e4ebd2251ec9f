"""SGD with momentum over flat buffers (reference: optim.SGD at
src/train.py:60-61 and src/train_dist.py:66, semantics torch/optim/sgd.py:358-380).

`buf = momentum*buf + grad; param -= lr*buf`, dampening 0. The momentum buffer starts at
zero, which gives the same numbers as torch's "first step copies the gradient" rule.
Hyper-parameters sit in a small device tensor so a captured graph sees changes.
"""
from __future__ import annotations

import torch

from .. import _native as N


class FlatSGD:
    def __init__(self, param: torch.Tensor, grad: torch.Tensor, lr: float, momentum: float = 0.0,
                 weight_decay: float = 0.0, nesterov: bool = False, dampening: float = 0.0,
                 momentum_buffer: torch.Tensor = None):
        if dampening != 0.0:
            raise ValueError("FlatSGD implements dampening=0 (all the reference uses)")
        if nesterov and momentum <= 0:
            raise ValueError("nesterov needs momentum > 0")
        if param.numel() % 4 != 0:
            raise ValueError("flat buffers must be padded to a multiple of 4 elements")
        self.param, self.grad = param, grad
        self.buf = momentum_buffer if momentum_buffer is not None else torch.zeros_like(param)
        self.hp = torch.tensor([lr, momentum, weight_decay, 1.0 if nesterov else 0.0],
                               dtype=torch.float32, device=param.device)

    def set_lr(self, lr: float):
        self.hp[0] = lr

    def step(self, grad_scale: float = 1.0):
        if self.param.is_cuda:
            N.call("b200_sgd_momentum", N.ptr(self.param), N.ptr(self.grad), N.ptr(self.buf),
                   self.param.numel(), N.ptr(self.hp), float(grad_scale), N.stream())
        else:
            lr, mu, wd, nest = self.hp.tolist()
            g = self.grad * grad_scale + wd * self.param
            self.buf.mul_(mu).add_(g)
            self.param.sub_(lr * (g + mu * self.buf if nest else self.buf))

    def zero_grad(self):
        """Kept for API parity (src/train.py:72); the fused kernels overwrite gradients."""
        self.grad.zero_()
