"""Flat parameter / gradient / momentum buffers.

All parameters of a module live in ONE contiguous device buffer (state_dict order,
every tensor start aligned to 16 bytes) and the `nn.Parameter`s become views into it,
so `state_dict()` stays torch-compatible (reference keys: src/train.py:84,
src/train_dist.py:163-164) while the optimizer and the gradient all-reduce see a single
float4-addressable array. Gradients get the same layout in a second buffer: it is the
DDP "bucket" (reference: DDP(model) at src/train_dist.py:63 keeps one 87,360-byte
bucket for `Net`; here it is 87,392 bytes with the alignment padding).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional

import torch
from torch import nn

ALIGN_BYTES = 16


@dataclass
class Slot:
    name: str
    offset: int  # in elements
    numel: int
    shape: torch.Size


def plan_layout(named_shapes, elem_size: int = 4, align_bytes: int = ALIGN_BYTES):
    """Offsets (in elements) for tensors packed in order with aligned starts.

    Returns (slots, total_elements); total is padded to the alignment too.
    """
    a = max(1, align_bytes // elem_size)
    slots: List[Slot] = []
    off = 0
    for name, shape in named_shapes:
        n = 1
        for s in shape:
            n *= int(s)
        slots.append(Slot(name, off, n, torch.Size(shape)))
        off = (off + n + a - 1) // a * a
    return slots, off


class FlatParams:
    """Re-homes a module's parameters into flat param/grad/momentum buffers.

    alloc(numel, dtype, device) -> 1-D tensor lets the caller place the buffers in
    symmetric (peer-mapped) memory; the default is a plain torch allocation.
    """

    def __init__(self, module: nn.Module, alloc: Optional[Callable] = None,
                 with_momentum: bool = True, grad_alloc: Optional[Callable] = None):
        params = [(n, p) for n, p in module.named_parameters()]
        if not params:
            raise ValueError("module has no parameters")
        dt = params[0][1].dtype
        dev = params[0][1].device
        for n, p in params:
            if p.dtype != dt or p.device != dev:
                raise ValueError("FlatParams needs one dtype and device for all parameters")
        self.module = module
        self.dtype, self.device = dt, dev
        self.slots, self.numel = plan_layout([(n, p.shape) for n, p in params],
                                             elem_size=params[0][1].element_size())
        alloc = alloc or (lambda n, dtype, device: torch.zeros(n, dtype=dtype, device=device))
        self.param = alloc(self.numel, dt, dev)
        self.grad = (grad_alloc or alloc)(self.numel, dt, dev)
        self.momentum = alloc(self.numel, torch.float32 if dt != torch.float32 else dt, dev) \
            if with_momentum else None
        self.param.zero_()
        self.grad.zero_()
        if self.momentum is not None:
            self.momentum.zero_()
        with torch.no_grad():
            for (n, p), s in zip(params, self.slots):
                view = self.param[s.offset:s.offset + s.numel].view(s.shape)
                view.copy_(p.data)
                p.data = view
                p.grad = self.grad[s.offset:s.offset + s.numel].view(s.shape)
        self._params = [p for _, p in params]

    def views(self, flat: torch.Tensor):
        return [flat[s.offset:s.offset + s.numel].view(s.shape) for s in self.slots]

    def offsets(self):
        return [s.offset for s in self.slots]

    def rebind_grads(self):
        """Point every p.grad back at its slot (after zero_grad(set_to_none=True))."""
        for p, s in zip(self._params, self.slots):
            p.grad = self.grad[s.offset:s.offset + s.numel].view(s.shape)

    def momentum_state_dict(self):
        """torch.optim.SGD-shaped state: {index: {'momentum_buffer': tensor}}."""
        return {i: {"momentum_buffer": v.clone()}
                for i, v in enumerate(self.views(self.momentum))}
