"""The widened MNIST MLP of BASELINE.json config 4 (784-4096-4096-10, bf16): the model that
stresses gradient all-reduce bandwidth (20,037,642 parameters = 38.2 MiB of bf16 gradients).

Every matrix product runs on the hand-written tcgen05 GEMM (ops/gemm.py): forward with
bias + ReLU in the epilogue, dgrad with the producer's ReLU mask in the epilogue, wgrad
writing straight into the layer's slot of the flat gradient bucket. Each layer is a
`torch.autograd.Function`; when a gradient sink is attached (parallel.DataParallel), its
backward stores dW/db into the bucket and tells the reducer the parameters are ready, so
the bucket's fused all-reduce + SGD kernel overlaps the backward of the layers below
(SURVEY §7.1 point 2; torch DDP's hook at src/train_dist.py:83).
"""
from __future__ import annotations

from typing import Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import _native as N
from ..ops import gemm


class LinearFn(torch.autograd.Function):
    """y = relu?(x W^T + b) on tcgen05; see module docstring for the gradient routing."""

    @staticmethod
    def forward(ctx, x, W, b, relu, mask_input_grad, premasked_grad, sink, iw, ib, out_pad,
                out_fp32):
        B = x.shape[0]
        n_out = W.shape[0]
        if out_pad:
            y = torch.zeros(B, out_pad, dtype=torch.float32 if out_fp32 else torch.bfloat16,
                            device=x.device)
            gemm.gemm_bf16(x, W, False, False, out=y, bias=b, relu=relu, n_valid=n_out)
        else:
            y = gemm.linear_fwd(x, W, b, relu=relu,
                                out_dtype=torch.float32 if out_fp32 else torch.bfloat16)
        ctx.save_for_backward(x, W, y if relu else None)
        ctx.cfg = (relu, mask_input_grad, premasked_grad, sink, iw, ib, n_out)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, W, y = ctx.saved_tensors
        relu, mask_in, premasked, sink, iw, ib, n_out = ctx.cfg
        if dy.dtype != torch.bfloat16:
            dy = dy.to(torch.bfloat16)
        if not dy.is_contiguous():
            dy = dy.contiguous()
        if relu and not premasked:
            dy = dy * (y > 0)
        dyv = dy[:, :n_out] if dy.shape[1] != n_out else dy   # cut padded logit columns
        dx = None
        # Without ping-pong buffers dgrad reads the W that the bucket kernel updates, so it must
        # be queued first. With them (DataParallel(pingpong=True)) the update goes to the other
        # buffer: wgrad and the bucket's all-reduce start first and dgrad overlaps them.
        late_dgrad = sink is not None and getattr(sink, "pingpong", False)
        if ctx.needs_input_grad[0] and not late_dgrad:
            dx = gemm.linear_dgrad(dyv, W, relu_mask=x if mask_in else None)
        if sink is not None:
            rs = sink.rs_target(iw) if hasattr(sink, "rs_target") else None
            # fused GEMM -> reduce-scatter: the wgrad epilogue stores each result vector into
            # the staging slot of the rank that owns it, over NVLink, tile by tile
            gemm.gemm_bf16(dyv, x, True, True, out=sink.view(iw), rs=rs)
            if rs is not None:
                sink.mark_pushed(rs)
            N.call("b200_colsum_bf16", N.ptr(dyv), dyv.stride(0), N.ptr(sink.view(ib)),
                   dyv.shape[0], n_out, N.stream())
            sink.ready(iw)
            sink.ready(ib)
            dW = db = None
            if ctx.needs_input_grad[0] and late_dgrad:
                dx = gemm.linear_dgrad(dyv, W, relu_mask=x if mask_in else None)
        else:
            dW = gemm.gemm_bf16(dyv, x, True, True)
            db = torch.empty(n_out, dtype=torch.bfloat16, device=x.device)
            N.call("b200_colsum_bf16", N.ptr(dyv), dyv.stride(0), N.ptr(db), dyv.shape[0], n_out,
                   N.stream())
        return dx, dW, db, None, None, None, None, None, None, None, None


class SoftmaxNLLFn(torch.autograd.Function):
    """Mean cross-entropy over padded fp32 logits; the gradient is produced in the same kernel."""

    @staticmethod
    def forward(ctx, logits, labels, ncls):
        B, ld = logits.shape
        loss = torch.empty((), dtype=torch.float32, device=logits.device)
        dlog = torch.empty(B, ld, dtype=torch.bfloat16, device=logits.device)
        N.call("b200_softmax_nll_bf16", N.ptr(logits), ld, ncls, N.ptr(labels), N.ptr(loss),
               N.ptr(dlog), ld, B, N.stream())
        ctx.save_for_backward(dlog)
        return loss

    @staticmethod
    def backward(ctx, g):
        (dlog,) = ctx.saved_tensors
        return dlog.float() * g, None, None


N.register_sigs({
    "b200_softmax_nll_bf16": [N.c_p, N.c_i, N.c_i, N.c_p, N.c_p, N.c_p, N.c_i, N.c_i, N.c_p],
    "b200_colsum_bf16": [N.c_p, N.c_ll, N.c_p, N.c_i, N.c_i, N.c_p],
    "b200_sgd_master_bf16": [N.c_p, N.c_p, N.c_p, N.c_p, N.c_ll, N.c_p, N.c_p],
})


class WideMLP(nn.Module):
    """784-4096-4096-10 with ReLU between layers (BASELINE.json config 4)."""

    supports_pingpong = True    # forward reads sink.live_param(i): see DataParallel(pingpong=True)

    def __init__(self, sizes: Sequence[int] = (784, 4096, 4096, 10), dtype=torch.bfloat16):
        super().__init__()
        self.sizes = tuple(sizes)
        self.layers = nn.ModuleList([nn.Linear(a, b) for a, b in zip(sizes[:-1], sizes[1:])])
        self.to(dtype)
        self.sink = None   # set by parallel.DataParallel / MLPEngine

    def logits_padded(self, x: torch.Tensor) -> torch.Tensor:
        nl = len(self.layers)
        h = x
        pp = self.sink is not None and getattr(self.sink, "pingpong", False)
        for i, lin in enumerate(self.layers):
            last = i == nl - 1
            # ping-pong mode: this step's weights are in the buffer the last update wrote
            W, b = (self.sink.live_param(2 * i), self.sink.live_param(2 * i + 1)) if pp \
                else (lin.weight, lin.bias)
            h = LinearFn.apply(h, W, b, not last, i > 0, not last, self.sink,
                               2 * i, 2 * i + 1, 16 if last else 0, last)
        return h

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """x: [B,784] (bf16 on CUDA). Returns fp32 logits [B,10]."""
        if not x.is_cuda:
            h = x.to(self.layers[0].weight.dtype)
            for i, lin in enumerate(self.layers):
                h = lin(h)
                if i < len(self.layers) - 1:
                    h = F.relu(h)
            return h.float()
        return self.logits_padded(x.to(torch.bfloat16).contiguous())[:, :self.sizes[-1]]

    def loss(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """Mean cross-entropy (CrossEntropyLoss, src/train_dist.py:67)."""
        if not x.is_cuda:
            return F.cross_entropy(self.forward(x), y)
        return SoftmaxNLLFn.apply(self.logits_padded(x.to(torch.bfloat16).contiguous()), y,
                                  self.sizes[-1])


class LocalSink:
    """Gradient sink for one GPU: dW/db land in the flat gradient buffer, nothing to reduce."""

    def __init__(self, flat):
        self.flat = flat
        self._views = flat.views(flat.grad)

    def view(self, i):
        return self._views[i]

    def ready(self, i):
        pass


class MasterSGD:
    """SGD-momentum for bf16 parameters with fp32 master weights (single GPU)."""

    def __init__(self, flat, lr, momentum, weight_decay=0.0):
        self.flat = flat
        self.master = flat.param.float().clone()
        self.mom = torch.zeros_like(self.master)
        self.hp = torch.tensor([lr, momentum, weight_decay, 0.0], dtype=torch.float32,
                               device=flat.param.device)

    def step(self):
        N.call("b200_sgd_master_bf16", N.ptr(self.flat.param), N.ptr(self.flat.grad),
               N.ptr(self.master), N.ptr(self.mom), self.flat.numel, N.ptr(self.hp), N.stream())
