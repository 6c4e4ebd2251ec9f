"""Autograd-facing bindings of the layered `Net` kernels (csrc/net_kernels.cu).

One `torch.autograd.Function` per fused layer of the reference forward pass
(src/model.py:16-22):

  conv1_pool_relu      = relu(max_pool2d(conv1(x), 2))                      (K1+K2 / K13)
  conv2_drop_pool_relu = relu(max_pool2d(dropout2d(conv2(x)), 2))           (K3+K4+K5 / K12)
  fc1_relu_drop        = dropout(relu(fc1(x)))                              (K7 / K11)
  fc2_log_softmax      = log_softmax(fc2(x))                                (K8+K9 / K10)
  nll_loss             = F.nll_loss (mean or sum)                           (K9)

Backward kernels write weight gradients with a deterministic two-stage reduction
(per-CTA partial rows in a workspace, summed in row order by the last CTA).
On CPU tensors the callers use the PyTorch oracle instead (models/oracle.py); on CUDA
tensors there is no fallback: a missing native library raises.
"""
from __future__ import annotations

from typing import Optional

import torch

from .. import _native as N

_WS = {}


class Workspace:
    """Per-device scratch for the backward reductions: partial rows + arrival counter."""

    def __init__(self, device):
        rows = N.lib().b200_net_max_rows()
        self.rows = rows
        self.scratch = torch.empty(N.lib().b200_net_scratch_floats(rows), dtype=torch.float32,
                                   device=device)
        self.counter = torch.zeros(4, dtype=torch.int32, device=device)


def workspace(device) -> Workspace:
    key = torch.device(device).index if torch.device(device).index is not None \
        else torch.cuda.current_device()
    ws = _WS.get(key)
    if ws is None:
        ws = _WS[key] = Workspace(torch.device("cuda", key))
    return ws


def _c(t: torch.Tensor) -> torch.Tensor:
    return t if t.is_contiguous() else t.contiguous()


class Conv1PoolRelu(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b):
        x, w, b = _c(x), _c(w), _c(b)
        B = x.shape[0]
        a1 = torch.empty(B, 10, 12, 12, dtype=torch.float32, device=x.device)
        pos1 = torch.empty(B, 10, 12, 12, dtype=torch.int16, device=x.device)
        N.call("b200_conv1_fwd", N.ptr(x), N.ptr(w), N.ptr(b), N.ptr(a1), N.ptr(pos1), B, N.stream())
        ctx.save_for_backward(x, a1, pos1)
        return a1

    @staticmethod
    def backward(ctx, da1):
        x, a1, pos1 = ctx.saved_tensors
        if ctx.needs_input_grad[0]:
            raise N.NativeError("conv1_pool_relu: gradient w.r.t. the input image is not "
                                "implemented (the reference never needs it)")
        B = x.shape[0]
        ws = workspace(x.device)
        dw = torch.empty(10, 1, 5, 5, dtype=torch.float32, device=x.device)
        db = torch.empty(10, dtype=torch.float32, device=x.device)
        N.call("b200_conv1_bwd", N.ptr(x), N.ptr(_c(da1)), N.ptr(a1), N.ptr(pos1), N.ptr(dw),
               N.ptr(db), N.ptr(ws.scratch), N.ptr(ws.counter), B, N.stream())
        return None, dw, db


class Conv2DropPoolRelu(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a1, w, b, mask, mask_mode, p, seed, step, impl="simt"):
        a1, w, b = _c(a1), _c(w), _c(b)
        B = a1.shape[0]
        a2 = torch.empty(B, 20, 4, 4, dtype=torch.float32, device=a1.device)
        pos2 = torch.empty(B, 20, 4, 4, dtype=torch.uint8, device=a1.device)
        if mask_mode == 1 and mask is None:
            mask = torch.empty(B, 20, dtype=torch.float32, device=a1.device)
        # "tc": TF32 implicit GEMM without im2col, TMA-fed (csrc/conv2_tc.cu);
        # "tcgen05": the first tensor-core version with an in-smem im2col (csrc/conv_tcgen05.cu);
        # "simt": fp32-exact CUDA-core kernel (csrc/net_kernels.cu)
        # "auto": the measured choice for the layered path (profiles/r2/conv2_bwd_bench.json,
        #         conv2_tc_bench.json): tensor-core backward always (7-14x the SIMT kernel),
        #         tensor-core forward from batch 2048 up
        fwd_tc = impl == "tc" or (impl == "auto" and B >= 2048)
        if fwd_tc:
            c4 = torch.empty(B, 4, 12, 12, 4, dtype=torch.float32, device=a1.device)
            N.call("b200_a1_to_c4", N.ptr(a1), N.ptr(c4), B, N.stream())
            wop = torch.empty(N.lib().b200_conv2_operand_floats(), dtype=torch.float32, device=a1.device)
            N.call("b200_conv2_w_to_operand", N.ptr(w), N.ptr(wop), N.stream())
            N.call("b200_conv2_fwd_tc2", N.ptr(c4), N.ptr(wop), N.ptr(b), N.ptr(mask), mask_mode, float(p),
                   int(seed), int(step), N.ptr(a2), N.ptr(pos2), B, N.stream())
        else:
            N.call("b200_conv2_fwd_tc" if impl == "tcgen05" else "b200_conv2_fwd", N.ptr(a1), N.ptr(w),
                   N.ptr(b), N.ptr(mask), mask_mode, float(p),
                   int(seed), int(step), N.ptr(a2), N.ptr(pos2), B, N.stream())
        ctx.tc = impl in ("tc", "auto")
        ctx.save_for_backward(a1, w, a2, pos2, mask if mask_mode else None)
        return a2

    @staticmethod
    def backward(ctx, da2):
        a1, w, a2, pos2, mask = ctx.saved_tensors
        B = a1.shape[0]
        if ctx.tc:
            return Conv2DropPoolRelu._backward_tc(ctx, a1, w, a2, pos2, mask, _c(da2)) + (None,) * 6
        ws = workspace(a1.device)
        dw = torch.empty_like(w)
        db = torch.empty(20, dtype=torch.float32, device=a1.device)
        da1 = torch.empty_like(a1) if ctx.needs_input_grad[0] else None
        N.call("b200_conv2_bwd", N.ptr(a1), N.ptr(w), N.ptr(_c(da2)), N.ptr(a2), N.ptr(pos2),
               N.ptr(mask), N.ptr(dw), N.ptr(db), N.ptr(da1), N.ptr(ws.scratch),
               N.ptr(ws.counter), B, N.stream())
        return da1, dw, db, None, None, None, None, None, None


def _conv2_backward_tc(ctx, a1, w, a2, pos2, mask, da2):
    """conv2 backward on tcgen05 (csrc/conv2_tc.cu): the weight/bias gradient as a pixel-reduction
    GEMM with MN-major operands in a swizzled [pixel][32 channels] layout, the data gradient as
    an implicit GEMM over a zero-padded dZ frame."""
    B, dev, lib = a2.shape[0], a2.device, N.lib()
    c32 = torch.empty(lib.b200_conv2_c32_a1_floats(B), dtype=torch.float32, device=dev)
    dz32 = torch.empty(lib.b200_conv2_c32_dz_floats(B), dtype=torch.float32, device=dev)
    N.call("b200_a1_to_c32", N.ptr(a1), N.ptr(c32), B, N.stream())
    need_da1 = bool(ctx.needs_input_grad[0])
    frame = None
    if need_da1:    # both dZ layouts from one pass over the pooled gradient
        frame = torch.empty(lib.b200_conv2_frame_floats(B), dtype=torch.float32, device=dev)
        N.call("b200_conv2_dz_both", N.ptr(da2), N.ptr(a2), N.ptr(pos2), N.ptr(mask), N.ptr(frame),
               N.ptr(dz32), B, N.stream())
    else:
        N.call("b200_conv2_dz_c32", N.ptr(da2), N.ptr(a2), N.ptr(pos2), N.ptr(mask), N.ptr(dz32), B, N.stream())
    dw = torch.empty_like(w)
    db = torch.empty(20, dtype=torch.float32, device=dev)
    N.call("b200_conv2_wgrad_tc", N.ptr(c32), N.ptr(dz32), N.ptr(dw), N.ptr(db), B, N.stream())
    da1 = None
    if need_da1:
        wd = torch.empty(lib.b200_conv2_dgrad_operand_floats(), dtype=torch.float32, device=dev)
        N.call("b200_conv2_w_to_dgrad_operand", N.ptr(w), N.ptr(wd), N.stream())
        da1 = torch.empty(B, 10, 12, 12, dtype=torch.float32, device=dev)
        N.call("b200_conv2_dgrad_tc", N.ptr(frame), N.ptr(wd), N.ptr(da1), B, N.stream())
    return da1, dw, db


Conv2DropPoolRelu._backward_tc = staticmethod(_conv2_backward_tc)


def conv2_backward_tc(a1, w, a2, pos2, mask, da2, need_da1: bool = True):
    """(da1, dw, db) of conv2 + Dropout2d + pool + ReLU on the tensor cores, from the layer's
    saved forward state (a2 and arg-max positions pos2 of either forward kernel)."""
    class _Ctx:
        needs_input_grad = (need_da1,)
    return _conv2_backward_tc(_Ctx, _c(a1), _c(w), _c(a2), _c(pos2), mask, _c(da2))


class Fc1ReluDrop(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a2, W, b, mask, mask_mode, p, seed, step):
        a2, W, b = _c(a2), _c(W), _c(b)
        B = a2.shape[0]
        h1 = torch.empty(B, 50, dtype=torch.float32, device=a2.device)
        if mask_mode == 1 and mask is None:
            mask = torch.empty(B, 50, dtype=torch.float32, device=a2.device)
        N.call("b200_fc1_fwd", N.ptr(a2), N.ptr(W), N.ptr(b), N.ptr(mask), mask_mode, float(p),
               int(seed), int(step), N.ptr(h1), B, N.stream())
        ctx.save_for_backward(a2, W, h1, mask if mask_mode else None)
        return h1

    @staticmethod
    def backward(ctx, dh1):
        a2, W, h1, mask = ctx.saved_tensors
        B = a2.shape[0]
        ws = workspace(a2.device)
        dW = torch.empty_like(W)
        db = torch.empty(50, dtype=torch.float32, device=a2.device)
        da2 = torch.empty_like(a2) if ctx.needs_input_grad[0] else None
        N.call("b200_fc1_bwd", N.ptr(a2), N.ptr(W), N.ptr(_c(dh1)), N.ptr(h1), N.ptr(mask),
               N.ptr(dW), N.ptr(db), N.ptr(da2), N.ptr(ws.scratch), N.ptr(ws.counter), B,
               N.stream())
        return da2, dW, db, None, None, None, None, None


class Fc2LogSoftmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, h1, W, b):
        h1, W, b = _c(h1), _c(W), _c(b)
        B = h1.shape[0]
        logp = torch.empty(B, 10, dtype=torch.float32, device=h1.device)
        N.call("b200_fc2_lsm_fwd", N.ptr(h1), N.ptr(W), N.ptr(b), N.ptr(logp), B, N.stream())
        ctx.save_for_backward(h1, W, logp)
        return logp

    @staticmethod
    def backward(ctx, dlogp):
        h1, W, logp = ctx.saved_tensors
        B = h1.shape[0]
        ws = workspace(h1.device)
        dW = torch.empty_like(W)
        db = torch.empty(10, dtype=torch.float32, device=h1.device)
        dh1 = torch.empty_like(h1) if ctx.needs_input_grad[0] else None
        N.call("b200_fc2_lsm_bwd", N.ptr(h1), N.ptr(W), N.ptr(logp), N.ptr(_c(dlogp)), N.ptr(dW),
               N.ptr(db), N.ptr(dh1), N.ptr(ws.scratch), N.ptr(ws.counter), B, N.stream())
        return dh1, dW, db


class NllLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logp, target, reduction):
        logp = _c(logp)
        target = _c(target.to(torch.int64))
        B = logp.shape[0]
        loss = torch.empty((), dtype=torch.float32, device=logp.device)
        N.call("b200_nll_fwd", N.ptr(logp), N.ptr(target), N.ptr(loss), B, reduction, N.stream())
        ctx.save_for_backward(target)
        ctx.reduction = reduction
        ctx.B = B
        return loss

    @staticmethod
    def backward(ctx, gloss):
        (target,) = ctx.saved_tensors
        dlogp = torch.empty(ctx.B, 10, dtype=torch.float32, device=target.device)
        N.call("b200_nll_bwd", N.ptr(_c(gloss.to(torch.float32))), N.ptr(target), N.ptr(dlogp),
               ctx.B, ctx.reduction, N.stream())
        return dlogp, None, None


def conv1_pool_relu(x, w, b):
    return Conv1PoolRelu.apply(x, w, b)


def conv2_drop_pool_relu(a1, w, b, mask=None, mask_mode=0, p=0.5, seed=0, step=0, impl="simt"):
    return Conv2DropPoolRelu.apply(a1, w, b, mask, mask_mode, p, seed, step, impl)


def fc1_relu_drop(a2, W, b, mask=None, mask_mode=0, p=0.5, seed=0, step=0):
    return Fc1ReluDrop.apply(a2, W, b, mask, mask_mode, p, seed, step)


def fc2_log_softmax(h1, W, b):
    return Fc2LogSoftmax.apply(h1, W, b)


_REDUCTIONS = {"mean": 0, "sum": 1}


def nll_loss(logp: torch.Tensor, target: torch.Tensor, reduction: str = "mean",
             size_average: Optional[bool] = None) -> torch.Tensor:
    """F.nll_loss for [B,10] log-probabilities (src/train.py:74,94).

    `size_average=False` (deprecated spelling used at src/train.py:94) means sum.
    """
    if size_average is not None:
        reduction = "mean" if size_average else "sum"
    if not logp.is_cuda:
        return torch.nn.functional.nll_loss(logp, target, reduction=reduction)
    if logp.shape[1] != 10 or logp.dtype != torch.float32:
        raise N.NativeError("nll_loss kernel handles fp32 [B,10] log-probabilities")
    return NllLoss.apply(logp, target, _REDUCTIONS[reduction])


def cross_entropy_on_log_probs(logp, target):
    """`CrossEntropyLoss()` applied to log-softmax output (src/train_dist.py:67,82).

    A second log_softmax of log-probabilities is the identity (their logsumexp is 0),
    so this equals NLL exactly; one kernel serves both scripts.
    """
    return nll_loss(logp, target, "mean")


def gather_normalize(data_u8: torch.Tensor, idx: Optional[torch.Tensor] = None) -> torch.Tensor:
    """u8 rows (optionally gathered by int32 idx) -> normalised fp32 [B,1,28,28] (K21)."""
    if not data_u8.is_cuda:
        from ..models.oracle import normalize_u8
        rows = data_u8 if idx is None else data_u8[idx.long()]
        return normalize_u8(rows.view(-1, 784))
    data_u8 = _c(data_u8)
    B = data_u8.shape[0] if idx is None else idx.shape[0]
    if idx is not None:
        idx = _c(idx.to(torch.int32))
    out = torch.empty(B, 1, 28, 28, dtype=torch.float32, device=data_u8.device)
    N.call("b200_gather_normalize", N.ptr(data_u8), N.ptr(idx), N.ptr(out), B, N.stream())
    return out
