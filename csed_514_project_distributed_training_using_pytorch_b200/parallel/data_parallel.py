"""DataParallel: the `DDP(model)` of the reference (src/train_dist.py:63), B200-native.

Construction: verifies that every rank holds the same parameter shapes,
re-homes the parameters and their gradients into flat symmetric buffers, and broadcasts
rank 0's values with the peer-memory broadcast kernel (K16), so ranks seeded differently
start identical. Backward (T3): post-accumulate-grad hooks mark gradients ready; the
reducer launches one fused all-reduce kernel per complete bucket on a side stream while
backward continues; `.grad` holds the mean over ranks before `optimizer.step()`.
Evaluation / `torch.no_grad()` forwards do no synchronisation and no bookkeeping (T4).
`no_sync()` skips the reduction for gradient accumulation.

For the reference network there is a faster door: `fused_engine()` returns a
`FusedNetEngine` whose kernel B does partial-row sum + all-reduce + SGD in one launch.
"""
from __future__ import annotations

import contextlib
from typing import Optional, Sequence

import torch
from torch import nn

from ..utils.flat import FlatParams
from .reducer import Reducer


def module_bytes(module: nn.Module) -> int:
    return sum(p.numel() * p.element_size() for p in module.parameters())


class DataParallel(nn.Module):
    def __init__(self, module: nn.Module, comm=None, bucket_of: Optional[Sequence[int]] = None,
                 fuse_sgd=None, algo="auto", broadcast_from: int = 0, pingpong: bool = False):
        """algo: "auto" | "oneshot" | "twoshot" | "nvls". "auto" (the default, like the
        reference's knob-free `DDP(model)`, src/train_dist.py:63) picks one-shot, two-shot or
        NVLS per bucket from the measured crossover table; "nvls" forces the in-switch reduction
        and falls back to "twoshot" when the machine does not offer multicast.
        pingpong (needs fuse_sgd): two parameter buffers. Step t reads buffer t & 1 and the
        bucket kernels write the updated weights into the other one, so a layer's weight
        gradient and its bucket's all-reduce + update can be queued BEFORE the layer's data
        gradient, which still reads the old weights (layers that know about it:
        models/mlp.py). `live_param(i)` is parameter i as the current step must read it;
        `sync_module()` copies the live buffer back into the module's own parameters."""
        super().__init__()
        if pingpong and not getattr(module, "supports_pingpong", False):
            # a module that reads its own nn.Parameters would never see the second buffer
            raise ValueError("pingpong=True needs a module whose layers read their weights "
                             "through sink.live_param() (models.mlp.WideMLP does)")
        if pingpong and fuse_sgd is None:
            raise ValueError("pingpong=True needs fuse_sgd: the bucket kernels write the other buffer")
        if comm is None:
            from .process_group import get_comm
            comm = get_comm()
        self.module = module
        self.comm = comm
        self._verify_shapes()
        grad_alloc = None
        self.nvls = False
        if algo in ("nvls", "auto") and hasattr(comm, "nvls_init") and comm.world_size > 1 \
                and module_bytes(module) >= (1 << 20):
            # gradient buckets live in the multicast-bound region whenever the machine offers
            # NVLS, so "auto" can pick the in-switch reduction per bucket from the measured table
            # (tiny models stay in the plain arena: their buckets are one-shot territory)
            if comm.nvls_init():
                grad_alloc = comm.alloc_nvls
                self.nvls = True
        if algo == "nvls" and not self.nvls:
            algo = "twoshot"
        self.flat = FlatParams(module, alloc=comm.alloc, grad_alloc=grad_alloc)
        # init_sync: rank 0's parameters (and buffers) win (torch DDP default)
        if comm.world_size > 1:
            comm.broadcast(self.flat.param, root=broadcast_from)
            for buf in module.buffers():
                comm.broadcast(buf, root=broadcast_from)
        self.reducer = Reducer(self.flat, comm, bucket_of=bucket_of, fuse_sgd=fuse_sgd, algo=algo)
        self._params = list(module.parameters())
        self._grad_views = self.flat.views(self.flat.grad)
        if hasattr(module, "sink"):
            module.sink = self       # native layers route their gradients through us
        self._require_backward_sync = True
        self._pass_open = False
        self._sunk = set()
        self.pingpong = bool(pingpong) and self.flat.param.is_cuda    # CPU/gloo: one buffer
        self._live = 0
        if self.pingpong:
            self.param_alt = comm.alloc(self.flat.param.numel(), self.flat.param.dtype)
            self.param_alt.copy_(self.flat.param)
            self._alt_leaves = [v.requires_grad_(True) for v in self.flat.views(self.param_alt)]
            self.reducer.update_target = self.param_alt      # step 0 reads flat.param
        for i, p in enumerate(self._params):
            if p.requires_grad:
                p.register_post_accumulate_grad_hook(self._make_hook(i))

    def _verify_shapes(self):
        shapes = [tuple(p.shape) for p in self.module.parameters()]
        all_shapes = self.comm.all_gather_object(shapes)
        for r, s in enumerate(all_shapes):
            if s != all_shapes[0]:
                raise RuntimeError(f"parameter shapes differ between rank 0 and rank {r}")

    def _make_hook(self, i: int):
        def hook(param):
            if not (self._require_backward_sync and self._pass_open):
                return
            if i in self._sunk:      # a native layer already delivered this gradient
                return
            # autograd may have rebound .grad to a fresh tensor: move it into the bucket
            s = self.flat.slots[i]
            slot = self.flat.grad[s.offset:s.offset + s.numel].view(s.shape)
            if param.grad is not None and param.grad.data_ptr() != slot.data_ptr():
                slot.copy_(param.grad)
                param.grad = slot
            self.reducer.mark_ready(i)
            if self.reducer.core.all_done():
                self.reducer.finalize_backward()
                self._end_pass()
        return hook

    def live_param(self, i: int) -> torch.Tensor:
        """Parameter i in the buffer the current step reads (ping-pong mode)."""
        return self._params[i] if self._live == 0 else self._alt_leaves[i]

    def sync_module(self):
        """Ping-pong mode: make the module's own parameters hold the live weights (before a
        checkpoint or an evaluation that does not go through `live_param`)."""
        if self.pingpong and self._live == 1:
            if self.reducer.is_cuda:
                torch.cuda.current_stream().wait_stream(self.reducer.stream)
            self.flat.param.copy_(self.param_alt)
            self._live = 0
            self.reducer.update_target = self.param_alt

    def _end_pass(self):
        self._pass_open = False
        if self.pingpong:       # the buckets of this pass wrote the other buffer: it is live now
            self._live ^= 1
            self.reducer.update_target = self.param_alt if self._live == 0 else self.flat.param

    def _begin_pass(self):
        if torch.is_grad_enabled() and self.training and self._require_backward_sync:
            # make sure the previous step's (fused) update is visible to this forward. Inside
            # a CUDA-graph capture the previous replay has already joined the side stream
            # (finalize_backward), and waiting on an un-captured stream is not allowed.
            if self.reducer.is_cuda and not torch.cuda.is_current_stream_capturing():
                torch.cuda.current_stream().wait_stream(self.reducer.stream)
            self.reducer.prepare_for_backward()
            self._pass_open = True
            self._sunk = set()

    def forward(self, *args, **kwargs):
        self._begin_pass()
        return self.module(*args, **kwargs)

    def loss(self, *args, **kwargs):
        """module.loss(...) with the same backward bookkeeping as forward()."""
        self._begin_pass()
        return self.module.loss(*args, **kwargs)

    # -- gradient sink: layers that write dW/db straight into the bucket --------------
    def view(self, i: int) -> torch.Tensor:
        """Slot of parameter i inside the flat gradient bucket."""
        return self._grad_views[i]

    def rs_target(self, i: int):
        """Fused GEMM -> reduce-scatter target for parameter i's gradient, or None."""
        if not (self._require_backward_sync and self._pass_open):
            return None
        return self.reducer.rs_target(i)

    def mark_pushed(self, target):
        self.reducer.mark_pushed(target)

    def ready(self, i: int):
        """Parameter i's gradient is final in its slot (called from a layer's backward)."""
        if not (self._require_backward_sync and self._pass_open):
            return
        self._sunk.add(i)
        self.reducer.mark_ready(i)
        if self.reducer.core.all_done():
            self.reducer.finalize_backward()
            self._end_pass()

    def finish_backward(self):
        """Call after `loss.backward()`: if a parameter received no gradient in this pass its
        bucket never filled, so launch what is still pending (zeros for that parameter, like
        torch DDP's find_unused_parameters) and join the reducer stream. A no-op when every
        bucket was launched from the hooks."""
        if self._pass_open:
            self.reducer.finalize_backward()
            self._end_pass()

    @contextlib.contextmanager
    def no_sync(self):
        """Accumulate gradients locally (no all-reduce) inside the context."""
        old = self._require_backward_sync
        self._require_backward_sync = False
        try:
            yield
        finally:
            self._require_backward_sync = old

    def zero_grad(self, set_to_none: bool = False):
        """Gradients live in the flat bucket: zero in place and keep the views bound."""
        self.flat.grad.zero_()
        self.flat.rebind_grads()

    def fused_engine(self, lr: float, momentum: float, seed: int = 1, **kw):
        """The fused engine for models.Net: the persistent step kernel (gradient exchange
        overlapped with backward, csrc/net_persist.cu) plus the two-kernel step with the
        all-reduce inside kernel B."""
        from ..ops.persist import PersistentNetEngine
        return PersistentNetEngine(self.module, lr=lr, momentum=momentum, seed=seed,
                                   comm=self.comm, flat=self.flat, **kw)

    def state_dict(self, *a, **k):
        """The UNWRAPPED module's state_dict: no `module.` prefix (src/train_dist.py:116,164)."""
        return self.module.state_dict(*a, **k)

    def load_state_dict(self, sd, *a, **k):
        return self.module.load_state_dict(sd, *a, **k)
