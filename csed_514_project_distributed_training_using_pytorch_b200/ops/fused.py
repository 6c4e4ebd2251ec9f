"""Fused training step for `Net`: two kernels per optimizer step.

Reference hot loop (src/train.py:71-76, src/train_dist.py:78-84):
    zero_grad -> model(data) -> loss -> backward -> [DDP all-reduce] -> optimizer.step
which on a GPU is ~40 library kernels plus a host sync for `loss.item()`.

Here the whole loop body is
    kernel A  `b200_net_fwdbwd`      forward + log_softmax/NLL + backward, one image per
                                     CTA iteration, activations never leave shared memory,
                                     gradient partial rows go to scratch (no zero_grad)
    kernel B  `b200_net_reduce_sgd`  fixed-order sum of the rows + SGD-momentum in place
                                     (single GPU), or `b200_comm_reduce_allreduce_sgd`
                                     (same, with the cross-GPU all-reduce over NVSwitch
                                     peer memory between the sum and the update)
optionally captured once as a CUDA graph per batch shape and replayed. The loss of every
step lands in a device ring buffer and is read back per log interval, not per step (K18).
"""
from __future__ import annotations

import ctypes

import torch

from .. import _native as N
from ..utils.flat import FlatParams

X_F32, X_U8, X_U8_GATHER, X_F32_GATHER = 0, 1, 2, 3


def net_layout():
    o = (ctypes.c_int * 9)()
    N.lib().b200_net_layout(o)
    return list(o)


class FusedNetEngine:
    """Owns the flat parameter/momentum buffers of a `Net` and runs fused steps on them.

    net:       models.Net on a CUDA device (its parameters become views of the flat buffer)
    lr, momentum: SGD hyper-parameters (reference: 0.01/0.5 single, 0.02/0.5 distributed);
               they live in device memory, so `set_lr` needs no graph re-capture
    comm:      parallel.comm.Communicator or None; when given, buffers are symmetric
               (peer-mapped) and kernel B all-reduces across ranks
    dropout:   True draws Philox masks per (seed, step, sample); masks can be injected
               per call for parity tests
    """

    def __init__(self, net, lr: float = 0.01, momentum: float = 0.5, seed: int = 1,
                 comm=None, ring: int = 4096, dropout: bool = True, flat: FlatParams = None):
        p0 = next(net.parameters())
        if not p0.is_cuda:
            raise N.NativeError("FusedNetEngine needs the model on a CUDA device")
        self.net = net
        self.device = p0.device
        self.comm = comm
        alloc = comm.alloc if comm is not None else None
        self.flat = flat if flat is not None else FlatParams(net, alloc=alloc)
        lay = net_layout()
        if self.flat.offsets() != lay[:8] or self.flat.numel != lay[8]:
            raise N.NativeError(f"flat layout {self.flat.offsets()} != kernel layout {lay}")
        self.numel = lay[8]
        self.max_rows = N.lib().b200_net_max_rows()
        dev = self.device
        self.partial = torch.empty(self.max_rows * self.numel, dtype=torch.float32, device=dev)
        self.loss_partial = torch.zeros(self.max_rows, dtype=torch.float32, device=dev)
        self.hp = torch.tensor([lr, momentum, 0.0, 0.0], dtype=torch.float32, device=dev)
        # ctl[0] = optimizer step counter, ctl[1] = step at which the current epoch started
        self.step_ctr = torch.zeros(2, dtype=torch.int64, device=dev)
        self.ring = ring
        # [0, ring): per-step losses; [ring]: always the latest step's loss (fixed address
        # for a per-step device->host copy)
        self.loss_ring = torch.zeros(ring + 1, dtype=torch.float32, device=dev)
        self.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        self.dropout = dropout
        self.sample_base = 0 if comm is None else comm.rank * (1 << 20)
        self.eval_loss = None
        self.eval_pred = None
        self.eval_out = torch.zeros(2, dtype=torch.float32, device=dev)
        self._graphs = {}
        self.launches = 0  # kernels launched through this engine (bench's gpu_launches)

    # -- hyper-parameters ----------------------------------------------------
    def set_lr(self, lr: float):
        self.hp[0] = lr

    def set_momentum(self, m: float):
        self.hp[1] = m

    # -- one step ------------------------------------------------------------
    def forward_backward(self, x, y, x_mode=X_F32, gidx=None, masks=None, batch=None,
                         idx_stride=0):
        """Kernel A only: gradient partial rows + loss partials for this batch.

        idx_stride != 0 (gather modes): the batch's indices are
        gidx[(step - epoch_start) * idx_stride : ... + B], read on the device from the step
        counter, so one captured graph walks a whole epoch of a device-resident dataset."""
        B = int(batch if batch is not None else (gidx.shape[0] if gidx is not None else x.shape[0]))
        if masks is not None:
            mode, m2, m3 = 2, masks[0], masks[1]
        elif self.dropout and self.net.training:
            mode, m2, m3 = 1, None, None
        else:
            mode, m2, m3 = 0, None, None
        N.call("b200_net_fwdbwd", N.ptr(self.flat.param), N.ptr(x), x_mode, N.ptr(y), N.ptr(gidx),
               B, N.ptr(self.partial), N.ptr(self.loss_partial), mode, N.ptr(m2), N.ptr(m3),
               self.seed, N.ptr(self.step_ctr), self.sample_base, int(idx_stride),
               int(gidx.numel()) if gidx is not None else 0, N.stream())
        self.launches += 1
        return B

    def begin_epoch(self):
        """Mark the current step as the start of an epoch (for idx_stride stepping)."""
        self.step_ctr[1:2].copy_(self.step_ctr[0:1])

    def reduce_and_update(self, B: int, write_grad: bool = False, update: bool = True):
        """Kernel B: sum partial rows (+ all-reduce) + SGD-momentum; loss -> ring."""
        rows = N.lib().b200_net_fused_rows(B)
        if self.comm is not None and self.comm.world_size > 1:
            self.comm.reduce_allreduce_sgd(self, rows, write_grad=write_grad, update=update)
        else:
            N.call("b200_net_reduce_sgd", N.ptr(self.partial), rows,
                   N.ptr(self.flat.param) if update else None,
                   N.ptr(self.flat.momentum) if update else None,
                   N.ptr(self.flat.grad) if write_grad else None, N.ptr(self.hp),
                   N.ptr(self.loss_partial), N.ptr(self.loss_ring), self.ring,
                   N.ptr(self.step_ctr), N.stream())
        self.launches += 1

    def step(self, x, y, x_mode=X_F32, gidx=None, masks=None, write_grad=False, batch=None,
             idx_stride=0):
        """One optimizer step on the current stream (no host sync)."""
        B = self.forward_backward(x, y, x_mode, gidx, masks, batch, idx_stride)
        self.reduce_and_update(B, write_grad=write_grad)

    # -- CUDA graph ----------------------------------------------------------
    def capture(self, x, y, x_mode=X_F32, gidx=None, batch=None, key=None, idx_stride=0,
                unroll: int = 1):
        """Capture one step reading the given static input tensors; returns the graph.

        Replay with `graph.replay()` after copying new data into x / y (/ gidx).
        Kernel arguments are baked in; step counter, learning rate and the loss ring
        index live in device memory, so replays advance them correctly.

        unroll > 1 captures that many consecutive steps in one graph. It only makes sense
        with device-side index stepping (idx_stride != 0), where every step finds its own
        batch; the programmatic-dependent-launch edges then span step boundaries too.
        """
        s = torch.cuda.Stream(device=self.device)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            # warm-up launch outside capture (sets function attributes, faults in code)
            saved = self._snapshot()
            self.step(x, y, x_mode, gidx, batch=batch, idx_stride=idx_stride)
            self._restore(saved)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        launches_before = self.launches
        with torch.cuda.graph(g, stream=s):
            for _ in range(max(1, unroll)):
                self.step(x, y, x_mode, gidx, batch=batch, idx_stride=idx_stride)
        g.kernels_per_replay = self.launches - launches_before
        g.steps_per_replay = max(1, unroll)
        self.launches = launches_before
        # First launch of an instantiated graph uploads it to the device (tens of
        # microseconds for a 32-kernel graph). Pay that here, not in the caller's first
        # timed replay: one throw-away replay with the training state put back afterwards.
        # Every rank captures at the same point, so the collective inside kernel B matches.
        with torch.cuda.stream(s):
            saved = self._snapshot()
            g.replay()
            self._restore(saved)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize(self.device)
        if key is not None:
            self._graphs[key] = g
        return g

    def _snapshot(self):
        return (self.flat.param.clone(), self.flat.momentum.clone(), self.step_ctr.clone(),
                self.loss_ring.clone())

    def _restore(self, saved):
        self.flat.param.copy_(saved[0])
        self.flat.momentum.copy_(saved[1])
        self.step_ctr.copy_(saved[2])
        self.loss_ring.copy_(saved[3])

    # -- losses --------------------------------------------------------------
    def steps_done(self) -> int:
        return int(self.step_ctr[0].item())

    def last_losses(self, n: int) -> torch.Tensor:
        """The last n per-step mean losses (host sync; call per log interval, not per step)."""
        done = self.steps_done()
        n = min(n, done, self.ring)
        idx = [(done - n + i) % self.ring for i in range(n)]
        return self.loss_ring[idx].cpu()

    @property
    def last_loss(self) -> torch.Tensor:
        """1-element device view holding the most recent step's loss."""
        return self.loss_ring[self.ring:self.ring + 1]

    # -- evaluation (K19) ------------------------------------------------------
    def evaluate_batch(self, x, y, x_mode=X_F32, gidx=None, accumulate=False, logp_out=None,
                       batch=None):
        """Forward in eval mode + NLL sum + correct count into self.eval_out (device)."""
        B = int(batch if batch is not None else (gidx.shape[0] if gidx is not None else x.shape[0]))
        if self.eval_loss is None or self.eval_loss.numel() < B:
            self.eval_loss = torch.empty(B, dtype=torch.float32, device=self.device)
            self.eval_pred = torch.empty(B, dtype=torch.int32, device=self.device)
        N.call("b200_net_eval", N.ptr(self.flat.param), N.ptr(x), x_mode, N.ptr(y), N.ptr(gidx), B,
               N.ptr(logp_out), N.ptr(self.eval_loss), N.ptr(self.eval_pred),
               N.ptr(self.eval_out), 1 if accumulate else 0, N.stream())
        self.launches += 2
        return self.eval_out
