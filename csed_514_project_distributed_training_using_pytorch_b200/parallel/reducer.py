"""Gradient reducer: buckets, ready counting, stream wiring.

torch DDP's C++ Reducer (reached from `DDP(model)` at src/train_dist.py:63) copies ready
gradients into buckets, all-reduces each full bucket with NCCL while backward continues,
and copies the means back before `optimizer.step()`. This reducer keeps that contract but:

* gradients are VIEWS into one flat symmetric buffer, so there is no copy-in/copy-out;
* the bucket plan and ready counting are native C++ (csrc/reducer.cpp);
* a complete bucket is reduced by ONE hand-written kernel on a side stream (one-shot or
  two-shot over NVSwitch peer memory, chosen by size), which with `fuse_sgd` also applies
  the 1/world scale and the SGD-momentum update, sharding the optimizer state across ranks
  in the two-shot case (csrc/comm.cu: allreduce_kernel);
* the next forward waits on the comm stream, not the host.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence

import torch

from .. import _native as N
from ..utils import nvtx

c_p, c_i, c_ll = C.c_void_p, C.c_int, C.c_longlong
N.register_sigs({
    "b200_plan_buckets": [C.POINTER(c_ll), c_i, c_ll, c_ll, C.POINTER(c_i)],
    "b200_plan_buckets_by_group": [c_i, C.POINTER(c_i), c_i, C.POINTER(c_i)],
    "b200_reducer_create": [C.POINTER(c_i), c_i, c_i],
    "b200_reducer_destroy": [c_p],
    "b200_reducer_reset": [c_p],
    "b200_reducer_mark_ready": [c_p, c_i],
    "b200_reducer_all_done": [c_p],
    "b200_reducer_pending_buckets": [c_p, C.POINTER(c_i), c_i],
    "b200_reducer_completion_order": [c_p, C.POINTER(c_i), c_i],
}, {"b200_reducer_create": c_p, "b200_reducer_destroy": None, "b200_reducer_reset": None})

FIRST_BUCKET_BYTES = 1 << 20    # torch/nn/parallel/distributed.py:1262-1276
BUCKET_CAP_BYTES = 25 << 20     # torch/nn/parallel/distributed.py:831-833


def plan_buckets(param_bytes: Sequence[int], first_cap: int = FIRST_BUCKET_BYTES,
                 cap: int = BUCKET_CAP_BYTES) -> List[int]:
    """bucket index per parameter (registration order), torch-DDP rule, native planner."""
    n = len(param_bytes)
    arr = (c_ll * n)(*[int(b) for b in param_bytes])
    out = (c_i * n)()
    nb = N.lib().b200_plan_buckets(arr, n, first_cap, cap, out)
    if nb < 0:
        N.check(nb, "b200_plan_buckets")
    return list(out)


def plan_buckets_by_group(n_params: int, boundaries: Sequence[int]) -> List[int]:
    """One bucket per explicit parameter group (e.g. per layer), numbered back-to-front."""
    b = (c_i * len(boundaries))(*[int(x) for x in boundaries])
    out = (c_i * n_params)()
    nb = N.lib().b200_plan_buckets_by_group(n_params, b, len(boundaries), out)
    if nb < 0:
        N.check(nb, "b200_plan_buckets_by_group")
    return list(out)


class ReducerCore:
    """Thin RAII wrapper over the native ready-counting core."""

    def __init__(self, bucket_of: Sequence[int]):
        self.bucket_of = list(bucket_of)
        self.nbuckets = max(self.bucket_of) + 1
        arr = (c_i * len(self.bucket_of))(*self.bucket_of)
        self._h = N.lib().b200_reducer_create(arr, len(self.bucket_of), self.nbuckets)
        if not self._h:
            raise N.NativeError("b200_reducer_create failed: " +
                                (N.lib().b200_last_error() or b"").decode())

    def reset(self):
        N.lib().b200_reducer_reset(self._h)

    def mark_ready(self, i: int) -> int:
        return N.lib().b200_reducer_mark_ready(self._h, i)

    def all_done(self) -> bool:
        return bool(N.lib().b200_reducer_all_done(self._h))

    def pending_buckets(self) -> List[int]:
        out = (c_i * self.nbuckets)()
        k = N.lib().b200_reducer_pending_buckets(self._h, out, self.nbuckets)
        return list(out[:k])

    def completion_order(self) -> List[int]:
        out = (c_i * self.nbuckets)()
        k = N.lib().b200_reducer_completion_order(self._h, out, self.nbuckets)
        return list(out[:k])

    def __del__(self):
        try:
            if self._h:
                N.lib().b200_reducer_destroy(self._h)
                self._h = None
        except Exception:
            pass


class Reducer:
    """Bucketed, overlapped gradient all-reduce over a FlatParams layout.

    flat:     utils.flat.FlatParams whose param/grad buffers come from comm.alloc
    comm:     communicator
    bucket_of: bucket index per parameter; None = torch DDP's size rule
    fuse_sgd: (lr, momentum, weight_decay) -> the bucket kernel also updates the weights
              (then `optimizer.step()` must NOT be called for these parameters)
    """

    def __init__(self, flat, comm, bucket_of: Optional[Sequence[int]] = None, fuse_sgd=None,
                 algo="auto"):
        self.flat, self.comm, self.algo = flat, comm, algo
        esz = flat.param.element_size()
        a = max(1, 16 // esz)
        sizes = [((s.numel + a - 1) // a * a) * esz for s in flat.slots]
        self.bucket_of = list(bucket_of) if bucket_of is not None else plan_buckets(sizes)
        self.core = ReducerCore(self.bucket_of)
        # contiguous element range of every bucket inside the flat buffers
        self.ranges = []
        for b in range(self.core.nbuckets):
            idx = [i for i, x in enumerate(self.bucket_of) if x == b]
            if idx != list(range(idx[0], idx[-1] + 1)):
                raise ValueError("a bucket must hold consecutively registered parameters")
            lo = flat.slots[idx[0]].offset
            last = flat.slots[idx[-1]]
            hi = (last.offset + last.numel + a - 1) // a * a
            self.ranges.append((lo, hi))
        self.is_cuda = flat.param.is_cuda
        self.stream = torch.cuda.Stream(device=flat.param.device) if self.is_cuda else None
        self.fuse = fuse_sgd is not None
        self.update_target = None      # set by DataParallel(pingpong=True)
        self.launched: List[int] = []
        self.sync_enabled = True
        self.kernel_launches = 0
        self.pushed = {}        # bucket -> vectors already scattered by a fused GEMM epilogue
        # let wgrad GEMMs scatter straight into the owners' staging (B200DP_FUSE_RS=0 disables)
        self.fuse_rs = os.environ.get("B200DP_FUSE_RS", "1") != "0"
        if self.fuse:
            lr, mu, wd = fuse_sgd
            dev = flat.param.device
            self.hp = torch.tensor([lr, mu, wd, 0.0], dtype=torch.float32, device=dev)
            W, r = comm.world_size, comm.rank
            self.state = []
            for (lo, hi) in self.ranges:
                nbytes = (hi - lo) * esz
                two_shot = self._algo_for(nbytes) in (1, 2) and W > 1   # sharded optimizer state
                if two_shot:   # this rank owns slice r of the bucket: shard master + momentum
                    nvec = nbytes // 16
                    sl = (nvec + W - 1) // W
                    per_vec = 16 // esz
                    e_lo = min(hi - lo, r * sl * per_vec)
                    e_hi = min(hi - lo, (r + 1) * sl * per_vec)
                    master = flat.param[lo + e_lo:lo + e_hi].float().clone()
                    if master.numel() == 0:
                        master = torch.zeros(per_vec, dtype=torch.float32, device=dev)
                    st = (master, torch.zeros_like(master), r * sl * per_vec)
                else:          # every rank updates its full replica redundantly
                    if flat.param.dtype == torch.float32:
                        master = flat.param[lo:hi]
                    else:
                        master = flat.param[lo:hi].float().clone()
                    st = (master, torch.zeros(hi - lo, dtype=torch.float32, device=dev), 0)
                self.state.append(st)

    def _algo_for(self, nbytes: int) -> int:
        if self.algo == "auto":
            if not hasattr(self.comm, "choose_algo"):
                return 0
            nvls_ok = hasattr(self.comm, "nvls_offset_of") and \
                self.comm.nvls_offset_of(self.flat.grad) is not None
            return self.comm.choose_algo(nbytes, nvls_ok)
        return {"oneshot": 0, "twoshot": 1, "nvls": 2}[self.algo]

    def set_lr(self, lr: float):
        self.hp[0] = lr

    # -- one backward pass ------------------------------------------------------
    def prepare_for_backward(self):
        self.core.reset()
        self.launched = []
        self.pushed = {}

    def rs_target(self, param_index: int):
        """If parameter i's gradient may be scattered to the owner ranks by the GEMM that
        produces it (fused GEMM -> reduce-scatter), return the b200_gemm_bf16_rs arguments.
        Conditions: CUDA, world > 1, two-shot algorithm, and the parameter sits at the start
        of its bucket with a whole number of 16-byte vectors."""
        if not (self.fuse_rs and self.is_cuda and self.sync_enabled and self.comm.world_size > 1
                and hasattr(self.comm, "rs_target")):
            return None
        b = self.bucket_of[param_index]
        lo, hi = self.ranges[b]
        slot = self.flat.slots[param_index]
        esz = self.flat.grad.element_size()
        nbytes = (hi - lo) * esz
        if slot.offset != lo or (slot.numel * esz) % 16 or self._algo_for(nbytes) != 1 \
                or self.flat.grad.dtype != torch.bfloat16:
            return None
        t = dict(self.comm.rs_target(nbytes, channel=b + 1))
        t["base_vec"] = 0
        t["_bucket"] = b
        t["_vecs"] = slot.numel * esz // 16
        return t

    def mark_pushed(self, target: dict):
        self.pushed[target["_bucket"]] = target["_vecs"]

    def mark_ready(self, param_index: int):
        """Called when parameter i's gradient is final (autograd hook or layer backward)."""
        if not self.sync_enabled:
            return
        b = self.core.mark_ready(param_index)
        if b == -2:
            raise RuntimeError(f"parameter {param_index} marked ready twice in one backward "
                               "(gradient accumulation needs no_sync())")
        if b >= 0:
            self._launch_bucket(b)

    def _launch_bucket(self, b: int):
        lo, hi = self.ranges[b]
        g = self.flat.grad[lo:hi]
        nbytes = (hi - lo) * g.element_size()
        algo = self._algo_for(nbytes)
        nvtx.push(f"bucket {b} all-reduce ({nbytes} B, algo {algo})")
        if self.is_cuda:
            self.stream.wait_stream(torch.cuda.current_stream())
            ctx = torch.cuda.stream(self.stream)
        else:
            import contextlib
            ctx = contextlib.nullcontext()
        with ctx:
            if self.fuse and self.is_cuda:
                master, mom, base = self.state[b]
                # ping-pong mode: the updated weights go to the buffer the step does NOT read
                tgt = self.update_target if self.update_target is not None else self.flat.param
                self.comm.all_reduce(g, op="mean", algo=algo, channel=b + 1,
                                     push_from=self.pushed.get(b, 0),
                                     sgd=(tgt[lo:hi], master, mom, self.hp, base))
            elif self.is_cuda:
                self.comm.all_reduce(g, op="mean", algo=algo, channel=b + 1,
                                     push_from=self.pushed.get(b, 0))
            else:
                self.comm.all_reduce(g, op="mean", algo="auto")
                if self.fuse:   # CPU/gloo path: plain SGD on the averaged bucket
                    master, mom, _ = self.state[b]
                    lr, mu, wd, _n = self.hp.tolist()
                    gg = g.float() + wd * master
                    mom.mul_(mu).add_(gg)
                    master.sub_(lr * mom)
                    if master.data_ptr() != self.flat.param[lo:hi].data_ptr():
                        self.flat.param[lo:hi].copy_(master)
        nvtx.pop()
        self.kernel_launches += 1
        self.launched.append(b)

    def finalize_backward(self):
        """Launch anything still pending (unused parameters contribute zeros) and make the
        compute stream wait for the reductions (torch DDP's finalize_backward)."""
        if not self.sync_enabled:
            return
        for b in self.core.pending_buckets():
            if b not in self.launched:
                self._launch_bucket(b)
        if self.is_cuda:
            torch.cuda.current_stream().wait_stream(self.stream)
