"""Persistent training engine for `Net`: many optimizer steps per kernel launch.

Reference loop (src/train_dist.py:78-89): per batch `zero_grad -> model(data) -> loss ->
backward -> [DDP all-reduce overlapped with backward, :63,:83] -> optimizer.step ->
loss.item()`. `PersistentNetEngine.run_steps(..., nsteps)` executes `nsteps` of those bodies in
ONE cooperative launch of `net_persist_kernel` (csrc/net_persist.cu): worker CTAs do
forward/backward out of shared memory, service CTAs form the fc gradients as a batch GEMM,
sum the conv partial rows, exchange every bucket with the peer GPUs as soon as it is final
(fc first, conv2 next, conv1 last), apply SGD-momentum and publish the new weights, all
while the workers are still in backward. No kernel boundary, graph or host work between steps.

`HostFeeder` (the e2e path) streams batches from pinned host memory into a device ring on a
copy stream and tells the running kernel, step by step, what has landed; the kernel writes
each step's loss straight into the caller's pinned buffer.

The two-kernel `FusedNetEngine.step` stays available (single steps, ragged last batches,
gradient-only calls inside CUDA graphs); both paths share parameters, momentum, step
counter, loss ring, dropout streams and the peer staging, so they can be mixed freely.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from .. import _native as N
from .fused import FusedNetEngine, X_F32, X_U8, X_U8_GATHER, X_F32_GATHER  # noqa: F401

c_p, c_i, c_u, c_sz, c_ull, c_ll = (C.c_void_p, C.c_int, C.c_uint, C.c_size_t, C.c_ulonglong,
                                    C.c_longlong)


class _PersistHost(C.Structure):   # mirrors B200PersistHost in csrc/net_persist.cu
    _fields_ = [("params", c_p), ("mom", c_p), ("grad_out", c_p), ("hp", c_p),
                ("x", c_p), ("labels", c_p), ("gidx", c_p),
                ("x_mode", c_i), ("B", c_i), ("nsteps", c_i), ("update", c_i),
                ("idx_stride", c_i), ("n_idx", c_i),
                ("ring_slots", c_i), ("ring_pos", c_u),
                ("feed_ready", c_p), ("feed_consumed", c_p), ("feed_base", c_u), ("mask_mode", c_i),
                ("scratch", c_p),
                ("loss_ring", c_p), ("ring", c_i), ("multi", c_i),
                ("host_loss", c_p),
                ("step_ctr", c_p),
                ("seed", c_ull),
                ("sample_base", c_u), ("pad0", c_i),
                ("m2buf", c_p), ("m3buf", c_p),
                ("comm", c_p),
                ("stage_off", c_sz),
                ("seq_ctr", c_p),
                ("err", c_p),
                ("timeout_ns", c_ull),
                ("dbg", c_p),
                ("label_slots", c_i), ("label_pos", c_u),
                ("epoch_begin", c_i), ("pad1", c_i)]


N.register_sigs({
    "b200_net_persist_host_size": [],
    "b200_net_persist_plan": [c_i, C.POINTER(c_i)],
    "b200_net_persist_scratch_floats": [c_i],
    "b200_net_persist_smem_bytes": [],
    "b200_net_persist_dbg_words": [C.POINTER(c_i), C.POINTER(c_i)],
    "b200_net_persist_run": [C.POINTER(_PersistHost), c_p],
    "b200_feeder_create": [C.POINTER(_PersistHost), c_i, C.POINTER(c_p)],
    "b200_feeder_run": [c_p, c_p, c_p, c_i, c_p, c_p],
    "b200_feeder_launches": [c_p],
    "b200_feeder_destroy": [c_p],
}, {"b200_net_persist_scratch_floats": c_ll, "b200_feeder_launches": c_ll,
    "b200_net_persist_dbg_words": c_ll})


def persist_plan(batch: int) -> dict:
    """How the persistent kernel lays a batch out over the SMs (host-only query)."""
    o = (c_i * 9)()
    N.check(N.lib().b200_net_persist_plan(int(batch), o), "b200_net_persist_plan")
    keys = ("grid", "workers", "split", "rows", "images_per_worker", "service", "conv2_ctas",
            "w3_ctas", "w3_cols")
    return dict(zip(keys, list(o)))


class PersistentNetEngine(FusedNetEngine):
    """FusedNetEngine + `run_steps`: nsteps optimizer steps in one persistent launch."""

    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        if C.sizeof(_PersistHost) != N.lib().b200_net_persist_host_size():
            raise N.NativeError("B200PersistHost layout mismatch between Python and the library")
        self._pscratch = {}
        self._perr = torch.zeros(4, dtype=torch.int32, device=self.device)
        self._feeders = {}
        self._hcache = {}
        self._dbg = None
        if self.comm is not None and self.comm.world_size > 1:
            self.comm.net_channel()      # staging + sequence counter shared with kernel B

    # ------------------------------------------------------------------ plumbing
    def _scratch_for(self, B: int) -> torch.Tensor:
        s = self._pscratch.get(B)
        if s is None:
            n = int(N.lib().b200_net_persist_scratch_floats(B))
            s = self._pscratch[B] = torch.zeros(n, dtype=torch.float32, device=self.device)
        return s

    def _host_args(self, B: int, x_mode: int, mask_mode: int) -> _PersistHost:
        h = _PersistHost()
        h.params, h.mom = self.flat.param.data_ptr(), self.flat.momentum.data_ptr()
        h.hp = self.hp.data_ptr()
        h.x_mode, h.B, h.update = x_mode, B, 1
        h.scratch = self._scratch_for(B).data_ptr()
        h.loss_ring, h.ring = self.loss_ring.data_ptr(), self.ring
        h.step_ctr = self.step_ctr.data_ptr()
        h.seed, h.sample_base, h.mask_mode = self.seed, self.sample_base, mask_mode
        h.err = self._perr.data_ptr()
        h.dbg = self._dbg.data_ptr() if self._dbg is not None else None
        comm = self.comm
        if comm is not None and comm.world_size > 1:
            n = comm.net_channel()
            h.multi, h.comm = 1, C.cast(comm._blob, c_p)
            h.stage_off, h.seq_ctr = n["stage"], n["seq"]
        return h

    def _mask_mode(self, masks) -> int:
        if masks is not None:
            return 2
        return 1 if (self.dropout and self.net.training) else 0

    # ------------------------------------------------------------------ steps
    def run_steps(self, x, y, x_mode=X_F32, gidx=None, batch=None, nsteps: int = 1,
                  idx_stride: int = 0, masks=None, write_grad: bool = False, update: bool = True,
                  begin_epoch: bool = False):
        """`nsteps` optimizer steps on the current stream, one launch, no host sync.

        Gather modes (X_U8_GATHER / X_F32_GATHER): x, y are the device-resident dataset and
        step s of the epoch reads gidx[(step - epoch_start) * idx_stride + i] exactly like
        `step(..., idx_stride=...)`; call `begin_epoch()` first.
        Direct modes (X_F32 / X_U8): x is [nsteps * batch, 784], y [nsteps * batch]; step s
        reads rows [s * batch, (s + 1) * batch).
        write_grad: the averaged gradient of the last step goes to `flat.grad`;
        update=False leaves parameters and momentum untouched (nsteps must be 1);
        begin_epoch=True marks the first of these steps as the start of an epoch inside the
        launch (same effect as calling `begin_epoch()` first, without the extra device op)."""
        gather = x_mode in (X_U8_GATHER, X_F32_GATHER)
        if batch is None:
            batch = (gidx.shape[0] if gather else x.shape[0] // max(1, nsteps))
        B = int(batch)
        if not update and nsteps != 1:
            raise N.NativeError("update=False only makes sense for a single step")
        mm = self._mask_mode(masks)
        key = (B, x_mode, mm, self._dbg is None)
        h = self._hcache.get(key)            # the argument block is built once per shape
        if h is None:
            h = self._hcache[key] = self._host_args(B, x_mode, mm)
        h.epoch_begin = 1 if begin_epoch else 0
        h.m2buf = h.m3buf = None
        h.x, h.labels, h.gidx = N.ptr(x), N.ptr(y), N.ptr(gidx)
        h.nsteps, h.update = int(nsteps), 1 if update else 0
        h.grad_out = self.flat.grad.data_ptr() if write_grad else None
        h.idx_stride = int(idx_stride)
        h.n_idx = int(gidx.numel()) if gidx is not None else 0
        h.ring_slots, h.ring_pos = (0, 0) if gather else (int(nsteps), 0)
        if masks is not None:
            h.m2buf, h.m3buf = masks[0].data_ptr(), masks[1].data_ptr()
        N.check(N.lib().b200_net_persist_run(C.byref(h), N.stream()), "b200_net_persist_run")
        self.launches += 1
        return B

    def enable_phase_stamps(self, on: bool = True):
        """Debug aid: per-CTA, per-step clock64/globaltimer stamps of the persistent kernel's
        phases (scripts/persist_phases.py turns them into a timeline). Returns the buffer
        viewed as [148, steps, phases, 2] int64."""
        self._hcache = {}
        if not on:
            self._dbg = None
            return None
        st, ph = c_i(0), c_i(0)
        n = int(N.lib().b200_net_persist_dbg_words(C.byref(st), C.byref(ph)))
        self._dbg = torch.zeros(n, dtype=torch.int64, device=self.device)
        return self._dbg.view(148, st.value, ph.value, 2)

    def check_errors(self):
        """Raise if a bounded wait inside the persistent kernel timed out (host sync)."""
        if self.comm is not None and self.comm.world_size > 1:
            self.comm.check_errors()
        e = self._perr.cpu().tolist()
        if e[0] != 0:
            raise N.NativeError(f"persistent step kernel aborted: wait #{e[2]} timed out "
                                f"(peer {e[1]}, sequence {e[3] & 0xffffffff})")

    # ------------------------------------------------------------------ host feed
    def feeder(self, batch: int, x_dtype=torch.float32, ring_slots: int = 32) -> "HostFeeder":
        key = (batch, x_dtype, ring_slots)
        f = self._feeders.get(key)
        if f is None:
            f = self._feeders[key] = HostFeeder(self, batch, x_dtype, ring_slots)
        return f


class HostFeeder:
    """Streams batches from pinned host memory into a running persistent kernel."""

    def __init__(self, engine: PersistentNetEngine, batch: int, x_dtype=torch.float32,
                 ring_slots: int = 32):
        self.engine, self.batch, self.x_dtype = engine, int(batch), x_dtype
        mode = X_U8 if x_dtype == torch.uint8 else X_F32
        mask_mode = 1 if (engine.dropout and engine.net.training) else 0
        self._tmpl = engine._host_args(self.batch, mode, mask_mode)
        h = c_p()
        N.check(N.lib().b200_feeder_create(C.byref(self._tmpl), int(ring_slots), C.byref(h)),
                "b200_feeder_create")
        self._h = h
        self._pinned_ok = set()
        self.kernels_per_run = 1

    def run(self, hx: torch.Tensor, hy: torch.Tensor, steps: int, hloss: torch.Tensor = None):
        """hx [steps*batch, 784] and hy [steps*batch] int64 in PINNED host memory; hloss pinned
        float32 [steps] or None. Asynchronous: synchronise before reading hloss."""
        # is_pinned() is a driver query per tensor: remember storages that passed
        for t in (hx, hy, hloss):
            if t is None:
                continue
            key = (t.untyped_storage().data_ptr(), t.untyped_storage().nbytes())
            if key not in self._pinned_ok:
                if not t.is_pinned():
                    raise N.NativeError("HostFeeder.run needs pinned host tensors")
                if len(self._pinned_ok) > 64:
                    self._pinned_ok.clear()
                self._pinned_ok.add(key)
        if hx.dtype != self.x_dtype or hy.dtype != torch.int64:
            raise N.NativeError("HostFeeder.run: dtype mismatch")
        if hx.shape[0] < steps * self.batch or hy.shape[0] < steps * self.batch:
            raise N.NativeError("HostFeeder.run: not enough samples for the requested steps")
        if hloss is not None and (hloss.dtype != torch.float32 or hloss.numel() < steps):
            raise N.NativeError("HostFeeder.run: loss buffer must be float32 with >= steps elements")
        N.check(N.lib().b200_feeder_run(self._h, hx.data_ptr(), hy.data_ptr(), int(steps),
                                        N.ptr(hloss), N.stream()), "b200_feeder_run")
        self.engine.launches += 1

    def launches(self) -> int:
        return int(N.lib().b200_feeder_launches(self._h)) if self._h else 0

    def close(self):
        if self._h:
            N.lib().b200_feeder_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def persistent_enabled() -> bool:
    """B200DP_PERSIST=0 routes the trainer back to the two-kernel graphs."""
    return os.environ.get("B200DP_PERSIST", "1") != "0"
