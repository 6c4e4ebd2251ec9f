"""Python handle on the native step runner (csrc/runner.cu).

`NativeRunner(engine, batch, x_dtype)` captures the engine's fused step once per staging
slot inside the native library and `run(hx, hy, steps, hloss)` submits whole sequences of
steps from pinned host memory with copy/compute overlap and no Python per step.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native as N

c_p, c_i, c_u, c_sz, c_ull = C.c_void_p, C.c_int, C.c_uint, C.c_size_t, C.c_ulonglong

N.register_sigs({
    "b200_runner_config_size": [],
    "b200_runner_create": [c_p, C.POINTER(c_p)],
    "b200_runner_run": [c_p, c_p, c_p, c_i, c_p, c_p],
    "b200_runner_destroy": [c_p],
})


class _RunnerConfig(C.Structure):
    _fields_ = [("params", c_p), ("mom", c_p), ("partial", c_p), ("loss_partial", c_p),
                ("hp", c_p), ("loss_ring", c_p), ("ctl", c_p), ("ring", c_i), ("batch", c_i),
                ("x_mode", c_i), ("mask_mode", c_i), ("seed", c_ull), ("sample_base", c_u),
                ("multi", c_i), ("comm", c_p), ("stage_off", c_sz), ("flag_off", c_sz),
                ("seq_ctr", c_p), ("arrive_ctr", c_p), ("protocol", c_i), ("pad_", c_i)]


class NativeRunner:
    def __init__(self, engine, batch: int, x_dtype=torch.float32):
        lib = N.lib()
        if C.sizeof(_RunnerConfig) != lib.b200_runner_config_size():
            raise N.NativeError("RunnerConfig layout mismatch between Python and the library")
        self.engine = engine
        self.batch = batch
        self.x_dtype = x_dtype
        eng = engine
        # the kernels must have run once outside capture; leave the training state untouched
        snap = eng._snapshot()
        x = torch.zeros(batch, 784, dtype=x_dtype, device=eng.device)
        y = torch.zeros(batch, dtype=torch.int64, device=eng.device)
        eng.step(x, y, 1 if x_dtype == torch.uint8 else 0, batch=batch)
        eng._restore(snap)
        torch.cuda.synchronize(eng.device)
        cfg = _RunnerConfig()
        cfg.params, cfg.mom = eng.flat.param.data_ptr(), eng.flat.momentum.data_ptr()
        cfg.partial, cfg.loss_partial = eng.partial.data_ptr(), eng.loss_partial.data_ptr()
        cfg.hp, cfg.loss_ring, cfg.ctl = eng.hp.data_ptr(), eng.loss_ring.data_ptr(), eng.step_ctr.data_ptr()
        cfg.ring, cfg.batch = eng.ring, batch
        cfg.x_mode = 1 if x_dtype == torch.uint8 else 0
        cfg.mask_mode = 1 if (eng.dropout and eng.net.training) else 0
        cfg.seed, cfg.sample_base = eng.seed, eng.sample_base
        comm = eng.comm
        if comm is not None and comm.world_size > 1:
            comm.host_barrier()
            n = comm._net
            cfg.multi, cfg.comm = 1, C.cast(comm._blob, c_p)
            cfg.stage_off, cfg.flag_off = n["stage"], n["flags"]
            cfg.seq_ctr, cfg.arrive_ctr = n["seq"], n["arrive"]
            import os
            cfg.protocol = int(os.environ.get("B200DP_NET_PROTOCOL", "0"))
        self._cfg = cfg
        h = c_p()
        N.check(lib.b200_runner_create(C.byref(cfg), C.byref(h)), "b200_runner_create")
        self._h = h
        self.kernels_per_step = 2

    def run(self, hx: torch.Tensor, hy: torch.Tensor, steps: int, hloss: torch.Tensor = None):
        """hx [steps*batch, 784] and hy [steps*batch] int64 in PINNED host memory; hloss
        pinned float32 [steps] or None. Asynchronous: synchronise before reading hloss."""
        if not (hx.is_pinned() and hy.is_pinned()) or (hloss is not None and not hloss.is_pinned()):
            raise N.NativeError("NativeRunner.run needs pinned host tensors")
        if hx.dtype != self.x_dtype or hy.dtype != torch.int64:
            raise N.NativeError("NativeRunner.run: dtype mismatch")
        if hx.shape[0] < steps * self.batch or hy.shape[0] < steps * self.batch:
            raise N.NativeError("NativeRunner.run: not enough samples for the requested steps")
        N.check(N.lib().b200_runner_run(self._h, hx.data_ptr(), hy.data_ptr(), steps,
                                        N.ptr(hloss), N.stream()), "b200_runner_run")
        self.engine.launches += steps * self.kernels_per_step

    def chunks_launched(self) -> int:
        """How many 16-step graphs (copies as graph nodes) have been launched; the remaining
        steps went through the per-step submission path."""
        f = N.lib().b200_runner_chunks_launched
        f.restype, f.argtypes = C.c_longlong, [c_p]
        return int(f(self._h)) if self._h else 0

    def close(self):
        if self._h:
            N.lib().b200_runner_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
