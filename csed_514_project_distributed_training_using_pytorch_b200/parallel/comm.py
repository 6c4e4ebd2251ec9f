"""Communicators: the distributed runtime.

Reference: `torch.distributed.init_process_group(backend="gloo", rank, world_size)` with
MASTER_ADDR/MASTER_PORT rendezvous (src/train_dist.py:144-146, src/run1.py:19-23), then
implicit broadcast/all-reduce through DDP and explicit `dist.send/recv` (src/run1.py:13,16).

Two implementations behind one interface:

* `PeerComm` (CUDA): one process per GPU. Every rank cudaMallocs a symmetric arena, the
  cudaIpc handles are exchanged over the bootstrap store, and every rank maps every arena.
  Collectives are hand-written kernels (csrc/comm.cu) that store into peers' memory over
  NVLink/NVSwitch and synchronise with release/acquire flags. torch.distributed is used
  only for bootstrap (rendezvous + handle exchange), never for tensor data.
* `GlooComm` (CPU): the same interface on torch.distributed/gloo, so sampler, reducer,
  checkpoint and launcher logic run on a machine without a GPU.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional

import torch
import torch.distributed as dist

from .. import _native as N

c_p, c_i, c_u, c_f, c_sz = C.c_void_p, C.c_int, C.c_uint, C.c_float, C.c_size_t
c_ull = C.c_ulonglong

N.register_sigs({
    "b200_symm_alloc": [c_sz, C.POINTER(c_p), C.c_char_p],
    "b200_symm_open": [C.c_char_p, C.POINTER(c_p)],
    "b200_symm_close": [c_p],
    "b200_symm_free": [c_p],
    "b200_device_can_access_peer": [c_i, c_i, C.POINTER(c_i)],
    "b200_comm_struct_size": [],
    "b200_comm_make": [c_p, c_i, c_i, C.POINTER(c_p), c_p, c_ull],
    "b200_comm_barrier": [c_p, c_sz, c_p, c_p],
    "b200_comm_broadcast": [c_p, c_i, c_sz, c_sz, c_sz, c_p, c_p, c_p],
    "b200_comm_send": [c_p, c_i, c_p, c_sz, c_sz, c_sz, c_sz, c_sz, c_p, c_p],
    "b200_comm_recv": [c_p, c_i, c_p, c_sz, c_sz, c_sz, c_sz, c_sz, c_p, c_p],
    "b200_comm_sendrecv": [c_p, c_i, c_p, c_sz, c_i, c_p, c_sz, c_sz, c_sz, c_sz, c_sz, c_p, c_p, c_p],
    "b200_nvls_supported": [c_i, C.POINTER(c_i)],
    "b200_nvls_granularity": [c_i, c_sz, C.POINTER(c_sz)],
    "b200_nvls_create": [c_i, c_sz, C.POINTER(c_ull), C.POINTER(c_i)],
    "b200_nvls_import": [c_i, C.POINTER(c_ull)],
    "b200_nvls_add_device": [c_ull, c_i],
    "b200_nvls_bind_map": [c_ull, c_i, c_sz, C.POINTER(c_p), C.POINTER(c_p)],
    "b200_nvls_allreduce": [c_p, c_p, c_sz, c_sz, c_i, c_sz, c_i, c_p, c_p, c_f, c_i, c_sz, c_p, c_p,
                            c_p, c_sz, c_i, c_p],
    "b200_comm_net_stage_bytes": [],
    "b200_comm_net_flag_bytes": [],
    "b200_comm_net_reduce_allreduce_sgd": [c_p, c_p, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_p,
                                           c_sz, c_sz, c_p, c_p, c_i, c_p],
    "b200_comm_allreduce": [c_p, c_i, c_i, c_i, c_p, c_p, c_sz, c_sz, c_sz, c_sz, c_sz, c_i, c_p,
                            c_p, c_f, c_p, c_p, c_p, c_sz, c_i, c_sz, c_p],
}, {"b200_comm_net_stage_bytes": c_sz, "b200_comm_net_flag_bytes": c_sz})

MAX_WORLD = 8
_DTYPE_CODE = {torch.float32: 0, torch.bfloat16: 1}


class CommError(RuntimeError):
    pass


_CROSSOVER = None


def _crossover_table() -> dict:
    """profiles/allreduce_crossover.json (written by scripts/sweep_report.py from the measured
    sweep): per world size, the largest payload for which the one-shot kernel won."""
    global _CROSSOVER
    if _CROSSOVER is None:
        import json
        from pathlib import Path
        p = Path(__file__).resolve().parents[2] / "profiles" / "allreduce_crossover.json"
        try:
            _CROSSOVER = json.loads(p.read_text())
        except Exception:
            _CROSSOVER = {}
    return _CROSSOVER


def choose_from_table(world: int, nbytes: int, nvls_ok: bool = False) -> int:
    """Algorithm for a bucket of `nbytes` on `world` ranks from the measured table
    (profiles/allreduce_crossover.json, written from the sweep): 0 tagged one-shot while the
    bytes every rank must absorb stay small, 2 NVLS (reduction inside the NVSwitch) in the
    range where it won and only if the bucket lives in the multicast region (`nvls_ok`),
    1 two-shot otherwise. World sizes without a measured row use the nearest measured one."""
    if "B200DP_ONESHOT_MAX_KB" not in os.environ:
        tab_all = {k: v for k, v in _crossover_table().items() if k.isdigit()}
        tab = tab_all.get(str(world))
        if tab is None and tab_all:
            tab = tab_all[str(min((int(k) for k in tab_all), key=lambda k: abs(k - world)))]
        if tab is not None:
            if nbytes <= tab["oneshot_max_bytes"]:
                return 0
            if nvls_ok and "nvls_min_bytes" in tab and \
                    tab["nvls_min_bytes"] <= nbytes <= tab.get("nvls_max_bytes", 1 << 62):
                return 2
            return 1
    budget = int(os.environ.get("B200DP_ONESHOT_MAX_KB", "256")) << 10
    return 0 if nbytes * max(1, world - 1) <= budget else 1


class _RawCudaMemory:
    """Exposes a raw device allocation to torch through __cuda_array_interface__."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1",
                                         "data": (ptr, False), "version": 3, "strides": None}


class BaseComm:
    rank: int
    world_size: int
    device: torch.device

    def alloc(self, numel: int, dtype=torch.float32, device=None) -> torch.Tensor:
        raise NotImplementedError

    def barrier(self): raise NotImplementedError
    def broadcast(self, tensor, root=0): raise NotImplementedError
    def all_reduce(self, tensor, op="mean", algo="auto"): raise NotImplementedError
    def send(self, tensor, dst): raise NotImplementedError
    def recv(self, tensor, src): raise NotImplementedError
    def check_errors(self): pass
    def close(self): pass

    def host_barrier(self):
        if dist.is_initialized() and self.world_size > 1:
            dist.barrier(group=self._cpu_group())

    def _cpu_group(self):
        return getattr(self, "cpu_group", None)

    def all_gather_object(self, obj):
        if self.world_size == 1 or not dist.is_initialized():
            return [obj]
        out = [None] * self.world_size
        dist.all_gather_object(out, obj, group=self._cpu_group())
        return out

    def max_over_ranks(self, value: float) -> float:
        return max(self.all_gather_object(float(value)))


class GlooComm(BaseComm):
    """CPU tensors over torch.distributed/gloo (the reference's own backend)."""

    def __init__(self, rank: int, world_size: int, cpu_group=None):
        self.rank, self.world_size = rank, world_size
        self.device = torch.device("cpu")
        self.cpu_group = cpu_group

    def alloc(self, numel, dtype=torch.float32, device=None):
        return torch.zeros(numel, dtype=dtype)

    def barrier(self):
        self.host_barrier()

    def broadcast(self, tensor, root=0):
        if self.world_size > 1:
            dist.broadcast(tensor, src=root, group=self.cpu_group)
        return tensor

    def all_reduce(self, tensor, op="mean", algo="auto"):
        if self.world_size > 1:
            dist.all_reduce(tensor, op=dist.ReduceOp.SUM, group=self.cpu_group)
        if op == "mean":
            tensor.div_(self.world_size)
        return tensor

    def send(self, tensor, dst):
        dist.send(tensor, dst=dst, group=self.cpu_group)

    def recv(self, tensor, src):
        dist.recv(tensor, src=src, group=self.cpu_group)
        return tensor

    def sendrecv(self, send_tensor, dst, recv_tensor, src):
        reqs = [dist.isend(send_tensor, dst=dst, group=self.cpu_group),
                dist.irecv(recv_tensor, src=src, group=self.cpu_group)]
        for r in reqs:
            r.wait()
        return recv_tensor


class PeerComm(BaseComm):
    """Peer-memory communicator for one GPU per process on an NVSwitch box."""

    def __init__(self, rank: int, world_size: int, device: torch.device, cpu_group=None,
                 arena_bytes: Optional[int] = None, timeout_s: Optional[float] = None):
        if world_size > MAX_WORLD:
            raise CommError(f"world_size {world_size} > {MAX_WORLD} (one NVSwitch box)")
        self.rank, self.world_size, self.device = rank, world_size, torch.device(device)
        self.cpu_group = cpu_group
        self.lib = N.lib()
        torch.cuda.set_device(self.device)
        if arena_bytes is None:
            arena_bytes = int(os.environ.get("B200DP_ARENA_MB", "512")) << 20
        if timeout_s is None:
            timeout_s = float(os.environ.get("B200DP_COMM_TIMEOUT_S", "20"))
        self.arena_bytes = arena_bytes
        base = c_p()
        handle = C.create_string_buffer(64)
        N.check(self.lib.b200_symm_alloc(arena_bytes, C.byref(base), handle), "b200_symm_alloc")
        self.local_base = base.value
        handles = self.all_gather_object(handle.raw)
        self.bases = []
        for r, h in enumerate(handles):
            if r == rank:
                self.bases.append(self.local_base)
            else:
                p = c_p()
                N.check(self.lib.b200_symm_open(h, C.byref(p)), f"b200_symm_open(rank {r})")
                self.bases.append(p.value)
        self._arena = torch.as_tensor(_RawCudaMemory(self.local_base, arena_bytes),
                                      device=self.device)
        self._top = 0
        # local (non-symmetric) control words
        self.err = torch.zeros(4, dtype=torch.int32, device=self.device)
        self._ctr = torch.zeros(256, dtype=torch.int32, device=self.device)
        self._ctr_top = 0
        self._blob = C.create_string_buffer(self.lib.b200_comm_struct_size())
        arr = (c_p * world_size)(*self.bases)
        N.check(self.lib.b200_comm_make(self._blob, rank, world_size, arr, self.err.data_ptr(),
                                        int(timeout_s * 1e9)), "b200_comm_make")
        # fixed regions every communicator has
        self._bar_flags = self._region(MAX_WORLD * 4)
        self._bar_seq = self._counter()
        self._bc_flags = self._region(2 * MAX_WORLD * 4)
        self._bc_seq, self._bc_arrive = self._counter(), self._counter()
        self.box_bytes = 1 << 20
        self._box = self._region(MAX_WORLD * self.box_bytes)
        self._p2p_flags = self._region(MAX_WORLD * 4)
        self._p2p_acks = self._region(MAX_WORLD * 4)
        self._sent = torch.zeros(MAX_WORLD, dtype=torch.int32, device=self.device)
        self._recvd = torch.zeros(MAX_WORLD, dtype=torch.int32, device=self.device)
        self._nvls = None     # lazily: NVLS (multicast) region, False if unsupported
        self._net = None      # lazily: staging/flags of the fused Net kernel B
        self._ar = None       # lazily: generic all-reduce channel
        self._bounce: Dict[int, torch.Tensor] = {}
        torch.cuda.synchronize(self.device)
        self.host_barrier()   # every arena mapped and zeroed before anyone signals

    # -- arena ---------------------------------------------------------------
    def _region(self, nbytes: int, align: int = 256) -> int:
        off = (self._top + align - 1) // align * align
        if off + nbytes > self.arena_bytes:
            raise CommError(f"symmetric arena exhausted ({self.arena_bytes} B): raise "
                            "B200DP_ARENA_MB")
        self._top = off + nbytes
        return off

    def _counter(self) -> int:
        """Device address of a fresh local u32 counter."""
        i = self._ctr_top
        self._ctr_top += 1
        if self._ctr_top > self._ctr.numel():
            raise CommError("out of control counters")
        return self._ctr.data_ptr() + 4 * i

    def alloc(self, numel: int, dtype=torch.float32, device=None) -> torch.Tensor:
        """A zeroed 1-D tensor inside the symmetric arena. Every rank must make the same
        sequence of alloc calls, so the same offset names the same buffer on all ranks."""
        esz = torch.empty((), dtype=dtype).element_size()
        off = self._region(numel * esz)
        t = self._arena[off:off + numel * esz].view(dtype)
        t.zero_()
        return t

    def offset_of(self, t: torch.Tensor) -> Optional[int]:
        off = t.data_ptr() - self.local_base
        if 0 <= off and off + t.numel() * t.element_size() <= self.arena_bytes:
            return off
        return None

    # -- health --------------------------------------------------------------
    def check_errors(self):
        """Raise if a kernel's bounded wait on a peer timed out (host sync)."""
        e = self.err.cpu().tolist()
        if e[0] == 2:    # the persistent step kernel gave up on one of its own counters
            raise CommError(f"rank {self.rank}: persistent step kernel aborted, internal wait "
                            f"#{e[2]} timed out (sequence {e[3] & 0xffffffff})")
        if e[0] != 0:
            where = {1: "barrier", 2: "broadcast(root waits)", 3: "broadcast(wait root)",
                     4: "send(wait ack)", 5: "recv", 6: "net all-reduce", 7: "one-shot all-reduce",
                     8: "two-shot reduce-scatter", 9: "two-shot all-gather",
                     10: "NVLS all-reduce (inputs ready)", 11: "NVLS all-reduce (results)",
                     12: "persistent step kernel (gradient exchange)"
                     }.get(e[2], str(e[2]))
            raise CommError(f"rank {self.rank}: timed out waiting for rank {e[1]} in {where} "
                            f"(sequence {e[3] & 0xffffffff})")

    # -- collectives ---------------------------------------------------------
    def barrier(self):
        N.check(self.lib.b200_comm_barrier(self._blob, self._bar_flags, self._bar_seq, N.stream()),
                "b200_comm_barrier")

    def broadcast(self, tensor: torch.Tensor, root: int = 0) -> torch.Tensor:
        """In-place broadcast of root's tensor (K16). Arena tensors are read by the peers
        directly; others go through a symmetric bounce buffer."""
        if self.world_size == 1:
            return tensor
        nbytes = tensor.numel() * tensor.element_size()
        off = self.offset_of(tensor) if tensor.is_contiguous() else None
        if off is not None and off % 16 == 0 and nbytes % 16 == 0:
            N.check(self.lib.b200_comm_broadcast(self._blob, root, off, nbytes, self._bc_flags,
                                                 self._bc_seq, self._bc_arrive, N.stream()),
                    "b200_comm_broadcast")
            return tensor
        pad = (nbytes + 15) // 16 * 16
        bounce = self._bounce_buffer(pad)
        flat = bounce[:nbytes].view(tensor.dtype)
        if self.rank == root:
            flat.copy_(tensor.reshape(-1))
        N.check(self.lib.b200_comm_broadcast(self._blob, root, self.offset_of(bounce), pad,
                                             self._bc_flags, self._bc_seq, self._bc_arrive,
                                             N.stream()), "b200_comm_broadcast")
        if self.rank != root:
            tensor.copy_(flat.view(tensor.shape))
        return tensor

    def _bounce_buffer(self, nbytes: int) -> torch.Tensor:
        key = 1 << max(12, (nbytes - 1).bit_length())
        b = self._bounce.get(key)
        if b is None:
            b = self._bounce[key] = self.alloc(key, torch.uint8)
        return b

    def send(self, tensor: torch.Tensor, dst: int):
        """Blocking-in-stream send (K20): payload into dst's mailbox, then a release flag."""
        buf, nbytes = self._p2p_payload(tensor, copy_in=True)
        N.check(self.lib.b200_comm_send(self._blob, dst, buf.data_ptr(), nbytes, self._box,
                                        self.box_bytes, self._p2p_flags, self._p2p_acks,
                                        self._sent.data_ptr(), N.stream()), "b200_comm_send")

    def recv(self, tensor: torch.Tensor, src: int) -> torch.Tensor:
        buf, nbytes = self._p2p_payload(tensor, copy_in=False)
        N.check(self.lib.b200_comm_recv(self._blob, src, buf.data_ptr(), nbytes, self._box,
                                        self.box_bytes, self._p2p_flags, self._p2p_acks,
                                        self._recvd.data_ptr(), N.stream()), "b200_comm_recv")
        if buf.data_ptr() != tensor.data_ptr():
            tensor.copy_(buf[:tensor.numel() * tensor.element_size()].view(tensor.dtype)
                         .view(tensor.shape))
        return tensor

    def sendrecv(self, send_tensor: torch.Tensor, dst: int, recv_tensor: torch.Tensor, src: int):
        """Send to dst and receive from src in ONE launch (ring shift / pairwise swap).
        Use this instead of a concurrent send + recv pair on two streams: two launches on
        one GPU are not guaranteed to overlap, and the mailbox acks would deadlock."""
        sbuf, sbytes = self._p2p_payload(send_tensor, copy_in=True)
        rbuf, rbytes = self._p2p_payload(recv_tensor, copy_in=False)
        N.check(self.lib.b200_comm_sendrecv(self._blob, dst, sbuf.data_ptr(), sbytes, src,
                                            rbuf.data_ptr(), rbytes, self._box, self.box_bytes,
                                            self._p2p_flags, self._p2p_acks, self._sent.data_ptr(),
                                            self._recvd.data_ptr(), N.stream()),
                "b200_comm_sendrecv")
        if rbuf.data_ptr() != recv_tensor.data_ptr():
            recv_tensor.copy_(rbuf[:recv_tensor.numel() * recv_tensor.element_size()]
                              .view(recv_tensor.dtype).view(recv_tensor.shape))
        return recv_tensor

    def _p2p_payload(self, tensor, copy_in):
        nbytes = tensor.numel() * tensor.element_size()
        if tensor.is_contiguous() and nbytes % 16 == 0 and tensor.data_ptr() % 16 == 0:
            return tensor, nbytes
        pad = (nbytes + 15) // 16 * 16
        tmp = torch.zeros(pad, dtype=torch.uint8, device=self.device)
        if copy_in:
            tmp[:nbytes].copy_(tensor.contiguous().view(-1).view(torch.uint8))
        return tmp, pad

    # -- NVLS (NVLink SHARP): multicast region ---------------------------------------
    def nvls_init(self, nbytes: Optional[int] = None) -> bool:
        """Create the multicast object and this rank's bound memory (once). Returns False if
        the driver / fabric does not offer multicast on every rank (callers fall back to the
        two-shot kernel). Collective: every rank must call it at the same point."""
        if self._nvls is not None:
            return bool(self._nvls)
        import socket
        lib = self.lib
        ok = c_i(0)
        rc = lib.b200_nvls_supported(self.device.index or 0, C.byref(ok))
        mine = (rc == 0 and ok.value == 1 and self.world_size > 1
                and os.environ.get("B200DP_NVLS", "1") != "0")
        if not all(self.all_gather_object(bool(mine))):
            self._nvls = False
            return False
        nbytes = nbytes or (int(os.environ.get("B200DP_NVLS_MB", "128")) << 20)
        rounded = c_sz(0)
        N.check(lib.b200_nvls_granularity(self.world_size, nbytes, C.byref(rounded)),
                "b200_nvls_granularity")
        mc = c_ull(0)
        path = f"/tmp/b200dp_nvls_{os.environ.get('MASTER_PORT', '0')}_{os.getuid()}_{id(self) & 0xffff}"
        path = self.all_gather_object(path)[0]
        status = "ok"
        try:
            if self.rank == 0:
                fd = c_i(-1)
                N.check(lib.b200_nvls_create(self.world_size, rounded.value, C.byref(mc),
                                             C.byref(fd)), "b200_nvls_create")
                if os.path.exists(path):
                    os.unlink(path)
                srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                srv.bind(path)
                srv.listen(self.world_size)
                self.host_barrier()
                for _ in range(self.world_size - 1):
                    conn, _a = srv.accept()
                    socket.send_fds(conn, [b"mc"], [fd.value])
                    conn.close()
                srv.close()
                os.unlink(path)
                os.close(fd.value)
            else:
                self.host_barrier()
                cl = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                cl.connect(path)
                _msg, fds, _f, _a = socket.recv_fds(cl, 16, 1)
                cl.close()
                N.check(lib.b200_nvls_import(fds[0], C.byref(mc)), "b200_nvls_import")
                os.close(fds[0])
            N.check(lib.b200_nvls_add_device(mc.value, self.device.index or 0), "b200_nvls_add_device")
        except Exception as e:   # keep the ranks in step, then give up together
            status = f"{type(e).__name__}: {e}"
        if any(x != "ok" for x in self.all_gather_object(status)):
            self._nvls = False
            return False
        uc, mcp = c_p(), c_p()
        try:
            N.check(lib.b200_nvls_bind_map(mc.value, self.device.index or 0, rounded.value,
                                           C.byref(uc), C.byref(mcp)), "b200_nvls_bind_map")
        except Exception as e:
            status = f"{type(e).__name__}: {e}"
        if any(x != "ok" for x in self.all_gather_object(status)):
            self._nvls = False
            return False
        self._nvls = {"uc": uc.value, "mc": mcp.value, "bytes": rounded.value, "top": 0,
                      "arena": torch.as_tensor(_RawCudaMemory(uc.value, rounded.value),
                                               device=self.device)}
        self.host_barrier()
        return True

    def alloc_nvls(self, numel: int, dtype=torch.float32, device=None) -> torch.Tensor:
        """A zeroed tensor inside the NVLS region (same call order on every rank)."""
        if not self.nvls_init():
            raise CommError("NVLS is not available on this machine")
        nv = self._nvls
        esz = torch.empty((), dtype=dtype).element_size()
        off = (nv["top"] + 255) // 256 * 256
        if off + numel * esz > nv["bytes"]:
            raise CommError("NVLS region exhausted: raise B200DP_NVLS_MB")
        nv["top"] = off + numel * esz
        t = nv["arena"][off:off + numel * esz].view(dtype)
        t.zero_()
        return t

    def nvls_offset_of(self, t: torch.Tensor) -> Optional[int]:
        if not self._nvls:
            return None
        off = t.data_ptr() - self._nvls["uc"]
        if 0 <= off and off + t.numel() * t.element_size() <= self._nvls["bytes"]:
            return off
        return None

    # -- generic all-reduce ----------------------------------------------------
    GMAX = 256

    def _ar_channel(self, algo: int, src_bytes_needed: int, channel: int = 0):
        """Flags, counters and staging of all-reduce channel `channel`. Collectives on one
        channel are serialised by their sequence counter; independent streams of collectives
        (e.g. one per gradient bucket) use different channels. The tagged-word (one-shot) and
        bulk (two-shot) protocols get SEPARATE staging: a stale bulk payload word must never
        be mistaken for a sequence tag."""
        if self._ar is None:
            self._ar = {}
        ch = self._ar.get(channel)
        if ch is None:
            ch = self._ar[channel] = {"flags": self._region(2 * self.GMAX * MAX_WORLD * 4),
                                      "seq": self._counter(), "arrive": self._counter(),
                                      "stages": {}}
            torch.cuda.synchronize(self.device)
            self.host_barrier()
        st = ch["stages"].get(algo)
        if st is None or st["src_bytes"] < src_bytes_needed:
            floor_mb = "1" if algo == 0 else os.environ.get("B200DP_AR_STAGE_MB", "2")
            src_bytes = max(src_bytes_needed, int(floor_mb) << 20)
            src_bytes = (src_bytes + 255) // 256 * 256
            st = ch["stages"][algo] = {
                "src_bytes": src_bytes,
                "stage": self._region(2 * self.world_size * src_bytes)}
            torch.cuda.synchronize(self.device)
            self.host_barrier()
        return dict(ch, **st)

    def rs_target(self, nbytes: int, channel: int):
        """Where a producer kernel may scatter a bucket of `nbytes` directly (fused GEMM
        epilogue): the two-shot staging of `channel`. Returns the arguments of
        b200_gemm_bf16_rs."""
        nvec = nbytes // 16
        sl = (nvec + self.world_size - 1) // self.world_size
        ch = self._ar_channel(1, sl * 16, channel)
        return {"comm": self._blob, "stage_off": ch["stage"], "src_bytes": ch["src_bytes"],
                "slice_vec": sl, "seq": ch["seq"]}

    def choose_algo(self, nbytes: int, nvls_ok: bool = False) -> int:
        """0 one-shot, 1 two-shot, 2 NVLS for a bucket of `nbytes` (see `choose_from_table`)."""
        return choose_from_table(self.world_size, nbytes, nvls_ok)

    def all_reduce(self, tensor: torch.Tensor, op: str = "mean", algo="auto", blocks: int = 0,
                   sgd=None, channel: int = 0, push_from: int = 0):
        """In-place all-reduce of a contiguous fp32/bf16 tensor whose byte size is a multiple
        of 16. algo: "auto" | "oneshot" | "twoshot". sgd=(master, momentum, hp) fuses the
        SGD-momentum update: `tensor` is then the gradient bucket and the destination is
        sgd's parameter tensor (see parallel/reducer.py)."""
        if self.world_size == 1 and sgd is None:
            return tensor
        if tensor.dtype not in _DTYPE_CODE or not tensor.is_contiguous():
            raise CommError("all_reduce handles contiguous fp32/bf16 tensors")
        nbytes = tensor.numel() * tensor.element_size()
        if nbytes % 16:
            raise CommError("all_reduce size must be a multiple of 16 bytes (pad the bucket)")
        if algo == "nvls" or algo == 2:
            return self._all_reduce_nvls(tensor, op, sgd, channel, blocks)
        a = {"auto": self.choose_algo(nbytes, self.nvls_offset_of(tensor) is not None),
             "oneshot": 0, "twoshot": 1}[algo] if isinstance(algo, str) else int(algo)
        if a == 2:
            return self._all_reduce_nvls(tensor, op, sgd, channel, blocks)
        W = self.world_size
        nvec = nbytes // 16
        need = 2 * nbytes if a == 0 else ((nvec + W - 1) // W) * 16   # one-shot stages tagged words
        ch = self._ar_channel(a, need, channel)
        scale = 1.0 / W if op == "mean" else 1.0
        if sgd is None:
            out_t = tensor
            master = mom = hp = None
            upd, state_base = 0, 0
        else:
            out_t, master, mom, hp, state_base = sgd
            upd = 1
        out_off = self.offset_of(out_t)
        copy_back = None
        if a == 1 and out_off is None:
            bounce = self._bounce_buffer(nbytes)
            out_off = self.offset_of(bounce)
            copy_back = bounce
        N.check(self.lib.b200_comm_allreduce(
            self._blob, _DTYPE_CODE[tensor.dtype], a, upd, tensor.data_ptr(), out_t.data_ptr(),
            out_off or 0, nbytes, ch["stage"], ch["src_bytes"], ch["flags"], self.GMAX, ch["seq"],
            ch["arrive"], scale, N.ptr(master), N.ptr(mom), N.ptr(hp), int(state_base), int(blocks),
            int(push_from), N.stream()), "b200_comm_allreduce")
        if copy_back is not None:
            out_t.copy_(copy_back[:nbytes].view(out_t.dtype).view(out_t.shape))
        return out_t

    def _all_reduce_nvls(self, tensor, op, sgd, channel, blocks):
        """In-switch reduction: `tensor` must live in the NVLS region (alloc_nvls)."""
        off = self.nvls_offset_of(tensor)
        if off is None:
            raise CommError("all_reduce(algo='nvls') needs a tensor from comm.alloc_nvls()")
        nbytes = tensor.numel() * tensor.element_size()
        ch = self._ar_channel(1, 16, channel)
        scale = 1.0 / self.world_size if op == "mean" else 1.0
        if sgd is None:
            upd, out_off, master, mom, hp, base = 0, 0, None, None, None, 0
            out_t = tensor
        else:
            out_t, master, mom, hp, base = sgd
            upd, out_off = 1, self.offset_of(out_t)
            if out_off is None:
                raise CommError("fused SGD needs the parameters in the symmetric arena")
        N.check(self.lib.b200_nvls_allreduce(
            self._blob, self._nvls["mc"], off, nbytes, _DTYPE_CODE[tensor.dtype], ch["flags"],
            self.GMAX, ch["seq"], ch["arrive"], scale, upd, out_off, N.ptr(master), N.ptr(mom),
            N.ptr(hp), int(base), int(blocks), N.stream()), "b200_nvls_allreduce")
        return out_t

    # -- fused Net kernel B ------------------------------------------------------
    def net_channel(self) -> dict:
        """Staging, flags and sequence counter of the Net gradient exchange (collective on
        first use: every rank must reach it). Shared by kernel B of the two-kernel step and by
        the persistent kernel, whose tagged words carry the same sequence numbers."""
        if self._net is None:
            self._net = {
                "stage": self._region(self.lib.b200_comm_net_stage_bytes()),
                "flags": self._region(self.lib.b200_comm_net_flag_bytes()),
                "seq": self._counter(), "arrive": self._counter(),
            }
            torch.cuda.synchronize(self.device)
            self.host_barrier()
        return self._net

    def reduce_allreduce_sgd(self, eng, rows: int, write_grad: bool = False, update: bool = True):
        """Partial-row sum + one-shot all-reduce + 1/world + SGD-momentum in ONE kernel
        (SURVEY K17+K15) for ops.fused.FusedNetEngine."""
        n = self.net_channel()
        N.check(self.lib.b200_comm_net_reduce_allreduce_sgd(
            self._blob, N.ptr(eng.partial), rows,
            N.ptr(eng.flat.param) if update else None,
            N.ptr(eng.flat.momentum) if update else None,
            N.ptr(eng.flat.grad) if write_grad else None, N.ptr(eng.hp),
            N.ptr(eng.loss_partial), N.ptr(eng.loss_ring), eng.ring, N.ptr(eng.step_ctr),
            n["stage"], n["flags"], n["seq"], n["arrive"],
            int(os.environ.get("B200DP_NET_PROTOCOL", "0")), N.stream()),
            "b200_comm_net_reduce_allreduce_sgd")

    def close(self):
        torch.cuda.synchronize(self.device)
        self.host_barrier()
        for r, b in enumerate(self.bases):
            if r != self.rank and b:
                self.lib.b200_symm_close(b)
        self.bases = []
        # The local arena is intentionally not freed: torch tensors may still alias it.
