"""Process-group API (reference: src/train_dist.py:144-146, src/run1.py:19-24).

`init_process_group(backend, rank, world_size)` keeps the reference's call shape: every
process is told its rank and the world size and finds the others through
MASTER_ADDR/MASTER_PORT. The default address is 127.0.0.1 (the reference hard-codes a GCP
VPC address; the multi-machine setup is collapsed onto one 8xB200 box).

backend "b200" (default on CUDA): peer-memory communicator (parallel/comm.py: PeerComm);
torch.distributed (gloo) only bootstraps it. backend "gloo": CPU tensors over gloo, the
reference's own backend, used on machines without a GPU.
"""
from __future__ import annotations

import datetime
import os
from typing import Optional

import torch
import torch.distributed as dist

from .comm import BaseComm, GlooComm, PeerComm

_COMM: Optional[BaseComm] = None
_OWNS_TORCH_PG = False


def init_process_group(backend: str = "auto", rank: Optional[int] = None,
                       world_size: Optional[int] = None, master_addr: Optional[str] = None,
                       master_port: Optional[int] = None, device=None,
                       timeout_s: float = 300.0, arena_bytes: Optional[int] = None) -> BaseComm:
    """Join the job and return this process's communicator (also kept as the default)."""
    global _COMM, _OWNS_TORCH_PG
    if _COMM is not None:
        raise RuntimeError("process group already initialised; call destroy_process_group()")
    env = os.environ
    rank = int(env.get("RANK", 0)) if rank is None else int(rank)
    world_size = int(env.get("WORLD_SIZE", 1)) if world_size is None else int(world_size)
    if master_addr is not None:
        env["MASTER_ADDR"] = str(master_addr)
    if master_port is not None:
        env["MASTER_PORT"] = str(master_port)
    env.setdefault("MASTER_ADDR", "127.0.0.1")
    env.setdefault("MASTER_PORT", "29500")
    if backend == "auto":
        backend = "b200" if torch.cuda.is_available() and device != "cpu" else "gloo"
    cpu_group = None
    if world_size > 1 or dist.is_initialized():
        if not dist.is_initialized():
            # gloo carries the bootstrap objects only (rendezvous, handle exchange). Scripts
            # that time an NCCL baseline create their own NCCL group (scripts/_nccl.py).
            tb = "gloo"
            if backend == "b200":
                local = int(env.get("LOCAL_RANK", rank))
                torch.cuda.set_device(local if device is None else torch.device(device))
            dist.init_process_group(backend=tb, rank=rank, world_size=world_size,
                                    timeout=datetime.timedelta(seconds=timeout_s))
            _OWNS_TORCH_PG = True
        cpu_group = dist.new_group(backend="gloo") if backend == "b200" else None
    if backend == "b200":
        if device is None:
            device = torch.device("cuda", int(env.get("LOCAL_RANK", rank)) % max(1, torch.cuda.device_count()))
        _COMM = PeerComm(rank, world_size, torch.device(device), cpu_group=cpu_group,
                         arena_bytes=arena_bytes)
    elif backend == "gloo":
        _COMM = GlooComm(rank, world_size, cpu_group=cpu_group)
    else:
        raise ValueError(f"unknown backend {backend!r} (use 'b200' or 'gloo')")
    return _COMM


def get_comm() -> BaseComm:
    if _COMM is None:
        raise RuntimeError("init_process_group() has not been called")
    return _COMM


def is_initialized() -> bool:
    return _COMM is not None


def get_rank() -> int:
    return _COMM.rank if _COMM is not None else 0


def get_world_size() -> int:
    return _COMM.world_size if _COMM is not None else 1


def destroy_process_group():
    """Clean teardown (the reference never calls one: src/train_dist.py:161-164)."""
    global _COMM, _OWNS_TORCH_PG
    if _COMM is not None:
        _COMM.close()
        _COMM = None
    if _OWNS_TORCH_PG and dist.is_initialized():
        dist.destroy_process_group()
    _OWNS_TORCH_PG = False


# torch.distributed-shaped free functions (src/run1.py:13,16 use dist.send / dist.recv)
def send(tensor, dst: int):
    return get_comm().send(tensor, dst)


def recv(tensor, src: int):
    return get_comm().recv(tensor, src)


def broadcast(tensor, src: int = 0):
    return get_comm().broadcast(tensor, src)


def all_reduce(tensor, op: str = "sum", algo="auto"):
    return get_comm().all_reduce(tensor, op=op, algo=algo)


def barrier():
    return get_comm().barrier()
