"""Single-box launcher: spawn one rank per GPU (SURVEY §2.4 "Launcher").

The reference is launched by hand, one `python train_dist.py --local_rank R` per machine
(src/train_dist.py:120-122,141) and `run1.py` / `run2.py` per machine (src/run1.py:27-37).
Here `launch(fn, world_size)` spawns the ranks with the "spawn" start method
(src/run1.py:30), sets RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT, joins
them and propagates the first failure. `torchrun` works too: the CLIs read the same env.
"""
from __future__ import annotations

import os
import socket
import traceback
from typing import Callable, Sequence

import torch.multiprocessing as mp


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _entry(rank, fn, world_size, master_addr, master_port, args, err_q):
    os.environ.update({"RANK": str(rank), "LOCAL_RANK": str(rank), "WORLD_SIZE": str(world_size),
                       "MASTER_ADDR": master_addr, "MASTER_PORT": str(master_port)})
    try:
        fn(rank, world_size, *args)
    except BaseException:
        err_q.put((rank, traceback.format_exc()))
        raise


def launch(fn: Callable, world_size: int, args: Sequence = (), master_addr: str = "127.0.0.1",
           master_port: int = 0, timeout: float = None):
    """Run fn(rank, world_size, *args) in `world_size` spawned processes and join them."""
    port = master_port or free_port()
    ctx = mp.get_context("spawn")
    err_q = ctx.SimpleQueue()
    procs = []
    for r in range(world_size):
        p = ctx.Process(target=_entry, args=(r, fn, world_size, master_addr, port, tuple(args), err_q))
        p.start()
        procs.append(p)
    failed = None
    for p in procs:
        p.join(timeout)
        if p.is_alive():
            failed = failed or f"rank process {p.pid} timed out"
    if any(p.exitcode not in (0, None) for p in procs) or failed:
        for p in procs:           # exact PIDs we started, never a pattern
            if p.is_alive():
                p.terminate()
        msgs = []
        while not err_q.empty():
            r, tb = err_q.get()
            msgs.append(f"--- rank {r} ---\n{tb}")
        raise RuntimeError("launch failed: " + (failed or "") + "\n" + "\n".join(msgs) +
                           f"\nexit codes: {[p.exitcode for p in procs]}")
