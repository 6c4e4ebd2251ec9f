"""The reference's run1.py / run2.py (src/run1.py, src/run2.py): rank 0 adds 1 to a zero
tensor and sends it to rank 1, which receives it; both print `Rank r has data 1.`.

    python -m ...cli.p2p_demo --rank 0      # "run1.py" (on machine / GPU 0)
    python -m ...cli.p2p_demo --rank 1      # "run2.py" (on machine / GPU 1)
    python -m ...cli.p2p_demo --spawn       # both ranks on this box

Each invocation spawns its rank in a child process with the "spawn" start method like the
reference (src/run1.py:27-37). On GPUs the transfer is the mailbox kernel over NVLink.
"""
from __future__ import annotations

import argparse
import os

import torch
import torch.multiprocessing as mp


def run(rank, size, comm):
    """src/run1.py:8-17"""
    from .. import parallel as dist
    tensor = torch.zeros(1, device=comm.device)
    if rank == 0:
        tensor += 1
        dist.send(tensor, dst=1)        # send the tensor to process 1
    else:
        dist.recv(tensor, src=0)        # receive the tensor from process 0
    if comm.device.type == "cuda":
        torch.cuda.synchronize()
        comm.check_errors()
    print("Rank ", rank, " has data ", tensor[0].cpu())


def init_process(rank, size, fn, backend="auto", master_addr="127.0.0.1", master_port=29500):
    """src/run1.py:19-24: set the rendezvous env, join the group, run fn."""
    from .. import parallel as dist
    os.environ["MASTER_ADDR"] = master_addr
    os.environ["MASTER_PORT"] = str(master_port)
    os.environ.setdefault("LOCAL_RANK", str(rank))
    comm = dist.init_process_group(backend, rank=rank, world_size=size)
    try:
        fn(rank, size, comm)
    finally:
        dist.destroy_process_group()


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--rank", type=int, default=0)
    ap.add_argument("--size", type=int, default=2)
    ap.add_argument("--backend", default="auto", choices=["auto", "b200", "gloo"])
    ap.add_argument("--master_addr", default="127.0.0.1")
    ap.add_argument("--master_port", type=int, default=29500)
    ap.add_argument("--spawn", action="store_true", help="start every rank from this process")
    ns = ap.parse_args(argv)
    mp.set_start_method("spawn", force=True)
    ranks = range(ns.size) if ns.spawn else [ns.rank]
    procs = []
    for r in ranks:
        p = mp.Process(target=init_process,
                       args=(r, ns.size, run, ns.backend, ns.master_addr, ns.master_port))
        p.start()
        procs.append(p)
    for p in procs:
        p.join()
    if any(p.exitcode != 0 for p in procs):
        raise SystemExit(1)


if __name__ == "__main__":
    main()
