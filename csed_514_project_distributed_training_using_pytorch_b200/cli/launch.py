"""Spawn N ranks of `cli.train_dist` on this box (the reference starts them by hand, one per
machine: src/train_dist.py:120-122).

    python -m csed_514_project_distributed_training_using_pytorch_b200.cli.launch --nproc 8 -- \
        --epochs 6 --per_rank_batch 64
"""
from __future__ import annotations

import argparse
import sys

from ..parallel.launch import launch


def _rank_main(rank, world, argv):
    from . import train_dist
    train_dist.main(list(argv) + ["--rank", str(rank), "--world_size", str(world)])


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    rest = []
    if "--" in argv:
        i = argv.index("--")
        argv, rest = argv[:i], argv[i + 1:]
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--nproc", type=int, default=2)
    ap.add_argument("--master_port", type=int, default=0)
    ns = ap.parse_args(argv)
    launch(_rank_main, ns.nproc, args=(rest,), master_port=ns.master_port)


if __name__ == "__main__":
    main()
