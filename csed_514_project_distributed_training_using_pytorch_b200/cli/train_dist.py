"""One rank of distributed training: the reference's `python train_dist.py --local_rank R`
(src/train_dist.py:118-164).

    python -m csed_514_project_distributed_training_using_pytorch_b200.cli.train_dist \
        --rank R --world_size W [--master_addr 127.0.0.1 --master_port 29500]

`--local_rank` is accepted as an alias of `--rank`, which is how the reference uses it
(src/train_dist.py:121,141). Under torchrun the rank/world come from the environment.
Defaults are the reference's distributed literals: lr 0.02, momentum 0.5, epochs 6, global
batch 64 split as int(64 / world_size) per rank (src/train_dist.py:124-142); pass
`--per_rank_batch 64` for the weak-scaling configuration of BASELINE.json.
"""
from __future__ import annotations

import argparse

from .. import parallel
from ..trainer import Trainer
from ..utils.config import TrainConfig, add_config_args, apply_env, config_from_args
from ..utils.metrics import MetricsLogger


def main(argv=None):
    defaults = TrainConfig(epochs=6, learning_rate=0.02, world_size=2)
    ap = add_config_args(argparse.ArgumentParser(description=__doc__), defaults)
    ap.add_argument("--metrics_dir", default="")
    ap.add_argument("--backend", default="auto", choices=["auto", "b200", "gloo"])
    ns = ap.parse_args(argv)
    cfg = config_from_args(ns, defaults)
    if ns.local_rank and not ns.rank:
        cfg.rank = ns.local_rank          # the reference's (mis)use of --local_rank as the rank
    cfg = apply_env(cfg)
    backend = ns.backend
    if cfg.device == "cpu":
        backend = "gloo"
    comm = parallel.init_process_group(backend, rank=cfg.rank, world_size=cfg.world_size,
                                       master_addr=cfg.master_addr, master_port=cfg.master_port)
    try:
        tr = Trainer(cfg, comm=comm, logger=MetricsLogger(ns.metrics_dir, rank=cfg.rank),
                     distributed_style=True)
        tr.load_data()
        log = tr.fit()
        log.close()
    finally:
        parallel.destroy_process_group()


if __name__ == "__main__":
    main()
