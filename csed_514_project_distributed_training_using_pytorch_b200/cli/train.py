"""Single-process training: the reference's `python train.py` (src/train.py).

    python -m csed_514_project_distributed_training_using_pytorch_b200.cli.train [--epochs 3 ...]

Defaults are the reference's literals (n_epochs=3, batch 64/1000, lr 0.01, momentum 0.5,
log_interval 10, seed 1; src/train.py:12-19). Prints the same lines, writes
results/model.pth, results/optimizer.pth and images/train_test_curve.csv (+ .png when
matplotlib is available). `--device cpu` runs the PyTorch oracle path (BASELINE config 1).
"""
from __future__ import annotations

import argparse

from ..trainer import Trainer
from ..utils.config import TrainConfig, add_config_args, config_from_args
from ..utils.metrics import MetricsLogger


def main(argv=None):
    ap = add_config_args(argparse.ArgumentParser(description=__doc__), TrainConfig())
    ap.add_argument("--metrics_dir", default="", help="write metrics.jsonl here")
    ns = ap.parse_args(argv)
    cfg = config_from_args(ns)
    tr = Trainer(cfg, logger=MetricsLogger(ns.metrics_dir), distributed_style=False)
    tr.load_data()
    log = tr.fit()
    log.close()
    return log


if __name__ == "__main__":
    main()
