"""One config object with the reference's literals as defaults (SURVEY §5 "Config / flag
system"): the reference hard-codes these at module level (src/train.py:12-19,
src/train_dist.py:124-145); here they are fields and CLI flags.
"""
from __future__ import annotations

import argparse
import os
from dataclasses import asdict, dataclass, fields


@dataclass
class TrainConfig:
    # src/train.py:12-19 (single) / src/train_dist.py:124-139 (distributed)
    epochs: int = 3                 # n_epochs=3 single, epochs=6 distributed
    batch_size_train: int = 64      # GLOBAL batch in the reference's distributed script
    batch_size_test: int = 1000
    learning_rate: float = 0.01     # 0.01 single, 0.02 distributed
    momentum: float = 0.5
    log_interval: int = 10
    random_seed: int = 1
    sampler_seed: int = 42          # DistributedSampler(seed=42), src/train_dist.py:37
    # distributed (src/train_dist.py:131-146)
    world_size: int = 1
    rank: int = 0
    local_rank: int = 0
    master_addr: str = "127.0.0.1"  # reference hard-codes 10.128.0.2 (a GCP VPC address)
    master_port: int = 29500
    per_rank_batch: int = 0         # 0 = reference rule int(batch_size_train / world_size)
    # data / IO
    data_dir: str = ""              # directory with MNIST IDX files; empty -> synthetic data
    train_size: int = 60000
    test_size: int = 10000
    results_dir: str = "results"    # model.pth / optimizer.pth (src/train.py:84-85)
    images_dir: str = "images"      # loss curves (src/train.py:117, src/train_dist.py:56)
    checkpoint_every_log: bool = True   # src/train.py:84-85 saves at every log interval
    resume: str = ""                # path of a checkpoint to resume from (not in the reference)
    # execution
    device: str = "cuda"            # "cpu" runs the PyTorch oracle path (BASELINE config 1)
    fused: bool = True              # fused 2-kernel step; False = layered autograd path
    cuda_graph: bool = True
    conv2_impl: str = "simt"        # layered path only: "simt" fp32 | "tc" TF32 on tcgen05 | "auto"
    dropout: bool = True
    input_dtype: str = "u8"         # how the dataset is stored: "u8" (MNIST native) or "f32"
    feed: str = "resident"          # "resident": dataset in HBM, the kernel gathers its batch;
                                    # "host": dataset in pinned host memory, one H2D copy per step

    def rank_batch(self) -> int:
        if self.per_rank_batch > 0:
            return self.per_rank_batch
        return max(1, int(self.batch_size_train / max(1, self.world_size)))

    def to_dict(self):
        return asdict(self)


def add_config_args(parser: argparse.ArgumentParser, defaults: TrainConfig = None):
    d = defaults or TrainConfig()
    for f in fields(TrainConfig):
        val = getattr(d, f.name)
        name = "--" + f.name
        if isinstance(val, bool):
            parser.add_argument(name, type=lambda s: s.lower() in ("1", "true", "yes", "on"),
                                default=val, metavar="BOOL")
        else:
            parser.add_argument(name, type=type(val), default=val)
    return parser


def config_from_args(ns: argparse.Namespace, base: TrainConfig = None) -> TrainConfig:
    cfg = base or TrainConfig()
    for f in fields(TrainConfig):
        if hasattr(ns, f.name) and getattr(ns, f.name) is not None:
            setattr(cfg, f.name, getattr(ns, f.name))
    return cfg


def apply_env(cfg: TrainConfig) -> TrainConfig:
    """Fill rank/world/master from the torchrun-style environment when present."""
    env = os.environ
    if "RANK" in env:
        cfg.rank = int(env["RANK"])
    if "WORLD_SIZE" in env:
        cfg.world_size = int(env["WORLD_SIZE"])
    if "LOCAL_RANK" in env:
        cfg.local_rank = int(env["LOCAL_RANK"])
    if "MASTER_ADDR" in env:
        cfg.master_addr = env["MASTER_ADDR"]
    if "MASTER_PORT" in env:
        cfg.master_port = int(env["MASTER_PORT"])
    return cfg
