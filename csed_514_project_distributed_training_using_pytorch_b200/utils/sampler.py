"""DistributedSampler-equivalent sharding (reference: src/train_dist.py:33-37,72;
semantics pinned from torch/utils/data/distributed.py:102-134).

num_samples = ceil(N / world); total = num_samples * world; the epoch's permutation is
randperm(N) from a generator seeded with seed + epoch; it is padded by wrapping (or cut when
drop_last) to `total`; rank r takes indices[r:total:world]. The output is bit-equal to
torch's sampler (tests/test_cpu_core.py::test_sampler_bit_parity_with_torch) and comes back as
one int64 CPU tensor; the trainer converts it to int32 and uploads it once per epoch for the
device-side gather.
"""
from __future__ import annotations

import math
from typing import Iterator

import torch


class ShardSampler:
    def __init__(self, dataset_len: int, num_replicas: int = 1, rank: int = 0,
                 shuffle: bool = True, seed: int = 0, drop_last: bool = False):
        if not (0 <= rank < num_replicas):
            raise ValueError(f"invalid rank {rank} for world size {num_replicas}")
        self.n = int(dataset_len)
        self.num_replicas, self.rank = num_replicas, rank
        self.shuffle, self.seed, self.drop_last = shuffle, seed, drop_last
        self.epoch = 0
        if drop_last and self.n % num_replicas != 0:
            self.num_samples = math.ceil((self.n - num_replicas) / num_replicas)
        else:
            self.num_samples = math.ceil(self.n / num_replicas)
        self.total_size = self.num_samples * num_replicas

    def set_epoch(self, epoch: int) -> None:
        """src/train_dist.py:72 — changes the permutation of the next epoch."""
        self.epoch = int(epoch)

    def indices(self) -> torch.Tensor:
        """This rank's sample indices for the current epoch (int64, CPU)."""
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.n, generator=g)
        else:
            idx = torch.arange(self.n)
        if not self.drop_last:
            pad = self.total_size - idx.numel()
            if pad > 0:
                if pad <= idx.numel():
                    idx = torch.cat([idx, idx[:pad]])
                else:
                    reps = math.ceil(pad / idx.numel())
                    idx = torch.cat([idx, idx.repeat(reps)[:pad]])
        else:
            idx = idx[:self.total_size]
        return idx[self.rank:self.total_size:self.num_replicas].contiguous()

    def __iter__(self) -> Iterator[int]:
        return iter(self.indices().tolist())

    def __len__(self) -> int:
        return self.num_samples

    def batches(self, batch_size: int):
        """(start, size) of every batch of this rank's epoch; the short tail is kept
        (DataLoader default drop_last=False, src/train_dist.py:40-45)."""
        out, i = [], 0
        while i < self.num_samples:
            out.append((i, min(batch_size, self.num_samples - i)))
            i += batch_size
        return out
