#!/usr/bin/env python
"""Fixed cost of one persistent launch under the bench's conditions: CUDA-event time of
run_steps(K) for several K with (a) the same rows every run (L2-hot), (b) fresh rows of a 188 MB
dataset, (c) fresh rows while the NVML clock sampler runs."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from csed_514_project_distributed_training_using_pytorch_b200.models import Net
from csed_514_project_distributed_training_using_pytorch_b200.ops.persist import PersistentNetEngine, X_F32_GATHER
from csed_514_project_distributed_training_using_pytorch_b200.utils.clocks import ClockSampler

dev = torch.device("cuda", 0)
B = 64
g = torch.Generator().manual_seed(0)
n = 60000
x = torch.randn(n, 784, generator=g).to(dev)
y = torch.randint(0, 10, (n,), generator=g).to(dev)
perm = torch.randperm(n, generator=g).to(torch.int32).to(dev)
torch.manual_seed(1)
eng = PersistentNetEngine(Net().to(dev).train(), lr=0.02, momentum=0.5)
eng.run_steps(x, y, X_F32_GATHER, gidx=perm, batch=B, nsteps=8, idx_stride=B, begin_epoch=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def run(K, fresh, reps=7):
    ts = []
    for rep in range(reps):
        torch.cuda.synchronize()
        e0.record()
        eng.run_steps(x, y, X_F32_GATHER, gidx=perm, batch=B, nsteps=K, idx_stride=B,
                      begin_epoch=(not fresh) or rep == 0)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return ts


for label, fresh, sampler in (("hot rows", False, False), ("fresh rows", True, False),
                              ("fresh rows + NVML sampler", True, True)):
    s = ClockSampler(0, period_s=0.02).start() if sampler else None
    for K in (20, 80):
        ts = run(K, fresh)
        print(f"{label:28s} K={K:3d}: first {ts[0]:7.1f} us, median {sorted(ts)[len(ts)//2]:7.1f} us "
              f"({sorted(ts)[len(ts)//2]/K:.2f} us/step), all {[round(t) for t in ts]}")
    if s:
        s.stop()
