#!/usr/bin/env python
"""Fixed cost of one persistent launch: CUDA-event time of run_steps(K) for several K, fitted
as fixed + K * per_step (debug aid for the 20-step bench)."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from csed_514_project_distributed_training_using_pytorch_b200.models import Net
from csed_514_project_distributed_training_using_pytorch_b200.ops.persist import PersistentNetEngine, X_F32_GATHER

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
res = {}
for K in (1, 2, 5, 10, 20, 40, 80):
    ts, hs = [], []
    for rep in range(7):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record()
        eng.run_steps(x, y, X_F32_GATHER, gidx=perm, batch=B, nsteps=K, idx_stride=B, begin_epoch=True)
        t1 = time.perf_counter()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
        hs.append((t1 - t0) * 1e6)
    ts.sort(); hs.sort()
    res[K] = ts[len(ts) // 2]
    print(f"K={K:3d}: {ts[len(ts)//2]:8.1f} us total (min {ts[0]:.1f}), host call {hs[len(hs)//2]:.1f} us")
slope = (res[80] - res[20]) / 60
print(f"per step {slope:.2f} us, fixed {res[20] - 20 * slope:.1f} us")
