#!/usr/bin/env python
"""Where does the first 20-step launch after a short warm-up lose its time? Phase stamps
(globaltimer) of every step's start for consecutive launches, plus the CUDA-event time."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from csed_514_project_distributed_training_using_pytorch_b200.models import Net
from csed_514_project_distributed_training_using_pytorch_b200.ops.persist import (
    PersistentNetEngine, X_F32_GATHER, persist_plan)

dev = torch.device("cuda", 0)
B, K = 64, 20
g = torch.Generator().manual_seed(0)
n = 60000
x = torch.randn(n, 784, generator=g).to(dev)
y = torch.randint(0, 10, (n,), generator=g).to(dev)
perm = torch.randperm(n, generator=g).to(torch.int32).to(dev)
torch.manual_seed(1)
eng = PersistentNetEngine(Net().to(dev).train(), lr=0.02, momentum=0.5)
dbg = eng.enable_phase_stamps()
W = persist_plan(B)["workers"]
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
eng.run_steps(x, y, X_F32_GATHER, gidx=perm, batch=B, nsteps=5, idx_stride=B, begin_epoch=True)
torch.cuda.synchronize()
for rep in range(5):
    e0.record()
    eng.run_steps(x, y, X_F32_GATHER, gidx=perm, batch=B, nsteps=K, idx_stride=B)
    e1.record()
    torch.cuda.synchronize()
    gt = dbg.cpu()[..., 1].double()
    t0 = gt[:W, 0, 0].min()
    starts = (gt[:W, :K, 0].median(dim=0).values - t0) / 1e3
    end = (gt[:, K - 1, :].max() - t0) / 1e3
    per = (starts[1:] - starts[:-1])
    print(f"launch {rep}: events {e0.elapsed_time(e1)*1e3:.1f} us; first stamp -> last stamp {float(end):.1f} us; "
          f"step periods {[round(float(v), 1) for v in per]}")
