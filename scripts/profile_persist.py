#!/usr/bin/env python
"""Short run of the persistent step kernel for `ncu` (one GPU, 6 steps per launch)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from csed_514_project_distributed_training_using_pytorch_b200.models import Net
from csed_514_project_distributed_training_using_pytorch_b200.ops.persist import PersistentNetEngine, X_F32_GATHER

dev = torch.device("cuda", 0)
B, S = 64, 6
g = torch.Generator().manual_seed(0)
n = B * S * 4
x = torch.randn(n, 784, generator=g).to(dev)
y = torch.randint(0, 10, (n,), generator=g).to(dev)
perm = torch.randperm(n, generator=g).to(torch.int32).to(dev)
torch.manual_seed(1)
eng = PersistentNetEngine(Net().to(dev).train(), lr=0.02, momentum=0.5)
for i in range(3):
    eng.run_steps(x, y, X_F32_GATHER, gidx=perm, batch=B, nsteps=S, idx_stride=B, begin_epoch=(i == 0))
torch.cuda.synchronize()
eng.check_errors()
print("ok", eng.steps_done(), float(eng.last_losses(1)[0]))
