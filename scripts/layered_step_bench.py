#!/usr/bin/env python
"""One training step (forward, loss, backward, SGD) at large per-GPU batch: the layered autograd
path with conv2 on CUDA cores ("simt"), on tcgen05 ("tc"), the measured mix ("auto"), and the
persistent fused kernel. CUDA events, median of 15, clock record attached."""
import json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from csed_514_project_distributed_training_using_pytorch_b200.models import Net
from csed_514_project_distributed_training_using_pytorch_b200.ops import net_ops
from csed_514_project_distributed_training_using_pytorch_b200.ops.persist import PersistentNetEngine, X_F32
from csed_514_project_distributed_training_using_pytorch_b200.utils.clocks import ClockSampler

dev = torch.device("cuda", 0)


def med(fn, n=15):
    for _ in range(4):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(n):
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return round(sorted(ts)[n // 2], 1)


rows = {}
sampler = ClockSampler(0, period_s=0.02).start()
for B in (512, 1000, 4096, 16384):
    x = torch.randn(B, 1, 28, 28, device=dev)
    y = torch.randint(0, 10, (B,), device=dev)
    row = {}
    for impl in ("simt", "tc", "auto"):
        torch.manual_seed(1)
        net = Net(conv2_impl=impl).to(dev).train()
        opt = torch.optim.SGD(net.parameters(), lr=0.02, momentum=0.5)

        def step():
            opt.zero_grad(set_to_none=True)
            net_ops.nll_loss(net(x), y).backward()
            opt.step()
        row[f"layered_{impl}_us"] = med(step)
    torch.manual_seed(1)
    eng = PersistentNetEngine(Net().to(dev).train(), lr=0.02, momentum=0.5)
    xf = x.view(B, 784)
    row["persistent_fused_us"] = med(lambda: eng.run_steps(xf, y, X_F32, batch=B, nsteps=1))
    rows[B] = row
    print(B, row, flush=True)
res = {"what": "one optimizer step, eager launches (no graph) for the layered paths", "rows": rows,
       "clocks": sampler.stop()}
print(json.dumps(res))
if len(sys.argv) > 1:
    Path(sys.argv[1]).write_text(json.dumps(res, indent=1))
