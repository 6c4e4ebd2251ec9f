#!/usr/bin/env python
"""Evaluation of 10,000 images in batches of 1000 (src/train.py:87-104): the fused eval kernel
against the layered forward with conv2 on CUDA cores and on tcgen05. CUDA events, median of 15."""
import json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from csed_514_project_distributed_training_using_pytorch_b200.models import Net
from csed_514_project_distributed_training_using_pytorch_b200.ops import net_ops
from csed_514_project_distributed_training_using_pytorch_b200.ops.fused import FusedNetEngine, X_F32
from csed_514_project_distributed_training_using_pytorch_b200.utils.clocks import ClockSampler

dev = torch.device("cuda", 0)
torch.manual_seed(1)
x = torch.randn(10000, 784, device=dev)
y = torch.randint(0, 10, (10000,), device=dev)


def med(fn, n=15):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(n):
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return round(sorted(ts)[n // 2], 1)


out = {}
sampler = ClockSampler(0, period_s=0.02).start()
eng = FusedNetEngine(Net().to(dev).eval(), lr=0.01, momentum=0.5)


def fused():
    for s in range(0, 10000, 1000):
        eng.evaluate_batch(x[s:s + 1000], y[s:s + 1000], X_F32, accumulate=s > 0, batch=1000)
out["fused_eval_kernel_us"] = med(fused)
for impl in ("simt", "tc"):
    net = Net(conv2_impl=impl).to(dev).eval()

    def layered():
        with torch.no_grad():
            for s in range(0, 10000, 1000):
                logp = net(x[s:s + 1000].view(1000, 1, 28, 28))
                net_ops.nll_loss(logp, y[s:s + 1000], reduction="sum")
    out[f"layered_{impl}_us"] = med(layered)
out["clocks"] = sampler.stop()
print(json.dumps(out))
if len(sys.argv) > 1:
    Path(sys.argv[1]).write_text(json.dumps(out, indent=1))
