#!/usr/bin/env python
"""conv2 forward: TMA-fed implicit GEMM (csrc/conv2_tc.cu) vs the first tensor-core version
(in-smem im2col) vs the SIMT kernel; CUDA-event timed, clock record attached."""
import json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from csed_514_project_distributed_training_using_pytorch_b200.models import Net
from csed_514_project_distributed_training_using_pytorch_b200.ops import net_ops
from csed_514_project_distributed_training_using_pytorch_b200.utils.clocks import ClockSampler

dev = torch.device("cuda", 0)
torch.manual_seed(1)
net = Net().to(dev)
w, b = net.conv2.weight.detach(), net.conv2.bias.detach()
out = {}
sampler = ClockSampler(0, period_s=0.02).start()
for B in (64, 512, 1000, 4096):
    a1 = torch.relu(torch.randn(B, 10, 12, 12, device=dev))
    row = {}
    for impl in ("simt", "tcgen05", "tc"):
        f = lambda: net_ops.conv2_drop_pool_relu(a1, w, b, None, 0, impl=impl)
        for _ in range(5):
            f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ts = []
        for _ in range(20):
            e0.record(); f(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        ts.sort()
        row[impl + "_us"] = round(ts[len(ts) // 2], 2)
    out[B] = row
    print(B, row, flush=True)
res = {"what": "conv2 fwd + pool + relu, layered call incl. Python and output allocation "
               "(tc also includes the c4 layout kernel)", "rows": out, "clocks": sampler.stop()}
print(json.dumps(res))
if len(sys.argv) > 1:
    Path(sys.argv[1]).write_text(json.dumps(res, indent=1))
