#!/usr/bin/env python
"""conv2 backward: tcgen05 kernels (csrc/conv2_tc.cu) vs the SIMT kernel, per native call
(buffers preallocated, 20 back-to-back calls between two CUDA events), clock record attached."""
import json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from csed_514_project_distributed_training_using_pytorch_b200 import _native as N
from csed_514_project_distributed_training_using_pytorch_b200.models import Net
from csed_514_project_distributed_training_using_pytorch_b200.ops import net_ops
from csed_514_project_distributed_training_using_pytorch_b200.utils.clocks import ClockSampler

dev = torch.device("cuda", 0)
torch.manual_seed(1)
net = Net().to(dev)
w, b = net.conv2.weight.detach().contiguous(), net.conv2.bias.detach().contiguous()
lib = N.lib()
only = sys.argv[2] if len(sys.argv) > 2 else ""


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(5):
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3 / iters)
    return round(sorted(ts)[2], 2)


out = {}
sampler = ClockSampler(0, period_s=0.02).start()
import os
for B in ((int(os.environ.get('NCU_B', '1000')),) if only else (64, 256, 512, 1000, 4096, 16384)):
    a1 = torch.relu(torch.randn(B, 10, 12, 12, device=dev))
    da2 = torch.randn(B, 20, 4, 4, device=dev)
    a2 = torch.empty(B, 20, 4, 4, device=dev)
    pos2 = torch.empty(B, 20, 4, 4, dtype=torch.uint8, device=dev)
    s = N.stream()
    N.call("b200_conv2_fwd", N.ptr(a1), N.ptr(w), N.ptr(b), None, 0, 0.5, 0, 0, N.ptr(a2), N.ptr(pos2), B, s)
    ws = net_ops.workspace(dev)
    dw, db, da1 = torch.empty_like(w), torch.empty(20, device=dev), torch.empty_like(a1)
    c32 = torch.empty(lib.b200_conv2_c32_a1_floats(B), device=dev)
    dz32 = torch.empty(lib.b200_conv2_c32_dz_floats(B), device=dev)
    frame = torch.empty(lib.b200_conv2_frame_floats(B), device=dev)
    wd = torch.empty(lib.b200_conv2_dgrad_operand_floats(), device=dev)
    calls = {
        "simt_bwd": lambda: N.call("b200_conv2_bwd", N.ptr(a1), N.ptr(w), N.ptr(da2), N.ptr(a2), N.ptr(pos2), None,
                                   N.ptr(dw), N.ptr(db), N.ptr(da1), N.ptr(ws.scratch), N.ptr(ws.counter), B, s),
        "a1_to_c32": lambda: N.call("b200_a1_to_c32", N.ptr(a1), N.ptr(c32), B, s),
        "dz_c32": lambda: N.call("b200_conv2_dz_c32", N.ptr(da2), N.ptr(a2), N.ptr(pos2), None, N.ptr(dz32), B, s),
        "wgrad_tc": lambda: N.call("b200_conv2_wgrad_tc", N.ptr(c32), N.ptr(dz32), N.ptr(dw), N.ptr(db), B, s),
        "dz_frame": lambda: N.call("b200_conv2_dz_frame", N.ptr(da2), N.ptr(a2), N.ptr(pos2), None, N.ptr(frame), B, s),
        "dz_both": lambda: N.call("b200_conv2_dz_both", N.ptr(da2), N.ptr(a2), N.ptr(pos2), None, N.ptr(frame), N.ptr(dz32), B, s),
        "w_to_dgrad_operand": lambda: N.call("b200_conv2_w_to_dgrad_operand", N.ptr(w), N.ptr(wd), s),
        "dgrad_tc": lambda: N.call("b200_conv2_dgrad_tc", N.ptr(frame), N.ptr(wd), N.ptr(da1), B, s),
    }
    if only:                      # under ncu: run each kernel once after the operands exist
        for k in ("a1_to_c32", "dz_c32", "dz_frame", "w_to_dgrad_operand", "wgrad_tc", "dgrad_tc"):
            calls[k]()
        net_ops.conv2_drop_pool_relu(a1, w, b, None, 0, impl="tc")     # the forward kernel too
        torch.cuda.synchronize()
        continue
    row = {k + "_us": timed(f) for k, f in calls.items()}
    row["tc_bwd_total_us"] = round(sum(v for k, v in row.items()
                                      if not k.startswith(("simt", "dz_c32", "dz_frame"))), 2)   # dz_both replaces the two
    out[B] = row
    print(B, row, flush=True)
res = {"what": "conv2 backward (dw, db, da1) per native call; wgrad_tc includes its two memsets",
       "rows": out, "clocks": sampler.stop()}
print(json.dumps(res))
if len(sys.argv) > 1 and sys.argv[1] != "-":
    Path(sys.argv[1]).write_text(json.dumps(res, indent=1))
