"""Per-phase clock64 breakdown of the fused kernel A (debug aid)."""
import sys, ctypes
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from csed_514_project_distributed_training_using_pytorch_b200 import _native as N
from csed_514_project_distributed_training_using_pytorch_b200.models import Net
from csed_514_project_distributed_training_using_pytorch_b200.ops.fused import FusedNetEngine
N.register_sigs({"b200_net_set_phase_debug": [N.c_p]})
dev = torch.device("cuda", 0)
torch.manual_seed(1)
eng = FusedNetEngine(Net().to(dev).train())
B = 64
x = torch.randn(B, 1, 28, 28, device=dev); y = torch.randint(0, 10, (B,), device=dev)
for _ in range(5): eng.step(x, y)
buf = torch.zeros(148 * 16, dtype=torch.int64, device=dev)
N.call("b200_net_set_phase_debug", buf.data_ptr())
for _ in range(3): eng.step(x, y)
torch.cuda.synchronize()
N.call("b200_net_set_phase_debug", None)
t = buf.view(148, 16)[:2 * B].cpu()
# stamp k is taken after phase k; stamp 11 is unused (conv1_wgrad runs from stamp 10 to 12)
phases = [("start->loaded", 0, 1), ("conv1_fwd", 1, 2), ("conv2_fwd(+W3 wait)", 2, 3), ("fc1_fwd", 3, 4),
          ("fc2_fwd+softmax", 4, 5), ("fc2_bwd", 5, 6), ("fc1_bwd", 6, 7), ("build D", 7, 8),
          ("conv2 wgrad / dgrad T", 8, 9), ("gather", 9, 10), ("conv1_wgrad", 10, 12), ("tail", 12, 13)]
for role in (0, 1):
    rows = t[role::2]
    print(f"role {role}: total cycles median {float((rows[:,13]-rows[:,0]).double().median()):.0f}")
    for n, a, b in phases:
        if role == 0 and a >= 9: continue      # role 0 stops after the conv2 weight gradient
        col = (rows[:, b] - rows[:, a]).double()
        print(f"   {n:22s} median {float(col.median()):8.0f}  max {float(col.max()):8.0f}")

