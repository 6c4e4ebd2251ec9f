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
names = ["start->loaded", "conv1_fwd", "conv2_fwd(+W3 wait)", "fc1_fwd", "fc2_fwd+softmax", "fc2_bwd", "fc1_bwd", "build D", "conv2 wgrad / dgrad T", "gather", "-", "conv1_wgrad", "tail"]
for role in (0, 1):
    rows = t[role::2]
    d = rows[:, 1:14] - rows[:, 0:13]
    print(f"role {role}: total cycles median {float((rows[:,13]-rows[:,0]).double().median()):.0f}")
    for i, n in enumerate(names):
        col = d[:, i].double()
        if (role == 0 and i >= 9) or n == '-': continue
        print(f"   {n:22s} median {float(col.median()):8.0f}  max {float(col.max()):8.0f}")

