"""A handful of fused Net steps at per-GPU batch 64 (target for ncu captures)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from csed_514_project_distributed_training_using_pytorch_b200.models import Net
from csed_514_project_distributed_training_using_pytorch_b200.ops.fused import FusedNetEngine

dev = torch.device("cuda", 0)
torch.manual_seed(1)
net = Net().to(dev).train()
eng = FusedNetEngine(net)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
x = torch.randn(B, 1, 28, 28, device=dev)
y = torch.randint(0, 10, (B,), device=dev)
for _ in range(6):
    eng.step(x, y)
torch.cuda.synchronize()
print("ok", eng.last_losses(2).tolist())
