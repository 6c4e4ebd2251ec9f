"""A few tcgen05 GEMM launches (B=4096 fc2 forward / wgrad) and the tcgen05 conv2 (ncu target)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from csed_514_project_distributed_training_using_pytorch_b200.ops import gemm, net_ops
dev = torch.device("cuda", 0)
x = torch.randn(4096, 4096, device=dev).to(torch.bfloat16)
W = torch.randn(4096, 4096, device=dev).to(torch.bfloat16)
for _ in range(3):
    y = gemm.linear_fwd(x, W)
    dW = gemm.linear_wgrad(x, x)
a1 = torch.relu(torch.randn(1000, 10, 12, 12, device=dev))
w = torch.randn(20, 10, 5, 5, device=dev) * 0.1
b = torch.zeros(20, device=dev)
for _ in range(3):
    net_ops.conv2_drop_pool_relu(a1, w, b, impl="tcgen05")
torch.cuda.synchronize()
print("ok", float(y.float().abs().mean()))
