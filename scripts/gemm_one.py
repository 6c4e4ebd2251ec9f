#!/usr/bin/env python
"""One GEMM shape, one variant, a few calls (for ncu): gemm_one.py M N K [variant] [bn]."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from csed_514_project_distributed_training_using_pytorch_b200.ops import gemm
M, N, K = (int(v) for v in sys.argv[1:4])
if len(sys.argv) > 4:
    gemm.set_variant(sys.argv[4], int(sys.argv[5]) if len(sys.argv) > 5 else 0)
dev = torch.device("cuda", 0)
x = torch.randn(M, K, device=dev).to(torch.bfloat16)
W = torch.randn(N, K, device=dev).to(torch.bfloat16)
for _ in range(3):
    y = gemm.gemm_bf16(x, W, False, False)
ref = x @ W.t()
torch.cuda.synchronize()
print("max err", float((y.float() - ref.float()).abs().max()) / float(ref.float().abs().max()))
