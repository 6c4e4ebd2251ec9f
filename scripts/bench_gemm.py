"""TFLOP/s of the tcgen05 GEMM vs torch.matmul (cuBLAS) on the wide-MLP shapes (1 GPU)."""
import json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from csed_514_project_distributed_training_using_pytorch_b200.ops import gemm


def t(fn, iters=30, warm=5):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e-3


dev = torch.device("cuda", 0)
rows = []
for B in (64, 512, 4096):
    for (n, k) in ((4096, 784), (4096, 4096)):
        x = torch.randn(B, k, device=dev).to(torch.bfloat16)
        W = torch.randn(n, k, device=dev).to(torch.bfloat16)
        dy = torch.randn(B, n, device=dev).to(torch.bfloat16)
        fl = 2.0 * B * n * k
        r = {"B": B, "N": n, "K": k}
        r["fwd_ours_tflops"] = fl / t(lambda: gemm.linear_fwd(x, W)) / 1e12
        r["fwd_cublas_tflops"] = fl / t(lambda: x @ W.t()) / 1e12
        r["dgrad_ours_tflops"] = fl / t(lambda: gemm.linear_dgrad(dy, W)) / 1e12
        r["dgrad_cublas_tflops"] = fl / t(lambda: dy @ W) / 1e12
        r["wgrad_ours_tflops"] = fl / t(lambda: gemm.linear_wgrad(dy, x)) / 1e12
        r["wgrad_cublas_tflops"] = fl / t(lambda: dy.t() @ x) / 1e12
        rows.append({k_: (round(v, 1) if isinstance(v, float) else v) for k_, v in r.items()})
print(json.dumps(rows, indent=1))
