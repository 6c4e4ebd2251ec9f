"""TFLOP/s of every tcgen05 GEMM variant (one-tile CTA, CTA pair, persistent pair) against
cuBLAS on the wide-MLP shapes, plus a full-matrix error check against torch.matmul."""
import json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import os
from csed_514_project_distributed_training_using_pytorch_b200.ops import gemm
from csed_514_project_distributed_training_using_pytorch_b200.utils.clocks import ClockSampler


def t(fn, iters=20, warm=4):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e-3


dev = torch.device("cuda", 0)
variants = [("auto", 0), ("cta", 128), ("cta", 256), ("pair", 256), ("pair", 128), ("splitk", 256),
            ("persistent", 256), ("persistent", 128)]
shapes = [(4096, 4096, 4096), (8192, 8192, 4096), (512, 4096, 4096), (4096, 4096, 512),
          (4096, 784, 512), (512, 4096, 784)]
if len(sys.argv) > 1:
    variants = [v for v in variants if v[0] in sys.argv[1].split(",")]
rows = []
sampler = ClockSampler(0, period_s=0.02).start()
for (M, N, K) in shapes:
    x = torch.randn(M, K, device=dev).to(torch.bfloat16)
    W = torch.randn(N, K, device=dev).to(torch.bfloat16)
    xt = x.t().contiguous()      # [K, M]: MN-major A
    Wt = W.t().contiguous()      # [K, N]: MN-major B
    fl = 2.0 * M * N * K
    ref = x @ W.t()
    r = {"shape": f"{M}x{N}x{K}", "cublas": round(fl / t(lambda: x @ W.t()) / 1e12, 1)}
    for kind, bn in variants:
        gemm.set_variant(kind, bn)
        tag = f"{kind}{bn or ''}"
        try:
            r[tag + "_kk"] = round(fl / t(lambda: gemm.gemm_bf16(x, W, False, False)) / 1e12, 1)
            r[tag + "_mm"] = round(fl / t(lambda: gemm.gemm_bf16(xt, Wt, True, True)) / 1e12, 1)
            got = gemm.gemm_bf16(x, W, False, False)
            got2 = gemm.gemm_bf16(xt, Wt, True, True)
            den = float(ref.float().abs().max())
            r[tag + "_err"] = round(max(float((got.float() - ref.float()).abs().max()),
                                        float((got2.float() - ref.float()).abs().max())) / den, 4)
        except Exception as e:  # keep going: one broken variant must not hide the others
            r[tag + "_error"] = str(e)[:200]
            break
    gemm.set_variant("auto")
    rows.append(r)
    print(json.dumps(r), flush=True)
clocks = sampler.stop()
print(json.dumps({"clocks": clocks}), flush=True)
if os.environ.get("B200DP_OUT"):
    Path(os.environ["B200DP_OUT"]).write_text(json.dumps({"rows": rows, "clocks": clocks,
        "unit": "TFLOP/s bf16, CUDA events over 20 back-to-back calls after 4 warm-up calls",
        "columns": "_kk = both operands K-major, _mm = both MN-major, _err = max abs error / max |ref|"}, indent=1))
