"""Device-time the fused Net step pieces with CUDA events (1 GPU)."""
import sys, json, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from csed_514_project_distributed_training_using_pytorch_b200.models import Net
from csed_514_project_distributed_training_using_pytorch_b200.ops.fused import FusedNetEngine, X_F32, X_F32_GATHER


def ev_time(fn, iters=200, warm=20):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3  # us


def main():
    dev = torch.device("cuda", 0)
    torch.manual_seed(1)
    out = {}
    for B in (64, 128, 512):
        net = Net().to(dev).train()
        eng = FusedNetEngine(net, lr=0.01, momentum=0.5)
        x = torch.randn(B, 1, 28, 28, device=dev)
        y = torch.randint(0, 10, (B,), device=dev)
        out[f"B{B}_kernelA_us"] = ev_time(lambda: eng.forward_backward(x, y))
        out[f"B{B}_kernelB_us"] = ev_time(lambda: eng.reduce_and_update(B))
        out[f"B{B}_eager_step_us"] = ev_time(lambda: eng.step(x, y))
        g = eng.capture(x, y)
        out[f"B{B}_graph_step_us"] = ev_time(lambda: g.replay(), iters=1000)
    # epoch over a device-resident dataset, one graph replay per step
    N_ = 60000
    data = torch.randn(N_, 784, device=dev)
    labels = torch.randint(0, 10, (N_,), device=dev)
    perm = torch.randperm(N_, device=dev).to(torch.int32)
    net = Net().to(dev).train()
    eng = FusedNetEngine(net)
    g = eng.capture(data, labels, X_F32_GATHER, gidx=perm, batch=64, idx_stride=64)
    eng.begin_epoch()
    steps = N_ // 64
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for rep in range(2):
        eng.begin_epoch()
        a.record()
        for _ in range(steps):
            g.replay()
        b.record()
        torch.cuda.synchronize()
        out[f"epoch_resident_ms_rep{rep}"] = a.elapsed_time(b)
    out["epoch_last_losses"] = eng.last_losses(3).tolist()
    for U in (4, 16):
        gU = eng.capture(data, labels, X_F32_GATHER, gidx=perm, batch=64, idx_stride=64, unroll=U)
        for rep in range(2):
            eng.begin_epoch()
            a.record()
            for _ in range(steps // U):
                gU.replay()
            b.record()
            torch.cuda.synchronize()
        out[f"epoch_resident_unroll{U}_ms"] = a.elapsed_time(b) * steps / (steps // U * U)
    # eval 10k
    xe = torch.randn(1000, 1, 28, 28, device=dev)
    ye = torch.randint(0, 10, (1000,), device=dev)
    out["eval_B1000_us"] = ev_time(lambda: eng.evaluate_batch(xe, ye), iters=50, warm=5)
    # layered autograd path
    from csed_514_project_distributed_training_using_pytorch_b200.ops import net_ops
    net2 = Net().to(dev).train()
    x = torch.randn(64, 1, 28, 28, device=dev)
    y = torch.randint(0, 10, (64,), device=dev)
    def layered():
        for p in net2.parameters():
            p.grad = None
        loss = net_ops.nll_loss(net2(x), y)
        loss.backward()
    out["layered_fwd_bwd_us"] = ev_time(layered, iters=100, warm=10)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
