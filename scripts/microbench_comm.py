"""Multi-GPU microbenchmarks (run under torchrun): fused Net kernel B latency, generic
all-reduce sweep vs NCCL. Device-timed with CUDA events, max over ranks."""
import json, os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import torch.distributed as dist
from csed_514_project_distributed_training_using_pytorch_b200 import parallel as P
from csed_514_project_distributed_training_using_pytorch_b200.models import Net
from csed_514_project_distributed_training_using_pytorch_b200.parallel.data_parallel import DataParallel
from csed_514_project_distributed_training_using_pytorch_b200.utils.clocks import ClockSampler
from _nccl import nccl_group


def timed(comm, fn, iters, warm=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize(); comm.host_barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    comm.check_errors()
    return comm.max_over_ranks(a.elapsed_time(b) / iters * 1e3)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    os.environ.setdefault("B200DP_ARENA_MB", "2048")   # the sweep stages up to 64 MiB per source
    comm = P.init_process_group("b200", rank=rank, world_size=world)
    dev = comm.device
    out = {"world": world}
    sampler = ClockSampler(local, period_s=0.02).start() if rank == 0 else None
    ng = nccl_group()
    what = sys.argv[1] if len(sys.argv) > 1 else "net"
    if what in ("net", "all"):
        torch.manual_seed(1)
        ddp = DataParallel(Net().to(dev), comm)
        eng = ddp.fused_engine(0.02, 0.5)
        B = 64
        x = torch.randn(B, 1, 28, 28, device=dev); y = torch.randint(0, 10, (B,), device=dev)
        eng.step(x, y)
        out["kernelA_eager_us"] = timed(comm, lambda: eng.forward_backward(x, y), 200)
        out["kernelB_eager_us"] = timed(comm, lambda: eng.reduce_and_update(B), 200)
        # graphs of B only / A only / A+B
        def graph_of(fn):
            s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                fn()
            torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize(); comm.host_barrier()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                fn()
            return g
        gB = graph_of(lambda: eng.reduce_and_update(B))
        out["kernelB_graph_us"] = timed(comm, gB.replay, 1000)
        gA = graph_of(lambda: eng.forward_backward(x, y))
        out["kernelA_graph_us"] = timed(comm, gA.replay, 1000)
        gAB = graph_of(lambda: eng.step(x, y))
        out["step_graph_us"] = timed(comm, gAB.replay, 1000)
        out["barrier_us"] = timed(comm, comm.barrier, 500)
    if what in ("sweep", "all"):
        sizes = [4 << 10, 16 << 10, 64 << 10, 256 << 10, 1 << 20, 4 << 20, 16 << 20, 64 << 20, 256 << 20]
        rows = []
        os.environ.setdefault("B200DP_AR_STAGE_MB", "36")
        for nbytes in sizes:
            n = nbytes // 4
            row = {"bytes": nbytes}
            iters = 200 if nbytes <= (1 << 20) else (50 if nbytes <= (16 << 20) else 10)
            t = comm.alloc(n, torch.float32) if nbytes <= (64 << 20) else None
            if t is not None:
                t.normal_()
                if nbytes <= (4 << 20):
                    row["oneshot_us"] = timed(comm, lambda: comm.all_reduce(t, "sum", "oneshot"), iters)
                row["twoshot_us"] = timed(comm, lambda: comm.all_reduce(t, "sum", "twoshot"), iters)
            if comm.nvls_init(300 << 20) and nbytes <= (256 << 20):
                if "nv" not in globals():
                    globals()["nv"] = comm.alloc_nvls((256 << 20) // 4, torch.float32)
                tn = globals()["nv"][:n]
                row["nvls_us"] = timed(comm, lambda: comm.all_reduce(tn, "sum", "nvls"), iters)
            x = torch.randn(n, device=dev)
            row["nccl_us"] = timed(comm, lambda: dist.all_reduce(x, group=ng), iters)
            for k in ("oneshot_us", "twoshot_us", "nvls_us", "nccl_us"):
                if k in row:   # bus bandwidth = 2(N-1)/N * S / t
                    row[k.replace("_us", "_busGBs")] = 2 * (world - 1) / world * nbytes / (row[k] * 1e-6) / 1e9
            rows.append(row)
        out["sweep"] = rows
    if what in ("mlp", "all"):
        from csed_514_project_distributed_training_using_pytorch_b200.models.mlp import WideMLP
        torch.manual_seed(0)
        model = WideMLP().to(dev)
        ddp = DataParallel(model, comm, bucket_of=[2, 2, 1, 1, 0, 0], fuse_sgd=(0.02, 0.5, 0.0),
                           algo=os.environ.get("MLP_ALGO", "twoshot"))
        B = 512
        x = torch.randn(B, 784, device=dev).to(torch.bfloat16); y = torch.randint(0, 10, (B,), device=dev)
        def full():
            ddp.loss(x, y).backward()
        def compute_only():
            with ddp.no_sync():
                ddp.loss(x, y).backward()
        def graph_of(fn, join):
            s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                for _ in range(3): fn()
                if join: torch.cuda.current_stream().wait_stream(ddp.reducer.stream)
            torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize(); comm.host_barrier()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                fn()
                if join: torch.cuda.current_stream().wait_stream(ddp.reducer.stream)
            return g
        out["mlp_step_graph_us"] = timed(comm, graph_of(full, True).replay, 50)
        out["mlp_compute_only_graph_us"] = timed(comm, graph_of(compute_only, False).replay, 50)
        red = ddp.reducer
        for b in range(red.core.nbuckets):
            lo, hi = red.ranges[b]
            def one(b=b):
                red._launch_bucket(b)
                torch.cuda.current_stream().wait_stream(red.stream)
            out[f"bucket{b}_{(hi-lo)*2}B_allreduce_sgd_us"] = timed(comm, one, 50)
            g = ddp.flat.grad[lo:hi]
            out[f"bucket{b}_plain_twoshot_us"] = timed(comm, lambda: comm.all_reduce(g, "mean", "twoshot"), 50)
            xx = torch.randn(hi - lo, device=dev).to(torch.bfloat16)
            out[f"bucket{b}_nccl_us"] = timed(comm, lambda: dist.all_reduce(xx, group=ng), 50)
    if rank == 0:
        out["clocks"] = sampler.stop()
        print(json.dumps(out, indent=1))
        if os.environ.get("B200DP_OUT"):
            Path(os.environ["B200DP_OUT"]).write_text(json.dumps(out, indent=1))
    P.destroy_process_group()


if __name__ == "__main__":
    main()
