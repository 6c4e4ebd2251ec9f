"""BASELINE.json config 4: the 784-4096-4096-10 bf16 MLP under data parallelism.
ours  = tcgen05 GEMMs + per-layer buckets + fused two-shot all-reduce/SGD over peer memory
torch = nn.Linear (cuBLAS) + DistributedDataParallel (NCCL) + optim.SGD, same shapes
Run under torchrun. Device-timed (CUDA events), max over ranks.
  python -m torch.distributed.run --nproc-per-node N scripts/bench_mlp.py [--batch 512] [--steps 50]
"""
import argparse, json, os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import torch.distributed as dist
import torch.nn as nn
from csed_514_project_distributed_training_using_pytorch_b200 import parallel as P
from csed_514_project_distributed_training_using_pytorch_b200.models.mlp import WideMLP, LocalSink, MasterSGD
from csed_514_project_distributed_training_using_pytorch_b200.parallel.data_parallel import DataParallel
from csed_514_project_distributed_training_using_pytorch_b200.utils.flat import FlatParams
from csed_514_project_distributed_training_using_pytorch_b200.utils.clocks import ClockSampler
from _nccl import nccl_group

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=512)
ap.add_argument("--steps", type=int, default=50)
ap.add_argument("--warmup", type=int, default=10)
ap.add_argument("--skip-torch", action="store_true")
ap.add_argument("--algo", default="auto", choices=["twoshot", "nvls", "oneshot", "auto"])
ap.add_argument("--pingpong", action="store_true",
                help="two parameter buffers: wgrad + bucket kernel before dgrad (two steps per graph)")
a = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
os.environ.setdefault("B200DP_ARENA_MB", "1024")
comm = P.init_process_group("b200", rank=rank, world_size=world, device=dev) if world > 1 else None
B = a.batch
g = torch.Generator().manual_seed(rank)
x = torch.randn(B, 784, generator=g).to(torch.bfloat16).to(dev)
y = torch.randint(0, 10, (B,), generator=g).to(dev)
out = {"world": world, "per_gpu_batch": B, "grad_bytes_bf16": 20037642 * 2, "algo": a.algo,
       "pingpong": bool(a.pingpong)}
sampler = ClockSampler(local, period_s=0.02).start() if rank == 0 else None


def timed(fn, steps, warm):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    if world > 1: comm.host_barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    if world > 1:
        ms = comm.max_over_ranks(ms)
    return ms

# ---- ours
torch.manual_seed(0)
model = WideMLP().to(dev)
if comm is not None:
    ddp = DataParallel(model, comm, bucket_of=[2, 2, 1, 1, 0, 0], fuse_sgd=(0.02, 0.5, 0.0), algo=a.algo,
                       pingpong=a.pingpong)
    def ours_step():
        ddp.loss(x, y).backward()
    def ours_sync():
        torch.cuda.current_stream().wait_stream(ddp.reducer.stream)
else:
    flat = FlatParams(model, with_momentum=False)
    model.sink = LocalSink(flat)
    opt = MasterSGD(flat, 0.02, 0.5)
    def ours_step():
        model.loss(x, y).backward(); opt.step()
    def ours_sync(): pass
out["ours_eager_ms"] = timed(ours_step, a.steps, a.warmup)
ours_sync()
# whole step in a CUDA graph
try:
    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3): ours_step()
        ours_sync()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    if comm is not None: comm.host_barrier()
    graph = torch.cuda.CUDAGraph()
    per_graph = 2 if a.pingpong else 1      # ping-pong: an even and an odd step per graph
    if a.pingpong and comm is not None and ddp._live == 1:
        ours_step(); ours_sync()            # start the capture on the even buffer
    with torch.cuda.graph(graph, stream=s):
        for _ in range(per_graph):
            ours_step(); ours_sync()
    out["ours_graph_ms"] = timed(graph.replay, a.steps, a.warmup) / per_graph
except Exception as e:
    out["ours_graph_error"] = f"{type(e).__name__}: {str(e)[:300]}"
if comm is not None: comm.check_errors()
with torch.no_grad():
    out["ours_loss"] = float(model.loss(x, y))

# ---- torch DDP + NCCL + cuBLAS
if not a.skip_torch:
    torch.manual_seed(0)
    tm = nn.Sequential(nn.Linear(784, 4096), nn.ReLU(), nn.Linear(4096, 4096), nn.ReLU(), nn.Linear(4096, 10)).to(dev).to(torch.bfloat16)
    if world > 1:
        tm = nn.parallel.DistributedDataParallel(tm, device_ids=[local], process_group=nccl_group())
    topt = torch.optim.SGD(tm.parameters(), lr=0.02, momentum=0.5)
    lossf = nn.CrossEntropyLoss()
    def torch_step():
        topt.zero_grad(set_to_none=True)
        lossf(tm(x).float(), y).backward()
        topt.step()
    out["torch_ddp_eager_ms"] = timed(torch_step, a.steps, a.warmup)
for k in list(out):
    if k.endswith("_ms"):
        out[k.replace("_ms", "_img_per_s")] = round(B * world / (out[k] * 1e-3))
if rank == 0:
    out["clocks"] = sampler.stop()
    print(json.dumps(out, indent=1))
    if os.environ.get("B200DP_OUT"):
        Path(os.environ["B200DP_OUT"]).write_text(json.dumps(out, indent=1))
if comm is not None:
    P.destroy_process_group()
