"""Rank entry points for the multi-process tests (must be importable for mp spawn)."""
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

import torch

PKG = "csed_514_project_distributed_training_using_pytorch_b200"


def _pg():
    import importlib
    return importlib.import_module(PKG + ".parallel")


def p2p_demo(rank, world, backend, out_dir):
    """The reference's run1.py/run2.py: rank 0 sends tensor(1.) to rank 1 (src/run1.py:8-17)."""
    P = _pg()
    comm = P.init_process_group(backend=backend, rank=rank, world_size=world)
    t = torch.zeros(4, device=comm.device)
    if rank == 0:
        t += 1
        P.send(t, dst=1)
    elif rank == 1:
        P.recv(t, src=0)
    if comm.device.type == "cuda":
        torch.cuda.synchronize()
        comm.check_errors()
    torch.save(t.cpu(), os.path.join(out_dir, f"p2p_{rank}.pt"))
    P.destroy_process_group()


def collectives(rank, world, backend, out_dir):
    """barrier, broadcast, all_reduce (one-shot, two-shot; fp32, bf16), big send/recv."""
    P = _pg()
    comm = P.init_process_group(backend=backend, rank=rank, world_size=world)
    dev = comm.device
    res = {}
    comm.barrier()
    # broadcast of an arena tensor and of a plain tensor
    a = comm.alloc(1000 * 4)
    a.copy_(torch.arange(4000, dtype=torch.float32) * (rank + 1))
    comm.broadcast(a, root=0)
    b = torch.full((37,), float(rank), device=dev)
    comm.broadcast(b, root=world - 1)
    res["bc_arena"] = a.cpu()
    res["bc_plain"] = b.cpu()
    # all-reduce variants
    for name, dtype in (("f32", torch.float32), ("bf16", torch.bfloat16)):
        for algo in ("oneshot", "twoshot"):
            for n in (8, 4096 + 8, 300000):
                g = torch.Generator().manual_seed(100 + rank)
                x = torch.randn(n, generator=g).to(dtype)
                if backend == "gloo" and dtype == torch.bfloat16:
                    continue
                t = comm.alloc(n, dtype) if algo == "twoshot" else x.to(dev).clone()
                t.copy_(x)
                comm.all_reduce(t, op="mean", algo=algo)
                res[f"ar_{name}_{algo}_{n}"] = t.float().cpu().clone()
                # repeat on the same channel (slot reuse, sequence numbers)
                t.copy_(x)
                comm.all_reduce(t, op="sum", algo=algo)
                res[f"ar2_{name}_{algo}_{n}"] = t.float().cpu().clone()
    # multi-chunk send/recv ring: rank r -> r+1
    if world >= 2:
        n = (3 << 20) // 4 + 4  # > 1 MiB mailbox: exercises chunking + acks
        src = torch.arange(n, dtype=torch.float32, device=dev) + rank
        dst = torch.empty(n, dtype=torch.float32, device=dev)
        nxt, prv = (rank + 1) % world, (rank - 1) % world
        comm.sendrecv(src, nxt, dst, prv)
        res["ring"] = dst[:8].cpu()
        res["ring_ok"] = bool(torch.equal(dst.cpu(), torch.arange(n, dtype=torch.float32) + prv))
        # one-directional multi-chunk transfer 0 -> 1 (sender and receiver on different GPUs)
        big = torch.arange(n, dtype=torch.float32, device=dev) * 2.0
        if rank == 0:
            comm.send(big, 1)
        elif rank == 1:
            got = torch.empty_like(big)
            comm.recv(got, 0)
            res["oneway_ok"] = bool(torch.equal(got, big))
    if dev.type == "cuda":
        torch.cuda.synchronize()
        comm.check_errors()
    torch.save(res, os.path.join(out_dir, f"coll_{rank}.pt"))
    P.destroy_process_group()


def _small_mlp(seed):
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Linear(20, 64), torch.nn.ReLU(), torch.nn.Linear(64, 32),
                               torch.nn.ReLU(), torch.nn.Linear(32, 5))


def ddp_generic(rank, world, backend, out_dir, fuse, algo):
    """DataParallel over an arbitrary nn.Module through autograd hooks (T2, T3, T4)."""
    import importlib
    P = _pg()
    DP = importlib.import_module(PKG + ".parallel.data_parallel").DataParallel
    SGD = importlib.import_module(PKG + ".ops.sgd").FlatSGD
    comm = P.init_process_group(backend=backend, rank=rank, world_size=world)
    dev = comm.device
    model = _small_mlp(100 + rank).to(dev)       # ranks start DIFFERENT; DDP must sync them
    # three buckets (one per Linear) so overlap / ordering is exercised
    ddp = DP(model, comm, bucket_of=[2, 2, 1, 1, 0, 0],
             fuse_sgd=(0.05, 0.5, 0.0) if fuse else None, algo=algo)
    try:    # ping-pong buffers are refused for a module that reads its own parameters
        DP(model, comm, fuse_sgd=(0.05, 0.5, 0.0), pingpong=True)
        raise AssertionError("pingpong=True was accepted for a plain nn.Module")
    except ValueError:
        pass
    init = ddp.flat.param.clone().cpu()
    opt = None if fuse else SGD(ddp.flat.param, ddp.flat.grad, lr=0.05, momentum=0.5)
    losses = []
    for step in range(4):
        g = torch.Generator().manual_seed(1000 * step + rank)
        x = torch.randn(8, 20, generator=g).to(dev)
        y = torch.randint(0, 5, (8,), generator=g).to(dev)
        ddp.zero_grad()
        loss = torch.nn.functional.cross_entropy(ddp(x), y)
        loss.backward()
        if opt is not None:
            if dev.type == "cuda":
                torch.cuda.current_stream().wait_stream(ddp.reducer.stream)
            opt.step()
        losses.append(loss.item())
    # eval forward under no_grad does no sync/bookkeeping (T4)
    ddp.eval()
    with torch.no_grad():
        ddp(torch.zeros(2, 20, device=dev))
    assert not ddp._pass_open
    # no_sync: gradients stay local
    ddp.train()
    with ddp.no_sync():
        ddp.zero_grad()
        torch.nn.functional.cross_entropy(ddp(torch.full((2, 20), float(rank + 1), device=dev)),
                                          torch.zeros(2, dtype=torch.long, device=dev)).backward()
    local_grad = ddp.flat.grad.clone().cpu()
    if dev.type == "cuda":
        torch.cuda.current_stream().wait_stream(ddp.reducer.stream)
        torch.cuda.synchronize()
        comm.check_errors()
    torch.save({"init": init, "final": ddp.flat.param.clone().cpu(), "losses": losses,
                "local_grad": local_grad, "keys": list(ddp.state_dict().keys())},
               os.path.join(out_dir, f"ddp_{rank}.pt"))
    P.destroy_process_group()


def net_ddp_fused(rank, world, backend, out_dir, use_graph):
    """The reference's distributed loop body (src/train_dist.py:78-84) as two kernels."""
    import importlib
    P = _pg()
    DP = importlib.import_module(PKG + ".parallel.data_parallel").DataParallel
    Net = importlib.import_module(PKG + ".models").Net
    oracle = importlib.import_module(PKG + ".models.oracle")
    comm = P.init_process_group(backend=backend, rank=rank, world_size=world)
    dev = comm.device
    torch.manual_seed(1 + rank)
    ddp = DP(Net().to(dev), comm)
    init = ddp.flat.param.clone().cpu()
    eng = ddp.fused_engine(lr=0.02, momentum=0.5)
    B, steps = 16, 4
    sx = torch.empty(B, 1, 28, 28, device=dev)
    sy = torch.empty(B, dtype=torch.int64, device=dev)
    m2 = torch.empty(B, 20, device=dev)
    m3 = torch.empty(B, 50, device=dev)
    graph = None
    grads = []
    for step in range(steps):
        g = torch.Generator().manual_seed(1000 * step + rank)
        x = torch.randn(B, 1, 28, 28, generator=g)
        y = torch.randint(0, 10, (B,), generator=g)
        a, b = oracle.make_masks(B, g)
        sx.copy_(x); sy.copy_(y); m2.copy_(a); m3.copy_(b)
        if use_graph is True:
            if graph is None:
                s = torch.cuda.Stream()
                s.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(s):     # warm-up creates the comm regions
                    snap = eng._snapshot()
                    eng.step(sx, sy, masks=(m2, m3))
                    eng._restore(snap)
                torch.cuda.current_stream().wait_stream(s)
                torch.cuda.synchronize()
                comm.host_barrier()
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=s):
                    eng.step(sx, sy, masks=(m2, m3))
            graph.replay()
        elif use_graph == "persist":
            # the persistent kernel, one step per launch so the oracle's masks can be injected
            eng.run_steps(sx.view(B, 784), sy, 0, batch=B, nsteps=1, masks=(m2, m3), write_grad=True)
            grads.append(eng.flat.grad.clone().cpu())
        else:
            eng.step(sx, sy, masks=(m2, m3), write_grad=True)
            grads.append(eng.flat.grad.clone().cpu())
    torch.cuda.synchronize()
    comm.check_errors()
    if use_graph == "persist":
        eng.check_errors()
    torch.save({"init": init, "final": eng.flat.param.clone().cpu(), "grads": grads,
                "losses": eng.last_losses(steps), "steps": eng.steps_done()},
               os.path.join(out_dir, f"net_{rank}.pt"))
    P.destroy_process_group()


def net_ddp_persist_multi(rank, world, backend, out_dir):
    """24 optimizer steps in ONE persistent launch per rank (Philox dropout, device-resident
    shard), then the same 24 steps through the two-kernel step from the same start."""
    import importlib
    P = _pg()
    DP = importlib.import_module(PKG + ".parallel.data_parallel").DataParallel
    Net = importlib.import_module(PKG + ".models").Net
    comm = P.init_process_group(backend=backend, rank=rank, world_size=world)
    dev = comm.device
    B, steps = 64, 24
    g = torch.Generator().manual_seed(77 + rank)
    data = torch.randint(0, 256, (B * steps, 784), dtype=torch.uint8, generator=g).to(dev)
    labels = torch.randint(0, 10, (B * steps,), generator=g).to(dev)
    perm = torch.randperm(B * steps, generator=g).to(torch.int32).to(dev)
    out = {}
    for mode in ("persistent", "two-kernel"):
        torch.manual_seed(1 + rank)
        ddp = DP(Net().to(dev), comm)
        eng = ddp.fused_engine(lr=0.02, momentum=0.5, seed=3)
        eng.begin_epoch()
        if mode == "persistent":
            eng.run_steps(data, labels, 2, gidx=perm, batch=B, nsteps=10, idx_stride=B)
            eng.run_steps(data, labels, 2, gidx=perm, batch=B, nsteps=14, idx_stride=B)
        else:
            for _ in range(steps):
                eng.step(data, labels, 2, gidx=perm, batch=B, idx_stride=B)
        torch.cuda.synchronize()
        comm.check_errors()
        eng.check_errors()
        out[mode] = {"final": eng.flat.param.clone().cpu(), "mom": eng.flat.momentum.clone().cpu(),
                     "losses": eng.last_losses(steps), "steps": eng.steps_done(),
                     "launches": eng.launches}
        comm.host_barrier()
    torch.save(out, os.path.join(out_dir, f"netp_{rank}.pt"))
    P.destroy_process_group()


def late_rank(rank, world, backend, out_dir, what):
    """Failure detection: the last rank stays away from a collective for longer than
    the communicator's timeout. Every other rank's kernel must give up, and the host must get an
    exception that names the missing rank; tear-down must not hang."""
    import importlib
    import time
    os.environ["B200DP_COMM_TIMEOUT_S"] = "1.0"
    P = _pg()
    comm = P.init_process_group(backend=backend, rank=rank, world_size=world)
    CommError = importlib.import_module(PKG + ".parallel.comm").CommError
    NativeError = importlib.import_module(PKG + "._native").NativeError
    dev = comm.device
    late = world - 1
    err = None
    x = comm.alloc(4096) if what == "twoshot" else torch.zeros(4096, device=dev)
    x.fill_(float(rank + 1))
    eng = None
    if what == "persist":
        DP = importlib.import_module(PKG + ".parallel.data_parallel").DataParallel
        Net = importlib.import_module(PKG + ".models").Net
        torch.manual_seed(1)
        eng = DP(Net().to(dev), comm).fused_engine(lr=0.02, momentum=0.5)
        bx = torch.randn(4 * 16, 784, device=dev)
        by = torch.randint(0, 10, (4 * 16,), device=dev)
    # everyone healthy: no error (also creates the algorithm's staging, which is a collective
    # host-side step and must not be what the late rank holds up)
    comm.all_reduce(x.clone() if what != "twoshot" else x, op="sum",
                    algo="oneshot" if what == "persist" else what)
    x.fill_(float(rank + 1))
    torch.cuda.synchronize()
    comm.check_errors()
    comm.host_barrier()
    t0 = time.time()
    if rank == late:
        time.sleep(3.5)                                        # > timeout: the others give up
    try:
        if what == "persist":
            eng.run_steps(bx, by, 0, batch=16, nsteps=4)
            torch.cuda.synchronize()
            eng.check_errors()
        else:
            comm.all_reduce(x, op="sum", algo=what)
            torch.cuda.synchronize()
            comm.check_errors()
    except (CommError, NativeError) as e:
        err = str(e)
    waited = time.time() - t0
    torch.save({"err": err, "waited": waited}, os.path.join(out_dir, f"late_{rank}.pt"))
    P.destroy_process_group()


def protocol_stress(rank, world, backend, out_dir, iters):
    """Randomised arrival order on every collective (stands in for racecheck, which is closed on
    this pool): before each call every rank burns a random number of device cycles, so each
    protocol sees every interleaving of early and late peers; values are small integers, so
    every sum is exact and any torn, stale or missing word shows."""
    P = _pg()
    comm = P.init_process_group(backend=backend, rank=rank, world_size=world)
    dev = comm.device
    rng = torch.Generator().manual_seed(4242 + 17 * rank)
    sizes = [64, 4096, 1 << 16, 3 << 18]                     # floats: tagged, tagged, bulk, bulk
    bufs = {n: comm.alloc(n) for n in sizes}
    bad = []
    for it in range(iters):
        n = sizes[it % len(sizes)]
        algo = ("oneshot", "twoshot")[(it // len(sizes)) % 2]
        if algo == "oneshot" and n > (1 << 16):
            algo = "twoshot"
        delay = int(torch.randint(0, 400000, (1,), generator=rng))   # up to ~0.2 ms of spin
        torch.cuda._sleep(delay)
        x = bufs[n]
        x.copy_(torch.arange(n, device=dev, dtype=torch.float32).remainder_(97).add_(rank * 3 + it % 5))
        comm.all_reduce(x, op="sum", algo=algo)
        want = (torch.arange(n, device=dev, dtype=torch.float32).remainder_(97) * world
                + 3 * (world * (world - 1) // 2) + world * (it % 5))
        if not torch.equal(x, want):
            bad.append((it, algo, n, int((x != want).sum())))
        if it % 7 == 3:
            torch.cuda._sleep(int(torch.randint(0, 200000, (1,), generator=rng)))
            comm.barrier()
        if it % 11 == 5:
            root = it % world
            b = bufs[64]
            b.fill_(float(100 * rank + it))
            torch.cuda._sleep(int(torch.randint(0, 200000, (1,), generator=rng)))
            comm.broadcast(b, root=root)
            if not torch.equal(b, torch.full_like(b, float(100 * root + it))):
                bad.append((it, "broadcast", 64, -1))
    torch.cuda.synchronize()
    comm.check_errors()
    torch.save({"bad": bad, "iters": iters}, os.path.join(out_dir, f"stress_{rank}.pt"))
    P.destroy_process_group()


def mlp_ddp(rank, world, backend, out_dir, algo):
    """bf16 MLP on tcgen05 GEMMs under DataParallel: per-layer buckets, fused all-reduce +
    sharded master-weight SGD, overlap with backward."""
    import importlib
    P = _pg()
    DP = importlib.import_module(PKG + ".parallel.data_parallel").DataParallel
    WideMLP = importlib.import_module(PKG + ".models.mlp").WideMLP
    comm = P.init_process_group(backend=backend, rank=rank, world_size=world)
    dev = comm.device
    torch.manual_seed(50 + rank)
    model = WideMLP((784, 256, 128, 10)).to(dev)
    pingpong = algo.endswith("+pp")       # two parameter buffers: wgrad + bucket before dgrad
    algo = algo.replace("+pp", "")
    ddp = DP(model, comm, bucket_of=[2, 2, 1, 1, 0, 0], fuse_sgd=(0.1, 0.5, 0.0), algo=algo,
             pingpong=pingpong)
    init = ddp.flat.param.clone().float().cpu()
    losses = []
    nsteps = 5 if pingpong else 6         # odd: the live weights end up in the second buffer
    for step in range(nsteps):
        g = torch.Generator().manual_seed(1000 * step + rank)
        x = torch.randn(32, 784, generator=g).to(torch.bfloat16).to(dev)
        y = torch.randint(0, 10, (32,), generator=g).to(dev)
        loss = ddp.loss(x, y)
        loss.backward()
        losses.append(loss)
    torch.cuda.current_stream().wait_stream(ddp.reducer.stream)
    ddp.sync_module()                     # ping-pong: copy the live buffer into the module
    torch.cuda.synchronize()
    comm.check_errors()
    # fp32 master weights: whole buckets (one-shot) or this rank's shard of each (two-shot, NVLS)
    master = [(m.float().cpu(), int(base), int(lo), int(hi))
              for (m, _mom, base), (lo, hi) in zip(ddp.reducer.state, ddp.reducer.ranges)]
    torch.save({"init": init, "final": ddp.flat.param.clone().float().cpu(),
                "losses": [l.item() for l in losses], "master": master,
                "order": ddp.reducer.core.completion_order()},
               os.path.join(out_dir, f"mlp_{rank}.pt"))
    P.destroy_process_group()


def gemm_rs(rank, world, backend, out_dir, kind, bn):
    """Fused GEMM -> reduce-scatter: the wgrad tile vectors go straight into the owner ranks'
    two-shot staging over peer memory; the bucket kernel then skips its own scatter phase."""
    import importlib
    P = _pg()
    gemm = importlib.import_module(PKG + ".ops.gemm")
    comm = P.init_process_group(backend=backend, rank=rank, world_size=world)
    dev = comm.device
    M, N, K = 1024, 784, 520           # 784 = 3 x 256 + 16: a ragged last tile column
    g = torch.Generator().manual_seed(7 + rank)
    dy = torch.randn(K, M, generator=g).to(torch.bfloat16).to(dev)
    x = torch.randn(K, N, generator=g).to(torch.bfloat16).to(dev)
    bucket = torch.zeros(M * N, dtype=torch.bfloat16, device=dev)
    tgt = dict(comm.rs_target(bucket.numel() * 2, channel=3))
    tgt["base_vec"] = 0
    gemm.set_variant(kind, bn)
    for _ in range(3):                 # the staging slots alternate between rounds
        bucket.zero_()
        gemm.gemm_bf16(dy, x, True, True, out=bucket.view(M, N), rs=tgt)
        comm.all_reduce(bucket, op="mean", algo="twoshot", channel=3,
                        push_from=bucket.numel() * 2 // 16)
    gemm.set_variant("auto")
    torch.cuda.synchronize()
    comm.check_errors()
    torch.save({"got": bucket.float().cpu().view(M, N),
                "local": (dy.float().t() @ x.float()).cpu()},
               os.path.join(out_dir, f"rs_{rank}.pt"))
    P.destroy_process_group()


def trainer_dist(rank, world, backend, out_dir):
    """The reference's train_dist.py script body on `world` ranks (src/train_dist.py:118-164)."""
    import importlib
    P = _pg()
    Trainer = importlib.import_module(PKG + ".trainer").Trainer
    TrainConfig = importlib.import_module(PKG + ".utils.config").TrainConfig
    MetricsLogger = importlib.import_module(PKG + ".utils.metrics").MetricsLogger
    synthetic = importlib.import_module(PKG + ".utils.data").synthetic_mnist
    comm = P.init_process_group(backend=backend, rank=rank, world_size=world)
    cfg = TrainConfig(epochs=2, learning_rate=0.05, world_size=world, rank=rank, train_size=2048 + 24,
                      test_size=500, batch_size_test=250, results_dir=os.path.join(out_dir, "results"),
                      images_dir=os.path.join(out_dir, f"images{rank}"),
                      device="cpu" if backend == "gloo" else "cuda")
    tr = Trainer(cfg, comm=comm, logger=MetricsLogger(rank=rank, quiet=True), distributed_style=True)
    tr.set_data(synthetic(cfg.train_size, seed=0), synthetic(cfg.test_size, seed=1))
    log = tr.fit()
    if comm.device.type == "cuda":
        torch.cuda.synchronize()
        comm.check_errors()
    torch.save({"param": tr.flat.param.clone().cpu(), "test_losses": log.test_losses,
                "n_train_losses": len(log.train_losses)}, os.path.join(out_dir, f"tr_{rank}.pt"))
    P.destroy_process_group()


def nvls_allreduce(rank, world, backend, out_dir):
    """In-switch (multimem) all-reduce, fp32 and bf16, twice on the same buffers."""
    P = _pg()
    comm = P.init_process_group(backend=backend, rank=rank, world_size=world)
    dev = comm.device
    res = {"supported": bool(comm.nvls_init(32 << 20))}
    if res["supported"]:
        for name, dtype in (("f32", torch.float32), ("bf16", torch.bfloat16)):
            for n in (8, 4096 + 8, 1 << 20):
                t = comm.alloc_nvls(n, dtype)
                for k, op in enumerate(("mean", "sum")):
                    g = torch.Generator().manual_seed(100 + rank)
                    t.copy_(torch.randn(n, generator=g).to(dtype))
                    comm.all_reduce(t, op=op, algo="nvls")
                    res[f"{name}_{n}_{op}"] = t.float().cpu().clone()
        torch.cuda.synchronize()
        comm.check_errors()
    torch.save(res, os.path.join(out_dir, f"nvls_{rank}.pt"))
    P.destroy_process_group()
