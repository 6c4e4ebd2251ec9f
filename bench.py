#!/usr/bin/env python
"""Headline benchmark: MNIST `Net` training throughput (images/s, whole job) at N GPUs.

Metric/config (BASELINE.json): the reference's train_dist.py DDP loop on the src/model.py
MNIST net, per-GPU batch 64, fp32, synthetic 1x28x28 / 10-class data, random-init weights,
lr 0.02, momentum 0.5 (src/train_dist.py:124-127). One step = one optimizer step on every
rank (forward, loss, backward, gradient all-reduce, SGD-momentum). Weak scaling.

    python bench.py --gpus N --steps K --warmup W            # this framework
    python bench.py --impl reference --gpus N ...            # unmodified reference, baseline/_ref

For N > 1 launch with `python -m torch.distributed.run --nnodes=1 --nproc-per-node N
--master-addr 127.0.0.1 --master-port P bench.py --gpus N ...` (or just run it: it
re-launches itself that way when RANK is not set).

Prints ONE JSON line on rank 0. `value` is device-timed (CUDA events, max over ranks) with a
device-resident dataset larger than L2; `e2e` goes through the public API
(`Trainer.train_steps_from_host`) and includes, per step, the host->device copy of the
batch from pinned memory and a device->host read of the loss.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

PUBLISHED_SEC_PER_EPOCH = {1: 17.5, 2: 11.3, 4: 7.6, 8: 5.0}   # BASELINE.md §1 (read off a chart)
N_TRAIN = 60000
LR, MOMENTUM = 0.02, 0.5


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=938)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="per-GPU batch")
    ap.add_argument("--input", default="f32", choices=["f32", "u8"],
                    help="dtype of the images copied host->device in the e2e run")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-persist", action="store_true",
                    help="two-kernel step in 16-step CUDA graphs instead of the persistent kernel")
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--ref-cudnn", default="both", choices=["both", "on", "off"])
    return ap.parse_args()


def relaunch_if_needed(a):
    if a.gpus > 1 and "RANK" not in os.environ:
        from csed_514_project_distributed_training_using_pytorch_b200.parallel.launch import free_port
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
               f"--nproc-per-node={a.gpus}", "--master-addr", "127.0.0.1",
               "--master-port", str(free_port()), str(Path(__file__).resolve())] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))


def synthetic_host_dataset(rank, dtype):
    import torch
    g = torch.Generator().manual_seed(1234 + rank)
    y = torch.randint(0, 10, (N_TRAIN,), generator=g)
    if dtype == "u8":
        x = torch.randint(0, 256, (N_TRAIN, 784), dtype=torch.uint8, generator=g)
    else:
        x = torch.randn(N_TRAIN, 784, generator=g)
    return x, y


def result_line(a, impl, value, ms_per_step, e2e, clocks, launches, extra_cfg):
    pub = PUBLISHED_SEC_PER_EPOCH.get(a.gpus)
    base = (N_TRAIN / pub) if pub else None
    out = {
        "impl": impl,
        "metric": "mnist_net_train_images_per_sec",
        "value": value, "unit": "images/s", "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": (value / base) if (base and value) else None,
        "dtype": "fp32", "data": "synthetic",
        "sec_per_epoch_equiv": (N_TRAIN / value) if value else None,
        "config": dict({"model": "src/model.py Net (2 conv + 2 fc, 21,840 params)",
                        "global_batch": a.batch * a.gpus, "per_gpu_batch": a.batch,
                        "seq_len": None, "parallelism": f"dp{a.gpus}", "lr": LR,
                        "momentum": MOMENTUM,
                        "l2": "inputs larger than L2: every step reads a fresh batch of a "
                              "188 MB device-resident fp32 dataset; no flush"}, **extra_cfg),
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
    }
    print(json.dumps(out), flush=True)


# =============================================================================== ours
def wake_gpu(dev, ms: float = 40.0):
    """Five 18 us warm-up steps are 0.1 ms of work: keep the GPU busy with matmuls for `ms`
    milliseconds first, so the timed steps do not start from an idle GPU and the clock sampler
    has samples under load (both arms do this; it is not a training step, and it leaves the
    L2 full of foreign data)."""
    import torch
    a = torch.randn(4096, 4096, device=dev, dtype=torch.bfloat16)
    t0 = time.perf_counter()
    while (time.perf_counter() - t0) * 1e3 < ms:
        for _ in range(20):
            torch.mm(a, a)
        torch.cuda.synchronize()


def run_ours(a):
    import torch
    from csed_514_project_distributed_training_using_pytorch_b200 import parallel as P
    from csed_514_project_distributed_training_using_pytorch_b200.ops.fused import X_F32_GATHER
    from csed_514_project_distributed_training_using_pytorch_b200.trainer import Trainer
    from csed_514_project_distributed_training_using_pytorch_b200.utils.clocks import ClockSampler
    from csed_514_project_distributed_training_using_pytorch_b200.utils.config import TrainConfig

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if world != a.gpus:
        raise SystemExit(f"--gpus {a.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    comm = P.init_process_group("b200", rank=rank, world_size=world, device=dev) if world > 1 else None

    cfg = TrainConfig(learning_rate=LR, momentum=MOMENTUM, batch_size_train=a.batch * world,
                      per_rank_batch=a.batch, world_size=world, rank=rank,
                      cuda_graph=not a.no_graph)
    tr = Trainer(cfg, comm=comm, distributed_style=True)
    if a.no_persist:
        tr.persistent = False
    persistent = tr.persistent
    eng = tr.engine
    B, K, W = a.batch, a.steps, a.warmup

    # ---- device-timed: dataset resident in HBM (188 MB fp32 > 126 MB L2), one graph replay per step
    hx, hy = synthetic_host_dataset(rank, "f32")
    dx, dy = hx.to(dev), hy.to(dev)
    shard = N_TRAIN // world
    perm = torch.randperm(N_TRAIN, generator=torch.Generator().manual_seed(42))[rank::world][:shard]
    perm = perm.to(torch.int32).to(dev)
    steps_per_epoch = shard // B
    U = 16   # steps per captured graph (programmatic dependent launch spans step boundaries)
    if persistent:
        one = multi = None
    elif a.no_graph:
        def one():
            eng.step(dx, dy, X_F32_GATHER, gidx=perm, batch=B, idx_stride=B)
        multi = None
    else:
        one = eng.capture(dx, dy, X_F32_GATHER, gidx=perm, batch=B, idx_stride=B).replay
        multi = eng.capture(dx, dy, X_F32_GATHER, gidx=perm, batch=B, idx_stride=B, unroll=U).replay

    pos = [steps_per_epoch]   # position inside the epoch; starts at "epoch over"

    def run_steps(n):
        done = 0
        while done < n:
            new_epoch = pos[0] >= steps_per_epoch
            if new_epoch:
                pos[0] = 0
                if not persistent:
                    eng.begin_epoch()
            m = min(steps_per_epoch - pos[0], n - done)
            pos[0] += m
            i = 0
            if persistent:   # every step of this stretch in ONE launch of the persistent kernel
                eng.run_steps(dx, dy, X_F32_GATHER, gidx=perm, batch=B, nsteps=m, idx_stride=B,
                              begin_epoch=new_epoch)
                i = m
            while i < m:
                if multi is not None and m - i >= U:
                    multi(); i += U
                else:
                    one(); i += 1
            done += m

    def sync_all():
        torch.cuda.synchronize()
        if comm is not None:
            comm.host_barrier()

    def aligned_start(ev):
        """Ranks leave a host barrier tens of microseconds apart, which is several percent of a
        0.5 ms timed region: line them up on the device with the peer-memory barrier kernel
        and record the start event behind it."""
        torch.cuda._sleep(400_000)     # ~0.2 ms spin: the host enqueues what follows behind it, so the
        if comm is not None:           # region does not count the host's launch path with the GPU idle
            comm.barrier()
        ev.record()

    sampler = ClockSampler(local, period_s=0.02).start() if rank == 0 else None
    wake_gpu(dev)
    run_steps(W)
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = eng.launches
    aligned_start(e0)
    run_steps(K)
    e1.record()
    sync_all()
    ms = e0.elapsed_time(e1)
    if os.environ.get("B200DP_BENCH_DEBUG"):      # debug aid: repeat the timed region, print each
        import sys
        extra = []
        for _ in range(6):
            aligned_start(e0); run_steps(K); e1.record(); sync_all()
            extra.append(round(e0.elapsed_time(e1) * 1e3, 1))
        print(f"[bench debug] first {ms * 1e3:.1f} us, repeats {extra}", file=sys.stderr)
    if persistent:
        eng.check_errors()
    if comm is not None:
        comm.check_errors()
        ms = comm.max_over_ranks(ms)
    value = K * B * world / (ms / 1e3)
    launches = eng.launches - l0 if (persistent or a.no_graph) else K * 2
    kernels_per_step = launches / K

    # ---- end to end through the public API: pinned host -> device per step, loss -> host per step
    e2e = None
    if not a.skip_e2e:
        if a.input == "u8":
            hx, _ = synthetic_host_dataset(rank, "u8")
        hxp, hyp = hx.pin_memory(), hy.pin_memory()
        n_avail = N_TRAIN // B
        loss_pinned = torch.empty(max(K, W, 3), dtype=torch.float32).pin_memory()

        def host_steps(n):
            done, outs = 0, []
            while done < n:
                m = min(n_avail, n - done)
                outs.append(tr.train_steps_from_host(hxp, hyp, m, B,
                                                     loss_out=loss_pinned[done:done + m]))
                done += m
            return outs
        wake_gpu(dev, 10.0)
        for _ in range(max(W, 3)):      # one call per warm-up step: the host path warms up as well
            host_steps(1)
        sync_all()
        if comm is not None:
            comm.barrier()
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record()
        losses = host_steps(K)
        t_enq = time.perf_counter()
        e1.record()
        torch.cuda.synchronize()          # this rank's last loss is in host memory
        wall_ms = (time.perf_counter() - t0) * 1e3   # before the host barrier: gloo is not part of a step
        if os.environ.get("B200DP_BENCH_DEBUG"):
            import sys
            print(f"[bench debug] e2e: enqueue returned after {(t_enq - t0) * 1e6:.1f} us, wall {wall_ms * 1e3:.1f} us, "
                  f"events {e0.elapsed_time(e1) * 1e3:.1f} us", file=sys.stderr)
            reps = []
            for _ in range(5):
                ta = time.perf_counter(); e0.record(); host_steps(K); tb = time.perf_counter()
                e1.record(); torch.cuda.synchronize()
                reps.append((round((tb - ta) * 1e6, 1), round((time.perf_counter() - ta) * 1e6, 1),
                             round(e0.elapsed_time(e1) * 1e3, 1)))
            print(f"[bench debug] e2e repeats (enqueue, wall, events): {reps}", file=sys.stderr)
        sync_all()
        ems = max(e0.elapsed_time(e1), wall_ms)
        if comm is not None:
            ems = comm.max_over_ranks(ems)
        last = float(losses[-1][-1])
        e2e = {"value": K * B * world / (ems / 1e3), "unit": "images/s", "ms_per_step": ems / K,
               "h2d_bytes_per_step": B * 784 * hxp.element_size() + B * 8,
               "d2h_bytes_per_step": 4, "input_dtype": a.input, "last_loss": last,
               "api": "Trainer.train_steps_from_host",
               "submission": ("every step: its own H2D copies (images, labels) on a copy stream, a "
                              "32-bit stream write that tells the running persistent kernel the "
                              "batch has landed, and a 4-byte store of the step's loss by the kernel "
                              "into the caller's pinned buffer; one kernel launch per call")
               if persistent else
               ("every step: its own H2D copies (images, labels) and its own D2H copy of the "
                "loss; the kernels of 16 consecutive steps form one graph")}
    clocks = sampler.stop() if sampler is not None else None
    if rank == 0:
        result_line(a, "ours", value, ms / K, e2e, clocks, launches,
                    {"engine": "persistent kernel (csrc/net_persist.cu)" if persistent
                     else "two kernels per step in CUDA graphs",
                     "cuda_graph": (not a.no_graph) and not persistent,
                     "steps_per_launch": K if persistent else (1 if a.no_graph else U),
                     "kernels_per_step": kernels_per_step,
                     "timed_region_start": "in-stream peer barrier on every rank, then the start event",
                     "before_warmup": "40 ms of bf16 matmuls (GPU not idle, L2 flushed)",
                     "start_event": "recorded behind a 0.2 ms spin kernel, so the K-step launch is "
                                    "already queued when the region opens (device time, not the "
                                    "host's launch path; the e2e number includes that path)",
                     "last_losses": eng.last_losses(2).tolist()})
    if comm is not None:
        P.destroy_process_group()


# ========================================================================== reference
class _RefLoader:
    """A loader object for the reference's `main()`: yields W untimed + K timed batches and
    brackets the timed ones with barrier + synchronize + CUDA events. `from_host` copies
    each batch from pinned host memory (the reference's DataLoader(pin_memory=True) role)."""

    def __init__(self, x, y, batch, warmup, steps, dev, from_host):
        import torch
        self.torch = torch
        self.x, self.y, self.B, self.W, self.K, self.dev = x, y, batch, warmup, steps, dev
        self.from_host = from_host
        self.dataset = range(N_TRAIN)              # len(train_loader.dataset), src/train_dist.py:89
        self.sampler = self                        # train_loader.sampler.set_epoch(i), :72
        self.elapsed_ms = None
        self.wall_ms = None

    def set_epoch(self, e):
        pass

    def __len__(self):
        return self.W + self.K

    def _barrier(self):
        t = self.torch
        t.cuda.synchronize()
        if t.distributed.is_initialized() and t.distributed.get_world_size() > 1:
            t.distributed.barrier()
            t.cuda.synchronize()

    def __iter__(self):
        t = self.torch
        n_avail = self.x.shape[0] // self.B
        e0, e1 = t.cuda.Event(enable_timing=True), t.cuda.Event(enable_timing=True)
        for i in range(self.W + self.K):
            if i == self.W:
                self._barrier()
                self._t0 = time.perf_counter()
                e0.record()
            j = (i % n_avail) * self.B
            xb, yb = self.x[j:j + self.B], self.y[j:j + self.B]
            if self.from_host:
                xb = xb.to(self.dev, non_blocking=True)
                yb = yb.to(self.dev, non_blocking=True)
            yield xb.view(-1, 1, 28, 28), yb
        e1.record()
        t.cuda.synchronize()
        self.wall_ms = (time.perf_counter() - self._t0) * 1e3   # same rule as our arm: no barrier inside
        self._barrier()
        self.elapsed_ms = e0.elapsed_time(e1)


class _RefTestLoader:
    def __init__(self, dev):
        import torch
        self.dataset = range(1000)
        self.batch = (torch.randn(1000, 1, 28, 28, device=dev), torch.randint(0, 10, (1000,), device=dev))

    def __iter__(self):
        yield self.batch

    def __len__(self):
        return 1


def run_reference(a):
    ref_dir = ROOT / "baseline" / "_ref"
    if not (ref_dir / "train_dist.py").exists() and Path("/root/reference/src").exists() \
            and int(os.environ.get("LOCAL_RANK", 0)) == 0:
        # the install is git-ignored: redo it where the checkout is fresh (scripts/install_reference.sh)
        subprocess.call(["sh", str(ROOT / "scripts" / "install_reference.sh")],
                        stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    if not (ref_dir / "train_dist.py").exists():
        print(json.dumps({"impl": "reference",
                          "unavailable": "baseline/_ref is missing (see DESIGN.md: install from a "
                                         "/tmp copy with packaging metadata)"}), flush=True)
        return
    os.environ.setdefault("TQDM_DISABLE", "1")
    import types
    import torch
    try:
        import matplotlib.pyplot  # noqa: F401
    except Exception:
        # train_dist.py imports matplotlib at module scope only for plot(); the image has no
        # matplotlib, so an empty stand-in module satisfies the import. The reference file
        # itself is untouched.
        m = types.ModuleType("matplotlib")
        mp = types.ModuleType("matplotlib.pyplot")
        m.pyplot = mp
        sys.modules.setdefault("matplotlib", m)
        sys.modules.setdefault("matplotlib.pyplot", mp)
    sys.path.insert(0, str(ref_dir))
    import train_dist as ref                       # the unmodified reference module

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29531")
    # the reference's own call (src/train_dist.py:146) with the backend a GPU box needs
    torch.distributed.init_process_group(backend="nccl", rank=rank, world_size=world)
    from csed_514_project_distributed_training_using_pytorch_b200.utils.clocks import ClockSampler

    hx, hy = synthetic_host_dataset(rank, "f32")
    dx, dy = hx.to(dev), hy.to(dev)
    hxp, hyp = hx.pin_memory(), hy.pin_memory()
    B, K, W = a.batch, a.steps, a.warmup

    def one_run(cudnn_on, from_host):
        torch.backends.cudnn.enabled = cudnn_on    # the script sets False (src/train_dist.py:136)
        torch.manual_seed(1)
        # module-level names main() reads (set by the reference's __main__, :124-153)
        ref.learning_rate, ref.momentum = LR, MOMENTUM
        ref.train_losses, ref.train_counter, ref.test_losses = [], [], []
        ref.t0 = time.time()
        loader = _RefLoader(hxp if from_host else dx, hyp if from_host else dy, B, W, K, dev, from_host)
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):
            ref.main(rank=rank, epochs=1, model=ref.Net().to(dev), train_loader=loader,
                     test_loader=_RefTestLoader(dev))
        ms = loader.elapsed_ms
        wall = loader.wall_ms
        t = torch.tensor([ms, wall], device=dev)
        if world > 1:
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        return float(t[0]), float(t[1]), ref.train_losses[-1]

    variants = {"both": [False, True], "on": [True], "off": [False]}[a.ref_cudnn]
    sampler = ClockSampler(local, period_s=0.02).start() if rank == 0 else None
    wake_gpu(dev)
    res = {}
    for cud in variants:
        ms_dev, _, _ = one_run(cud, from_host=False)
        ms_e2e, wall_e2e, last = one_run(cud, from_host=True)
        res["cudnn_on" if cud else "cudnn_off"] = {
            "device_ms_per_step": ms_dev / K, "e2e_ms_per_step": max(ms_e2e, wall_e2e) / K,
            "images_per_s": K * B * world / (ms_dev / 1e3),
            "e2e_images_per_s": K * B * world / (max(ms_e2e, wall_e2e) / 1e3), "last_loss": last}
    clocks = sampler.stop() if sampler is not None else None
    best = max(res, key=lambda k: res[k]["images_per_s"])
    best_e2e = max(res, key=lambda k: res[k]["e2e_images_per_s"])
    if rank == 0:
        e2e = {"value": res[best_e2e]["e2e_images_per_s"], "unit": "images/s",
               "ms_per_step": res[best_e2e]["e2e_ms_per_step"],
               "h2d_bytes_per_step": B * 784 * 4 + B * 8, "d2h_bytes_per_step": 4,
               "variant": best_e2e, "api": "train_dist.main(rank, epochs, model, train_loader, test_loader)"}
        result_line(a, "reference", res[best]["images_per_s"], res[best]["device_ms_per_step"], e2e,
                    clocks, None,
                    {"reference_variants": res, "reported_variant": best,
                     "note": "unmodified baseline/_ref/train_dist.py main(): torch DDP + NCCL, "
                             "loss.item() every step; 'value' uses device-resident batches, "
                             "'e2e' copies each batch from pinned host memory"})
    torch.distributed.destroy_process_group()


def main():
    a = parse()
    relaunch_if_needed(a)
    if a.impl == "reference":
        try:
            run_reference(a)
        except Exception as e:  # the arm must not take the round down
            if int(os.environ.get("RANK", 0)) == 0:
                print(json.dumps({"impl": "reference",
                                  "unavailable": f"{type(e).__name__}: {str(e)[:200]}"}), flush=True)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
