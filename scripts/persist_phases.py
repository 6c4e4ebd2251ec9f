#!/usr/bin/env python
"""Timeline of the persistent step kernel from its phase stamps (debug aid).

    python scripts/persist_phases.py [--batch 64] [--steps 24] [--out profiles/persist_phases.md]

Every CTA stamps clock64 and %globaltimer after each phase of each step
(`PersistentNetEngine.enable_phase_stamps`). The report gives, for the steady-state steps,
the median duration of every worker phase (SM cycles -> us at the measured SM clock), the
service CTAs' wait / compute / exchange / publish times, and the step period seen by the
workers, so the critical path can be read off directly.
"""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

from csed_514_project_distributed_training_using_pytorch_b200.models import Net  # noqa: E402
from csed_514_project_distributed_training_using_pytorch_b200.ops.persist import (  # noqa: E402
    PersistentNetEngine, X_F32_GATHER, persist_plan)

WORKER = ["wait conv1 weights + image", "conv1 fwd", "conv2 fwd (+ wait conv2 weights)",
          "fc1 fwd (+ wait fc weights)", "fc2 fwd + loss",
          "fc2 bwd (dpre) + publish acts + fc1 bwd (data)", "build dense dconv2",
          "conv2 wgrad (half)", "conv2 dgrad T (half)", "conv2 dgrad gather (half)",
          "conv1 wgrad (half) + tagged stores",
          "bubble: conv2 slice sum + exchange + SGD + publish"]
SERVICE = ["wait for producers (conv1: poll + sum rows)", "reduce / batch GEMM",
           "exchange + SGD", "publish"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=24)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    import os
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    comm = None
    if world > 1:   # run under torchrun: every rank runs the launch, rank 0 reports its own CTAs
        from csed_514_project_distributed_training_using_pytorch_b200 import parallel as P
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
        comm = P.init_process_group("b200", rank=rank, world_size=world)
    dev = torch.device("cuda", torch.cuda.current_device())
    B, S = a.batch, a.steps
    plan = persist_plan(B)
    g = torch.Generator().manual_seed(0)
    n = B * S * 2
    x = torch.randn(n, 784, generator=g).to(dev)
    y = torch.randint(0, 10, (n,), generator=g).to(dev)
    perm = torch.randperm(n, generator=g).to(torch.int32).to(dev)
    torch.manual_seed(1)
    if comm is not None:
        from csed_514_project_distributed_training_using_pytorch_b200.parallel.data_parallel import DataParallel
        eng = DataParallel(Net().to(dev).train(), comm).fused_engine(lr=0.02, momentum=0.5)
    else:
        eng = PersistentNetEngine(Net().to(dev).train(), lr=0.02, momentum=0.5)
    eng.begin_epoch()
    eng.run_steps(x, y, X_F32_GATHER, gidx=perm, batch=B, nsteps=S, idx_stride=B)   # warm-up
    torch.cuda.synchronize()
    dbg = eng.enable_phase_stamps()
    eng.run_steps(x, y, X_F32_GATHER, gidx=perm, batch=B, nsteps=S, idx_stride=B)
    torch.cuda.synchronize()
    eng.check_errors()
    if comm is not None:
        comm.host_barrier()
    d = dbg.cpu()
    clk, gt = d[..., 0].double(), d[..., 1].double()
    W = plan["workers"]
    lo, hi = 4, min(S, d.shape[1]) - 1          # steady-state steps
    lines = [f"# Persistent step kernel: phase timeline, batch {B}, {S} steps per launch, "
             f"{world} GPU(s) (rank {rank}'s CTAs)", "",
             f"Plan: {plan}", ""]
    # SM clock from the two time bases (cycles per ns over the whole launch, CTA 0)
    cyc = clk[0, hi, 0] - clk[0, lo, 0]
    ns = gt[0, hi, 0] - gt[0, lo, 0]
    ghz = float(cyc / ns) if ns > 0 else 1.9
    period = float((gt[:W, hi, 0] - gt[:W, lo, 0]).median() / (hi - lo)) / 1e3
    lines += [f"SM clock {ghz:.3f} GHz; step period seen by the workers: **{period:.2f} us**", "",
              "## Workers (median over CTAs and steady-state steps)", "",
              "| phase | role 0 us | role 1 us |", "|---|---|---|"]
    nph = 13
    tot = [0.0, 0.0]
    for ph in range(nph):
        row = []
        for role in (0, 1):
            ctas = [c for c in range(W) if (c & 1) == role] if plan["split"] else list(range(W))
            t1 = clk[ctas][:, lo:hi, ph]
            t0 = clk[ctas][:, lo:hi, ph - 1] if ph > 0 else None
            if ph == 0:
                # stamp 0 of step s follows stamp 11 of step s-1
                prev = clk[ctas][:, lo - 1:hi - 1, 12]
                dur = (clk[ctas][:, lo:hi, 0] - prev)
                # phase 0 itself is "top of step"; report the gap end-of-step -> top as part of wait
                val = float(dur.median() / ghz / 1e3)
            else:
                val = float((t1 - t0).median() / ghz / 1e3)
            row.append(val)
            tot[role] += val
        name = "(end of previous step -> top of step)" if ph == 0 else WORKER[ph - 1]
        lines.append(f"| {name} | {row[0]:.2f} | {row[1]:.2f} |")
    lines.append(f"| **sum** | **{tot[0]:.2f}** | **{tot[1]:.2f}** |")
    lines += ["", "## Service CTAs (median over steady-state steps, us)", "",
              "| CTA | kind | " + " | ".join(SERVICE) + " |", "|---|---|" + "---|" * len(SERVICE)]
    kinds = (["conv1 half 0", "conv1 half 1", "small fc + loss"] + ["conv2"] * plan["conv2_ctas"]
             + ["W3"] * plan["w3_ctas"])
    for i, kind in enumerate(kinds):
        c = W + i
        vals = [float((clk[c, lo:hi, ph + 1] - clk[c, lo:hi, ph]).median() / ghz / 1e3)
                for ph in range(4)]
        lines.append(f"| {c} | {kind} | " + " | ".join(f"{v:.2f}" for v in vals) + " |")
    lines += ["", "## Inside the service compute (us after the wait ended: staging done, compute done)", ""]
    for i, kind in enumerate(kinds):
        c = W + i
        vals = [float((clk[c, lo:hi, ph] - clk[c, lo:hi, 1]).median() / ghz / 1e3) for ph in (5, 6)]
        lines.append(f"* CTA {c} {kind}: " + ", ".join(f"{v:.2f}" for v in vals))
    # when, inside the step, does each group's publish land relative to the workers' step start
    lines += ["", "## Publish times relative to the median worker step start (us, global timer)", ""]
    w0 = gt[:W, lo:hi, 0].median(dim=0).values
    for name, ctas in (("conv1", [W, W + 1]), ("small fc", [W + 2]),
                       ("conv2 (last CTA)", list(range(W + 3, W + 3 + plan["conv2_ctas"]))),
                       ("W3 (last CTA)", list(range(W + 3 + plan["conv2_ctas"], plan["grid"])))):
        if not ctas:
            continue
        pub = gt[ctas][:, lo:hi, 4].max(dim=0).values
        lines.append(f"* {name}: step s published {float((pub - w0).median()) / 1e3:.2f} us after "
                     "the workers started step s")
    wend = gt[:W, lo:hi, 11].max(dim=0).values
    lines.append(f"* last worker finishes step s {float((wend - w0).median()) / 1e3:.2f} us after "
                 "the median worker started it")
    # start-up: when did each of the first steps begin, relative to the first stamp of the launch
    t_first = gt[:W, 0, 0].min()
    starts = [float((gt[:W, i, 0].median() - t_first) / 1e3) for i in range(min(10, S))]
    lines += ["", "## Start of the first steps after the first CTA's first stamp (us): "
              + ", ".join(f"{v:.1f}" for v in starts)]
    # per-step trace of the link between two steps (global timer, us after the median worker
    # started the step): when the workers' conv1 rows were complete, when the conv1 service CTA
    # had them summed / exchanged / published, when the workers had the new conv1 weights
    lines += ["", "## Per-step trace of the conv1 link (us after the median worker started step s)", "",
              "| step | period | median worker: conv1 wgrad done | slowest worker: conv1 wgrad done | "
              "median worker: bubble done | service: rows summed | service: exchanged + SGD | "
              "service: published | median worker: has conv1 weights of step s (next step) |",
              "|---|---|---|---|---|---|---|---|---|"]
    for st in range(lo, min(hi, lo + 12)):
        o = gt[:W, st, 0].median()
        f = lambda v: f"{float(v - o) / 1e3:.2f}"
        nxt = gt[:W, st + 1, 1].median() if st + 1 < d.shape[1] else o
        lines.append(f"| {st} | {float(gt[:W, st + 1, 0].median() - o) / 1e3:.2f} | {f(gt[:W, st, 11].median())} | "
                     f"{f(gt[:W, st, 11].max())} | {f(gt[:W, st, 12].median())} | {f(gt[W, st, 1])} | "
                     f"{f(gt[W, st, 3])} | {f(gt[W, st, 4])} | {f(nxt)} |")
    text = "\n".join(lines) + "\n"
    if rank == 0:
        print(text)
    if a.out:      # every rank reports its own CTAs
        Path(a.out if rank == 0 else f"{a.out}.rank{rank}").write_text(text)
    if comm is not None:
        P.destroy_process_group()


if __name__ == "__main__":
    main()
