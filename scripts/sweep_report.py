"""profiles/allreduce_sweep.md from the raw sweep JSON (scripts/microbench_comm.py sweep)."""
import json, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
LINK = 770.0   # measured peer-copy GB/s per direction per GPU (B200_PROFILING.md); 900 nominal
out = ["# Gradient all-reduce sweep (BASELINE.json config 5)", "",
       "fp32 buffers, in place, device-timed with CUDA events (max over ranks), Python launch",
       "overhead included for every implementation alike (~10 us per call; the Net bucket's fused",
       "kernel is timed inside CUDA graphs elsewhere). Bus bandwidth = 2(N-1)/N x bytes / time.",
       f"Roofline denominator: {LINK:.0f} GB/s measured peer-copy bandwidth per direction per GPU",
       "(900 nominal). one-shot = tagged-word LL kernel, two-shot = reduce-scatter + all-gather",
       "by peer stores, nvls = multimem.ld_reduce / multimem.st through the NVSwitch.", ""]
for n, f in ((8, "allreduce_sweep_n8_raw.json"), (2, "allreduce_sweep_n2_raw.json")):
    p = ROOT / "profiles" / f
    if not p.exists():
        continue
    txt = p.read_text(); d = json.loads(txt[txt.find("{\n"):txt.rfind("}") + 1])
    out += [f"## {n} GPUs", "",
            "| bytes | one-shot us | two-shot us | nvls us | NCCL us | best ours / NCCL | best bus GB/s | fraction of measured link |",
            "|---|---|---|---|---|---|---|---|"]
    for r in d["sweep"]:
        ours = {k: r[k] for k in ("oneshot_us", "twoshot_us", "nvls_us") if k in r}
        best = min(ours, key=ours.get)
        bus = 2 * (n - 1) / n * r["bytes"] / (ours[best] * 1e-6) / 1e9
        g = lambda k: f"{r[k]:.1f}" if k in r else "-"
        out.append(f"| {r['bytes']} | {g('oneshot_us')} | {g('twoshot_us')} | {g('nvls_us')} | {g('nccl_us')} | "
                   f"{r['nccl_us'] / ours[best]:.2f}x ({best[:-3]}) | {bus:.1f} | {bus / LINK:.2f} |")
    out.append("")
# crossover table consumed by PeerComm.choose_algo: per world size, the largest payload (bytes)
# for which the tagged one-shot kernel measured faster than the two-shot kernel
table = {}
for n, f in ((8, "allreduce_sweep_n8_raw.json"), (2, "allreduce_sweep_n2_raw.json")):
    p = ROOT / "profiles" / f
    if not p.exists():
        continue
    txt = p.read_text(); d = json.loads(txt[txt.find("{\n"):txt.rfind("}") + 1])
    best = 0
    for r in d["sweep"]:
        if "oneshot_us" in r and "twoshot_us" in r and r["oneshot_us"] <= r["twoshot_us"]:
            best = max(best, r["bytes"])
    table[str(n)] = {"oneshot_max_bytes": best}
(ROOT / "profiles" / "allreduce_crossover.json").write_text(json.dumps(table, indent=1) + "\n")
(ROOT / "profiles" / "allreduce_sweep.md").write_text("\n".join(out) + "\n")
print("\n".join(out))
