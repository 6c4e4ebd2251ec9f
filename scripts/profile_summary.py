"""Turn an .ncu-rep into a committed markdown summary (profiles/): headline raw metrics per
kernel and the hottest source lines. Usage: profile_summary.py report.ncu-rep out.md [title]"""
import csv, io, subprocess, sys
rep, out = sys.argv[1], sys.argv[2]
title = sys.argv[3] if len(sys.argv) > 3 else rep
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.max", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio"]
lines = [f"# {title}", "", f"Source: `ncu --set full --clock-control none --import-source on` on one B200; report `{rep.split('/')[-1]}` (not committed, regenerate with the command in the file header of scripts/profile_summary.py).", ""]
for vals in rows[2:]:
    d = dict(zip(hdr, vals)); u = dict(zip(hdr, units))
    lines += [f"## `{d['Kernel Name'][:90]}`", "", "| metric | value | unit |", "|---|---|---|"]
    for k in want:
        if k in d and d[k] != "":
            lines.append(f"| {k} | {d[k]} | {u.get(k,'')} |")
    lines.append("")
src = subprocess.run([sys.executable, __file__.replace("profile_summary.py", "ncu_lines.py"), rep, "25"],
                     capture_output=True, text=True).stdout
lines += ["## Hottest source lines (all captured kernels; share of stall samples / of executed warp instructions)", "", "```", src.rstrip(), "```", ""]
open(out, "w").write("\n".join(lines))
print("wrote", out)
