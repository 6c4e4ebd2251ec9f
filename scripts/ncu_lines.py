"""Summarise an ncu report per CUDA source line: share of stall samples and of executed
instructions. Usage: python scripts/ncu_lines.py report.ncu-rep [top_n]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr = None; fpath = None; out = []
for r in rows:
    if not r: continue
    if r[0] == "File Path": fpath = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < 10 or r[0] in ("", "Function Name"): continue
    d = {}
    for k, v in zip(hdr, r):
        d.setdefault(k, v)
    try:
        smp = float(d["# Samples"]); ins = float(d["Instructions Executed"])
    except Exception:
        continue
    out.append((smp, ins, fpath, d["Line No"], r[1][:95]))
ts = sum(o[0] for o in out) or 1; ti = sum(o[1] for o in out) or 1
print(f"total samples {ts:.0f}, total warp-instructions {ti:.0f}")
for o in sorted(out, reverse=True)[:top]:
    print(f"{o[0]/ts*100:5.1f}% smp {o[1]/ti*100:5.1f}% ins  {o[2]}:{o[3]:>4}  {o[4]}")
