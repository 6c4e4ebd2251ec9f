#!/usr/bin/env python
"""Build the native library in-tree: every csrc/*.cu and csrc/*.cpp is compiled for
sm_100a with nvcc and linked into <package>/lib/libb200dp.so.

The library exposes a plain C ABI (see include/b200dp.h) and links cudart statically,
so it loads on a machine without a GPU or a CUDA driver (host-only entry points such
as the bucket planner still work there). Objects are cached by mtime under build/.

Usage: python build.py [--force] [--verbose] [--jobs N]
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
EXTRA = os.environ.get("B200DP_NVCC_EXTRA", "").split()      # e.g. -DB200_MBAR_UNBOUNDED
SUFFIX = os.environ.get("B200DP_LIB_SUFFIX", "")            # build a side-by-side variant
PKG = "csed_514_project_distributed_training_using_pytorch_b200"
SRC = ROOT / "csrc"
OBJ = ROOT / "build" / ("obj" + SUFFIX)
OUT = ROOT / PKG / "lib" / f"libb200dp{SUFFIX}.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "-Xcompiler", "-Wall",
    "-Xcompiler", "-Wno-unused-function",
    "--expt-relaxed-constexpr",
    "-I", str(ROOT / "include"),
    "-I", str(SRC),
]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def headers_mtime() -> float:
    hs = list(SRC.glob("*.cuh")) + list(SRC.glob("*.h")) + list((ROOT / "include").glob("*.h"))
    return max((h.stat().st_mtime for h in hs), default=0.0)


def compile_one(nvcc: str, src: Path, force: bool, verbose: bool, hdr_mtime: float) -> Path:
    obj = OBJ / (src.name + ".o")
    if (not force and obj.exists() and obj.stat().st_mtime >= src.stat().st_mtime
            and obj.stat().st_mtime >= hdr_mtime):
        return obj
    cmd = [nvcc, *NVCC_FLAGS, *EXTRA, "-c", str(src), "-o", str(obj)]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
    if verbose and (r.stdout or r.stderr):
        print(r.stdout, r.stderr, flush=True)
    return obj


def build(force: bool = False, verbose: bool = False, jobs: int | None = None) -> Path:
    nvcc = find_nvcc()
    OBJ.mkdir(parents=True, exist_ok=True)
    OUT.parent.mkdir(parents=True, exist_ok=True)
    srcs = sorted(list(SRC.glob("*.cu")) + list(SRC.glob("*.cpp")))
    if not srcs:
        raise RuntimeError("no sources under csrc/")
    hm = headers_mtime()
    with cf.ThreadPoolExecutor(max_workers=jobs or min(8, os.cpu_count() or 4)) as ex:
        objs = list(ex.map(lambda s: compile_one(nvcc, s, force, verbose, hm), srcs))
    newest = max(o.stat().st_mtime for o in objs)
    if force or not OUT.exists() or OUT.stat().st_mtime < newest:
        tmp = OUT.with_suffix(".so.tmp%d" % os.getpid())
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static",
               "-o", str(tmp), *map(str, objs), "-ldl", "-lpthread", "-lrt"]
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        os.replace(tmp, OUT)
    return OUT


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--jobs", type=int, default=None)
    a = ap.parse_args()
    out = build(a.force, a.verbose, a.jobs)
    print(f"built {out} ({out.stat().st_size} bytes)")
    sys.exit(0)
