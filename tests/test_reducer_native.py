"""The native (C++) bucket planner and reducer core, on CPU (csrc/reducer.cpp)."""
import pytest
import torch

from csed_514_project_distributed_training_using_pytorch_b200.parallel.reducer import (
    ReducerCore, plan_buckets, plan_buckets_by_group)


def test_planner_matches_torch_ddp_rule(native_lib):
    # Net: all eight tensors in one bucket (`[[0..7]]`, 87,360 B)
    net = [250 * 4, 10 * 4, 5000 * 4, 20 * 4, 16000 * 4, 50 * 4, 500 * 4, 10 * 4]
    assert plan_buckets(net) == [0] * 8
    # wide MLP in bf16: after the rebuild torch uses two buckets,
    # [fc3.b, fc3.w, fc2.b, fc2.w] then [fc1.b, fc1.w]
    mlp = [784 * 4096 * 2, 4096 * 2, 4096 * 4096 * 2, 4096 * 2, 4096 * 10 * 2, 10 * 2]
    assert plan_buckets(mlp) == [1, 1, 0, 0, 0, 0]
    # same rule as torch's own planner on an arbitrary list
    sizes = [3 << 20, 100, 30 << 20, 2 << 20, 512, 700 << 10, 400 << 10]
    params = [torch.empty(s // 4) for s in sizes]
    import torch.distributed as dist
    ref, _ = dist._compute_bucket_assignment_by_size(list(reversed(params)), [1 << 20, 25 << 20])
    want = [0] * len(sizes)
    for b, idxs in enumerate(ref):
        for i in idxs:
            want[len(sizes) - 1 - i] = b
    assert plan_buckets(sizes) == want
    assert plan_buckets_by_group(6, [0, 2, 4]) == [2, 2, 1, 1, 0, 0]
    with pytest.raises(Exception):
        plan_buckets_by_group(6, [1, 2])


def test_reducer_core_ready_counting(native_lib):
    core = ReducerCore([1, 1, 0, 0, 0, 0])
    core.reset()
    assert [core.mark_ready(i) for i in (5, 4, 3)] == [-1, -1, -1]
    assert core.pending_buckets() == [0, 1]
    assert core.mark_ready(2) == 0                 # bucket 0 complete
    assert core.mark_ready(2) == -2                # marked twice -> caller raises
    assert core.mark_ready(99) == -3
    assert not core.all_done()
    assert core.mark_ready(1) == -1 and core.mark_ready(0) == 1
    assert core.all_done() and core.completion_order() == [0, 1]
    core.reset()
    assert core.pending_buckets() == [0, 1] and core.completion_order() == []


def test_gemm_kernel_plan_host_only(native_lib):
    """Shape -> kernel choice of the tcgen05 GEMM (csrc/gemm_tcgen05.cu choose()); host-only."""
    from csed_514_project_distributed_training_using_pytorch_b200.ops import gemm
    try:
        # the wide-MLP products at per-GPU batch 512 (BASELINE.json config 4)
        assert gemm.plan(4096, 4096, 512) == ("persistent", 256)        # fc2 wgrad
        assert gemm.plan(4096, 4096, 512, rs=True, both_mn=True) == ("persistent", 256)
        assert gemm.plan(4096, 784, 512, rs=True, both_mn=True) == ("persistent", 256)   # fc1 wgrad
        # problems that fit one wave of 100-148 one-tile CTAs take the one-tile kernel
        # (in-kernel times: profiles/r2/gemm_small_ncu.md)
        assert gemm.plan(512, 4096, 4096) == ("cta", 128)               # fc2 forward / dgrad
        assert gemm.plan(512, 4096, 784) == ("cta", 128)
        assert gemm.plan(4096, 784, 512) == ("cta", 256)
        assert gemm.plan(1024, 4096, 4096) == ("persistent", 256)       # long K: CTA pairs again
        # the reduce-scatter epilogue of other layouts stays on the one-CTA kernel
        assert gemm.plan(4096, 4096, 512, rs=True, both_mn=False) == ("cta", 128)
        # skinny problems: one CTA per tile, tile width follows N
        assert gemm.plan(64, 4096, 4096) == ("cta", 128)
        assert gemm.plan(4096, 10, 4096) == ("cta", 16)
        assert gemm.plan(512, 40, 4096) == ("cta", 64)
        # pinning
        gemm.set_variant("pair", 128)
        assert gemm.plan(4096, 4096, 4096) == ("pair", 128)
        assert gemm.plan(64, 4096, 4096)[0] == "cta"                    # pairs need M > 128
        gemm.set_variant("cta", 256)
        assert gemm.plan(4096, 4096, 4096) == ("cta", 256)
        assert gemm.plan(100, 100, 100) == ("cta", 128)
        gemm.set_variant("splitk")
        assert gemm.plan(512, 512, 256) == ("splitk", 256)
        assert gemm.plan(4096, 4096, 4096) == ("persistent", 256)       # too many tiles to split
    finally:
        gemm.set_variant("auto")
    assert gemm.plan(4096, 4096, 4096) == ("persistent", 256)


def test_native_struct_layouts_match_python(native_lib):
    """ctypes mirrors of native structs have the library's size (host-only ABI check)."""
    import ctypes as C
    from csed_514_project_distributed_training_using_pytorch_b200 import _native as N
    from csed_514_project_distributed_training_using_pytorch_b200.ops.runner import _RunnerConfig
    lib = N.lib()
    assert C.sizeof(_RunnerConfig) == lib.b200_runner_config_size()
    # CommDev is passed by value to every collective kernel: it must fit the 4 KB of kernel
    # parameters comfortably and hold 8 peer base pointers
    n = lib.b200_comm_struct_size()
    assert 8 * 8 <= n <= 512


def test_header_declares_every_exported_entry_point(native_lib):
    """include/b200dp.h is the C contract: every b200_* symbol the library exports is declared
    there (INTEGRATION.md 'C ABI')."""
    import re
    import subprocess
    from pathlib import Path
    from csed_514_project_distributed_training_using_pytorch_b200 import _native as N
    root = Path(__file__).resolve().parents[1]
    header = (root / "include" / "b200dp.h").read_text()
    declared = set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", header))
    out = subprocess.run(["nm", "-D", "--defined-only", str(N.lib_path())], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("nm is not available")
    exported = {ln.split()[-1] for ln in out.stdout.splitlines() if " T b200_" in ln}
    assert exported, "no b200_* symbols exported?"
    missing = sorted(exported - declared)
    assert not missing, f"exported but not declared in include/b200dp.h: {missing}"
