"""NVTX ranges around the phases a timeline should show (SURVEY §5 tracing plan): one range
per training epoch / host-fed run, per evaluation pass, per gradient-bucket launch and per
collective. The reference has no tracing at all (src/train_dist.py prints wall-clock per epoch,
:111-113); here `nsys`/`ncu --nvtx` group the kernels under these names.

Ranges cost ~0.2 us each and are on by default; B200DP_NVTX=0 turns them into no-ops.
On a machine without CUDA they are always no-ops.
"""
from __future__ import annotations

import contextlib
import os

import torch

_ON = os.environ.get("B200DP_NVTX", "1") != "0"
_depth = 0


def enabled() -> bool:
    return _ON and torch.cuda.is_available()


def push(name: str) -> None:
    global _depth
    if enabled():
        torch.cuda.nvtx.range_push(name)
        _depth += 1


def pop() -> None:
    global _depth
    if enabled() and _depth > 0:
        torch.cuda.nvtx.range_pop()
        _depth -= 1


@contextlib.contextmanager
def range(name: str):   # noqa: A001 - mirrors torch.cuda.nvtx.range
    push(name)
    try:
        yield
    finally:
        pop()


def depth() -> int:
    """Open ranges (tests use it to check that every push has its pop)."""
    return _depth
