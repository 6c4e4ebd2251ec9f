"""ctypes binding to the in-tree native library (csrc/ -> lib/libb200dp.so).

The library is a plain C ABI (declared in include/b200dp.h): raw device pointers, sizes
and a CUDA stream handle go in, an integer status comes out. PyTorch stays the
framework layer: tensors are allocated by torch (or wrap our symmetric buffers) and
only their `data_ptr()` crosses the boundary, always on torch's current stream so the
calls compose with autograd, streams and CUDA-graph capture.

There is no silent fallback on a GPU: `lib()` raises if the .so is missing, and every
CUDA op in this package goes through it.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import torch

_LIB = None
_LIB_PATH = Path(os.environ.get("B200DP_LIB") or
                 (Path(__file__).resolve().parent / "lib" / "libb200dp.so"))

c_p = C.c_void_p
c_i = C.c_int
c_u = C.c_uint
c_f = C.c_float
c_ll = C.c_longlong
c_ull = C.c_ulonglong

# name -> argtypes (restype is always int unless listed in _RESTYPES)
_SIGS = {
    "b200_version": [],
    "b200_net_layout": [C.POINTER(c_i)],
    "b200_net_scratch_floats": [c_i],
    "b200_net_max_rows": [],
    "b200_net_fused_rows": [c_i],
    "b200_gather_normalize": [c_p, c_p, c_p, c_i, c_p],
    "b200_conv1_fwd": [c_p, c_p, c_p, c_p, c_p, c_i, c_p],
    "b200_conv1_bwd": [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_p],
    "b200_conv2_fwd": [c_p, c_p, c_p, c_p, c_i, c_f, c_ull, c_ull, c_p, c_p, c_i, c_p],
    "b200_conv2_fwd_tc": [c_p, c_p, c_p, c_p, c_i, c_f, c_ull, c_ull, c_p, c_p, c_i, c_p],
    "b200_a1_to_c4": [c_p, c_p, c_i, c_p],
    "b200_conv2_w_to_operand": [c_p, c_p, c_p],
    "b200_conv2_operand_floats": [],
    "b200_conv2_frame_floats": [c_i],
    "b200_conv2_dgrad_operand_floats": [],
    "b200_conv2_dz_frame": [c_p, c_p, c_p, c_p, c_p, c_i, c_p],
    "b200_conv2_w_to_dgrad_operand": [c_p, c_p, c_p],
    "b200_conv2_dgrad_tc": [c_p, c_p, c_p, c_i, c_p],
    "b200_conv2_wgrad_tc": [c_p, c_p, c_p, c_p, c_i, c_p],
    "b200_conv2_c32_a1_floats": [c_i],
    "b200_conv2_c32_dz_floats": [c_i],
    "b200_a1_to_c32": [c_p, c_p, c_i, c_p],
    "b200_conv2_dz_c32": [c_p, c_p, c_p, c_p, c_p, c_i, c_p],
    "b200_conv2_dz_both": [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_p],
    "b200_conv2_fwd_tc2": [c_p, c_p, c_p, c_p, c_i, c_f, c_ull, c_ull, c_p, c_p, c_i, c_p],
    "b200_conv2_bwd": [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_p],
    "b200_fc1_fwd": [c_p, c_p, c_p, c_p, c_i, c_f, c_ull, c_ull, c_p, c_i, c_p],
    "b200_fc1_bwd": [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_p],
    "b200_fc2_lsm_fwd": [c_p, c_p, c_p, c_p, c_i, c_p],
    "b200_fc2_lsm_bwd": [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_p],
    "b200_nll_fwd": [c_p, c_p, c_p, c_i, c_i, c_p],
    "b200_nll_bwd": [c_p, c_p, c_p, c_i, c_i, c_p],
    "b200_sgd_momentum": [c_p, c_p, c_p, c_ll, c_p, c_f, c_p],
    "b200_net_fwdbwd": [c_p, c_p, c_i, c_p, c_p, c_i, c_p, c_p, c_i, c_p, c_p, c_ull, c_p, c_u, c_i,
                        c_i, c_p],
    "b200_net_reduce_sgd": [c_p, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_p, c_p],
    "b200_net_eval": [c_p, c_p, c_i, c_p, c_p, c_i, c_p, c_p, c_p, c_p, c_i, c_p],
}
_RESTYPES = {"b200_last_error": C.c_char_p}


class NativeError(RuntimeError):
    pass


def lib_path() -> Path:
    return _LIB_PATH


def available() -> bool:
    return _LIB_PATH.exists()


def register_sigs(sigs: dict, restypes: dict | None = None) -> None:
    """Let other modules (comm, gemm, runtime) declare their entry points."""
    _SIGS.update(sigs)
    if restypes:
        _RESTYPES.update(restypes)
    if _LIB is not None:
        _declare(_LIB, sigs, restypes or {})


def _declare(lib, sigs, restypes) -> None:
    for name, argtypes in sigs.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            continue
        fn.argtypes = argtypes
        fn.restype = restypes.get(name, c_i)
    for name, rt in restypes.items():
        try:
            getattr(lib, name).restype = rt
        except AttributeError:
            pass


def lib():
    """Load (once) and return the native library; raise loudly if it is not built."""
    global _LIB
    if _LIB is None:
        if not _LIB_PATH.exists():
            if os.environ.get("B200DP_AUTOBUILD", "0") == "1":
                from importlib import import_module
                import sys
                sys.path.insert(0, str(_LIB_PATH.parents[2]))
                import_module("build").build()
            else:
                raise NativeError(
                    f"native library not built: {_LIB_PATH} is missing. Run `python build.py` "
                    "(or __graft_entry__.build()) in the repo root.")
        l = C.CDLL(str(_LIB_PATH))
        l.b200_last_error.restype = C.c_char_p
        _declare(l, _SIGS, _RESTYPES)
        _LIB = l
    return _LIB


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().b200_last_error()
        raise NativeError(f"{what or 'native call'} failed (code {rc}): "
                          f"{msg.decode() if msg else '?'}")


def call(name: str, *args) -> None:
    check(getattr(lib(), name)(*args), name)


def ptr(t) -> int | None:
    """Device (or host) address of a tensor; None passes a null pointer."""
    if t is None:
        return None
    return t.data_ptr()


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream
