"""bf16 GEMM on tcgen05 tensor cores (csrc/gemm_tcgen05.cu) and the three products of a
linear layer built from it. No operand is transposed in memory: forward
uses K-major/K-major operands, dgrad a MN-major weight, wgrad two MN-major operands.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from .. import _native as N
N_ = N   # `N` is also the conventional name of a GEMM dimension below

c_p, c_i, c_ll = C.c_void_p, C.c_int, C.c_longlong
c_sz = C.c_size_t
N.register_sigs({
    "b200_gemm_bf16": [c_p, c_i, c_ll, c_p, c_i, c_ll, c_p, c_i, c_ll, c_i, c_i, c_i, c_p, c_i,
                       c_p, c_ll, c_p],
    "b200_gemm_bf16_rs": [c_p, c_i, c_ll, c_p, c_i, c_ll, c_p, c_i, c_ll, c_i, c_i, c_i, c_p, c_i,
                          c_p, c_ll, c_p, c_sz, c_sz, c_sz, c_sz, c_p, c_p],
    "b200_gemm_set_variant": [c_i, c_i],
    "b200_gemm_plan": [c_i, c_i, c_i, c_i, c_i, C.POINTER(c_i), C.POINTER(c_i), C.POINTER(c_i)],
})

VARIANTS = {"auto": -1, "cta": 0, "pair": 1, "persistent": 2, "splitk": 3}


def plan(M: int, N: int, K: int = 4096, rs: bool = False, both_mn: bool = False):
    """(kind, bn) the library would pick for an M x N x K product: kind "cta" | "pair" |
    "persistent" | "splitk" (persistent with every tile split in two along K). Host-only
    (works without a GPU)."""
    k, b, sp = c_i(), c_i(), c_i()
    N_.call("b200_gemm_plan", M, N, K, int(rs), int(both_mn), C.byref(k), C.byref(b), C.byref(sp))
    name = {v: n for n, v in VARIANTS.items()}[k.value]
    return ("splitk" if sp.value == 2 else name), b.value


def set_variant(kind="auto", bn: int = 0):
    """Pin the GEMM kernel for tests / benchmarks: "cta" (one 128 x bn tile per CTA), "pair"
    (one 256 x bn tile per CTA pair, cta_group::2), "persistent" (CTA pairs looping over
    tiles with double-buffered TMEM), "splitk" (persistent, each tile as two K halves on two
    CTA pairs) or "auto" (shape-based choice, the default)."""
    N.call("b200_gemm_set_variant", VARIANTS[kind] if isinstance(kind, str) else int(kind), int(bn))


def _chk(t: torch.Tensor, name: str):
    if not t.is_cuda or t.dtype != torch.bfloat16 or t.dim() != 2 or t.stride(1) != 1:
        raise N.NativeError(f"gemm operand {name}: need a CUDA bf16 matrix with unit inner stride")
    if t.stride(0) % 8 or t.data_ptr() % 16:
        raise N.NativeError(f"gemm operand {name}: row pitch must be a multiple of 8 elements "
                            "and the base 16-byte aligned (TMA)")


def gemm_bf16(A: torch.Tensor, B: torch.Tensor, a_mn: bool = False, b_mn: bool = False,
              out: Optional[torch.Tensor] = None, out_dtype=torch.bfloat16,
              bias: Optional[torch.Tensor] = None, relu: bool = False,
              relu_mask: Optional[torch.Tensor] = None, n_valid: Optional[int] = None,
              rs: Optional[dict] = None):
    """C[M,N] = op(A) op(B)^T, fp32 accumulate.
    a_mn=False: A is [M,K]; True: A is [K,M].  b_mn=False: B is [N,K]; True: B is [K,N].
    n_valid limits N (and M for a_mn) to a prefix of a padded operand.
    rs (from PeerComm.rs_target + 'base_vec'): fused reduce-scatter epilogue, the result is
    scattered to the owner ranks' staging instead of being written to `out`."""
    _chk(A, "A"); _chk(B, "B")
    if a_mn:
        K, M = A.shape
    else:
        M, K = A.shape
    if b_mn:
        Kb, Nn = B.shape
    else:
        Nn, Kb = B.shape
    if Kb != K:
        raise N.NativeError(f"gemm: reduction sizes differ ({K} vs {Kb})")
    if n_valid is not None:
        Nn = n_valid
    if out is None:
        out = torch.empty(M, Nn, dtype=out_dtype, device=A.device)
    if out.dim() != 2 or out.stride(1) != 1 or out.shape[0] < M or out.shape[1] < Nn:
        raise N.NativeError("gemm: bad output tensor")
    c_fp32 = 1 if out.dtype == torch.float32 else 0
    if not c_fp32 and out.dtype != torch.bfloat16:
        raise N.NativeError("gemm: output must be bf16 or fp32")
    if bias is not None and (bias.dtype != torch.bfloat16 or bias.numel() < Nn):
        raise N.NativeError("gemm: bias must be bf16 with at least N elements")
    if relu_mask is not None and (relu_mask.dtype != torch.bfloat16 or relu_mask.stride(1) != 1):
        raise N.NativeError("gemm: relu_mask must be a bf16 matrix")
    if rs is not None:
        N.call("b200_gemm_bf16_rs", N.ptr(A), int(a_mn), A.stride(0), N.ptr(B), int(b_mn),
               B.stride(0), N.ptr(out), c_fp32, out.stride(0), M, Nn, K, None, 0, None, 0,
               rs["comm"], rs["stage_off"], rs["src_bytes"], rs["slice_vec"], rs["base_vec"],
               rs["seq"], N.stream())
        return out
    N.call("b200_gemm_bf16", N.ptr(A), int(a_mn), A.stride(0), N.ptr(B), int(b_mn), B.stride(0),
           N.ptr(out), c_fp32, out.stride(0), M, Nn, K, N.ptr(bias), int(relu), N.ptr(relu_mask),
           relu_mask.stride(0) if relu_mask is not None else 0, N.stream())
    return out


def linear_fwd(x, W, b=None, relu=False, out=None, out_dtype=torch.bfloat16):
    """y = x W^T + b (optionally ReLU): x [B,K], W [N,K]."""
    return gemm_bf16(x, W, False, False, out=out, out_dtype=out_dtype, bias=b, relu=relu)


def linear_dgrad(dy, W, relu_mask=None, out=None):
    """dx = dy W, zeroed where relu_mask <= 0: dy [B,N], W [N,K] -> [B,K]."""
    return gemm_bf16(dy, W, False, True, out=out, relu_mask=relu_mask)


def linear_wgrad(dy, x, out=None, n_valid=None):
    """dW = dy^T x: dy [B,N], x [B,K] -> [N,K] (rows of dy's padding can be cut with n_valid...
    here n_valid cuts the OUTPUT rows when dy carries padded columns)."""
    if n_valid is not None:
        dy = dy[:, :n_valid] if dy.shape[1] != n_valid else dy
    return gemm_bf16(dy, x, True, True, out=out)
