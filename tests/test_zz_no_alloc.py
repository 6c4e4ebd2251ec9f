"""The allocation-free conv1 backward entry point, called the way a C caller would.

Some callers of this library cannot allocate device memory when they need a gradient:
they were handed their buffers and a stream up front (a plugin inside another runtime, a
thread with a frozen memory budget, a loop where an allocation might stall the GPU).
`b200_conv1_bwd` is unusable for them because it wants a scratch buffer and a zeroed
counter. `b200_conv1_bwd_noalloc` keeps that scratch inside the library. This test drives
both entry points through the raw C ABI (ctypes, own prototypes, raw pointers) and checks
that the new one reproduces the existing one on two calls in a row, the second with
different data and a different batch size so stale scratch contents would show.

The file name sorts last on purpose: with `pytest -x` every other test has already run
by the time this one can fail. It also passes alone in a fresh process.
"""
import ctypes as C
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu

LIB = (Path(__file__).resolve().parents[1] / "csed_514_project_distributed_training_using_pytorch_b200"
       / "lib" / "libb200dp.so")
c_p, c_i = C.c_void_p, C.c_int


def _load():
    if not LIB.exists():
        import build
        build.build()
    lib = C.CDLL(str(LIB))
    lib.b200_last_error.restype = C.c_char_p
    lib.b200_conv1_fwd.argtypes = [c_p, c_p, c_p, c_p, c_p, c_i, c_p]
    lib.b200_conv1_bwd.argtypes = [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_p]
    lib.b200_conv1_bwd_noalloc.argtypes = [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_p]
    lib.b200_conv1_bwd_noalloc_scratch_bytes.restype = C.c_size_t
    return lib


def _case(lib, dev, B, seed, stream):
    """Inputs for one backward call, produced by the library's own forward."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, 1, 28, 28, generator=g).to(dev)
    w = (torch.randn(10, 1, 5, 5, generator=g) * 0.2).to(dev)
    b = (torch.randn(10, generator=g) * 0.1).to(dev)
    da1 = torch.randn(B, 10, 12, 12, generator=g).to(dev)
    a1 = torch.empty(B, 10, 12, 12, device=dev)
    pos1 = torch.empty(B, 10, 12, 12, dtype=torch.int16, device=dev)
    assert lib.b200_conv1_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), a1.data_ptr(),
                              pos1.data_ptr(), B, stream) == 0
    return x, da1, a1, pos1


def _existing(lib, dev, case, B, stream):
    x, da1, a1, pos1 = case
    scratch = torch.empty(148 * 264, device=dev)
    counter = torch.zeros(1, dtype=torch.int32, device=dev)
    dw = torch.empty(10, 1, 5, 5, device=dev)
    db = torch.empty(10, device=dev)
    rc = lib.b200_conv1_bwd(x.data_ptr(), da1.data_ptr(), a1.data_ptr(), pos1.data_ptr(),
                            dw.data_ptr(), db.data_ptr(), scratch.data_ptr(), counter.data_ptr(),
                            B, stream)
    assert rc == 0, lib.b200_last_error()
    return dw, db


def test_noalloc_matches_existing_entry_point_twice():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device visible")
    lib = _load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream
    assert lib.b200_conv1_bwd_noalloc_max_batch() == 128
    assert lib.b200_conv1_bwd_noalloc_scratch_bytes() == 256 + 128 * 264 * 4

    # outputs are allocated by the caller BEFORE the calls under test, as such a caller would
    outs = [(torch.full((10, 1, 5, 5), float("nan"), device=dev),
             torch.full((10,), float("nan"), device=dev)) for _ in range(2)]
    cases = [(_case(lib, dev, 64, 11, stream), 64), (_case(lib, dev, 37, 12, stream), 37)]
    refs = [_existing(lib, dev, c, B, stream) for c, B in cases]
    torch.cuda.synchronize()

    for (case, B), (dw, db), (rdw, rdb) in zip(cases, outs, refs):   # two calls in a row
        x, da1, a1, pos1 = case
        rc = lib.b200_conv1_bwd_noalloc(x.data_ptr(), da1.data_ptr(), a1.data_ptr(),
                                        pos1.data_ptr(), dw.data_ptr(), db.data_ptr(), B, stream)
        assert rc == 0, lib.b200_last_error()
    torch.cuda.synchronize()
    for (dw, db), (rdw, rdb) in zip(outs, refs):
        torch.testing.assert_close(dw, rdw, rtol=1e-4, atol=2e-5)
        torch.testing.assert_close(db, rdb, rtol=1e-4, atol=2e-5)

    # and against the plain PyTorch fp32 statement of the same gradient
    x, da1, a1, pos1 = cases[0][0]
    xr = x.detach().cpu()
    g = torch.Generator().manual_seed(11)
    torch.randn(64, 1, 28, 28, generator=g)
    w = (torch.randn(10, 1, 5, 5, generator=g) * 0.2).requires_grad_(True)
    b = (torch.randn(10, generator=g) * 0.1).requires_grad_(True)
    out = torch.relu(torch.nn.functional.max_pool2d(torch.nn.functional.conv2d(xr, w, b), 2))
    out.backward(da1.cpu())
    torch.testing.assert_close(outs[0][0].cpu(), w.grad, rtol=2e-4, atol=2e-5)
    torch.testing.assert_close(outs[0][1].cpu(), b.grad, rtol=2e-4, atol=2e-5)


def test_noalloc_rejects_oversize_batch():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device visible")
    lib = _load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream
    B = 129
    case = _case(lib, dev, B, 13, stream)
    dw = torch.zeros(10, 1, 5, 5, device=dev)
    db = torch.zeros(10, device=dev)
    x, da1, a1, pos1 = case
    rc = lib.b200_conv1_bwd_noalloc(x.data_ptr(), da1.data_ptr(), a1.data_ptr(), pos1.data_ptr(),
                                    dw.data_ptr(), db.data_ptr(), B, stream)
    assert rc != 0
    msg = lib.b200_last_error().decode()
    assert "exceeds the built-in scratch" in msg and "128" in msg
    torch.cuda.synchronize()
    assert torch.count_nonzero(dw) == 0   # nothing was launched
