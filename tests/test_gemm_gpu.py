"""tcgen05 bf16 GEMM vs a plain PyTorch fp32 reference of the same product."""
import pytest
import torch

from csed_514_project_distributed_training_using_pytorch_b200.ops import gemm

pytestmark = pytest.mark.gpu


def _ref(A, B, a_mn, b_mn):
    a = A.float().t() if a_mn else A.float()
    b = B.float() if b_mn else B.float().t()
    return a @ b


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 384, 512), (64, 4096, 784), (104, 200, 72),
                                   (512, 16, 4096), (16, 1024, 520), (304, 136, 1000)])
def test_gemm_all_majors(cuda_device, a_mn, b_mn, M, N, K):
    g = torch.Generator().manual_seed(M * 7 + N * 3 + K)
    A = torch.randn((K, M) if a_mn else (M, K), generator=g).to(torch.bfloat16).to(cuda_device)
    B = torch.randn((K, N) if b_mn else (N, K), generator=g).to(torch.bfloat16).to(cuda_device)
    want = _ref(A, B, a_mn, b_mn)
    got32 = gemm.gemm_bf16(A, B, a_mn, b_mn, out_dtype=torch.float32)
    torch.testing.assert_close(got32, want, rtol=2e-3, atol=2e-2 * (K ** 0.5) / 8)
    got16 = gemm.gemm_bf16(A, B, a_mn, b_mn)
    torch.testing.assert_close(got16.float(), want, rtol=1.6e-2, atol=2e-2 * (K ** 0.5) / 4)


def test_gemm_epilogues(cuda_device):
    g = torch.Generator().manual_seed(5)
    M, N, K = 192, 320, 256
    x = torch.randn(M, K, generator=g).to(torch.bfloat16).to(cuda_device)
    W = (torch.randn(N, K, generator=g) / 16).to(torch.bfloat16).to(cuda_device)
    b = torch.randn(N, generator=g).to(torch.bfloat16).to(cuda_device)
    y = gemm.linear_fwd(x, W, b, relu=True)
    want = torch.relu(x.float() @ W.float().t() + b.float())
    torch.testing.assert_close(y.float(), want, rtol=1.6e-2, atol=3e-2)
    # dgrad with the ReLU mask of the producing activation fused in
    dy = torch.randn(M, N, generator=g).to(torch.bfloat16).to(cuda_device)
    act = torch.randn(M, K, generator=g).to(torch.bfloat16).to(cuda_device)
    dx = gemm.linear_dgrad(dy, W, relu_mask=act)
    want_dx = (dy.float() @ W.float()) * (act.float() > 0)
    torch.testing.assert_close(dx.float(), want_dx, rtol=1.6e-2, atol=5e-2)
    # wgrad
    dW = gemm.linear_wgrad(dy, x)
    want_dW = dy.float().t() @ x.float()
    torch.testing.assert_close(dW.float(), want_dW, rtol=1.6e-2, atol=0.15)
    # output into a strided slice of a larger (flat-bucket-like) buffer, fp32
    big = torch.zeros(N, K + 8, dtype=torch.float32, device=cuda_device)
    gemm.gemm_bf16(dy, x, True, True, out=big[:, :K])
    torch.testing.assert_close(big[:, :K], want_dW, rtol=2e-3, atol=2e-2)
    assert torch.count_nonzero(big[:, K:]) == 0


@pytest.fixture
def restore_variant():
    yield
    gemm.set_variant("auto")


# (kind, bn): the one-tile kernel at its widest tile, the CTA-pair kernel, the persistent kernel
_VARIANTS = [("cta", 256), ("pair", 256), ("pair", 128), ("persistent", 256), ("persistent", 128),
             ("splitk", 256)]


@pytest.mark.parametrize("kind,bn", _VARIANTS)
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("M,N,K", [(512, 512, 256), (304, 264, 520), (1024, 384, 72)])
def test_gemm_variants(cuda_device, restore_variant, kind, bn, a_mn, b_mn, M, N, K):
    g = torch.Generator().manual_seed(M + 3 * N + 5 * K)
    A = torch.randn((K, M) if a_mn else (M, K), generator=g).to(torch.bfloat16).to(cuda_device)
    B = torch.randn((K, N) if b_mn else (N, K), generator=g).to(torch.bfloat16).to(cuda_device)
    want = _ref(A, B, a_mn, b_mn)
    gemm.set_variant(kind, bn)
    got32 = gemm.gemm_bf16(A, B, a_mn, b_mn, out_dtype=torch.float32)
    torch.testing.assert_close(got32, want, rtol=2e-3, atol=2e-2 * (K ** 0.5) / 8)
    got16 = gemm.gemm_bf16(A, B, a_mn, b_mn)
    torch.testing.assert_close(got16.float(), want, rtol=1.6e-2, atol=2e-2 * (K ** 0.5) / 4)


@pytest.mark.parametrize("kind,bn", [("persistent", 256), ("persistent", 128), ("pair", 256)])
def test_gemm_many_tiles_per_cluster(cuda_device, restore_variant, kind, bn):
    """More tiles than CTA pairs: every pair walks several tiles, so the TMEM double buffer,
    the barrier phases and the smem ring wrap around (K = 3 blocks against 6-8 stages)."""
    g = torch.Generator().manual_seed(11)
    M, N, K = 4096 + 40, 4096 + 24, 192
    x = (torch.randn(M, K, generator=g)).to(torch.bfloat16).to(cuda_device)
    W = (torch.randn(N, K, generator=g)).to(torch.bfloat16).to(cuda_device)
    b = torch.randn(N, generator=g).to(torch.bfloat16).to(cuda_device)
    gemm.set_variant(kind, bn)
    y = gemm.linear_fwd(x, W, b, relu=True)
    want = torch.relu(x.float() @ W.float().t() + b.float())
    torch.testing.assert_close(y.float(), want, rtol=1.6e-2, atol=8e-2)
    act = torch.randn(M, K, generator=g).to(torch.bfloat16).to(cuda_device)
    dy = torch.randn(M, N, generator=g).to(torch.bfloat16).to(cuda_device)
    dx = gemm.linear_dgrad(dy, W, relu_mask=act)      # output is 192 wide: 128-column tiles, 65 K blocks
    want_dx = (dy.float() @ W.float()) * (act.float() > 0)
    torch.testing.assert_close(dx.float(), want_dx, rtol=1.6e-2, atol=0.6)
    dW = gemm.linear_wgrad(dy[:, :512].contiguous(), dy[:, 512:1024].contiguous())
    want_dW = dy[:, :512].float().t() @ dy[:, 512:1024].float()
    torch.testing.assert_close(dW.float(), want_dW, rtol=1.6e-2, atol=0.6)


@pytest.mark.parametrize("kind,bn", [("cta", 128), ("persistent", 256)])
def test_gemm_bias_view_with_odd_offset(cuda_device, restore_variant, kind, bn):
    """A bias that is a view at a 2-byte offset (not 16-byte aligned) takes the scalar path."""
    g = torch.Generator().manual_seed(3)
    M, N, K = 512, 512, 128
    x = torch.randn(M, K, generator=g).to(torch.bfloat16).to(cuda_device)
    W = (torch.randn(N, K, generator=g) / 8).to(torch.bfloat16).to(cuda_device)
    bbuf = torch.randn(N + 8, generator=g).to(torch.bfloat16).to(cuda_device)
    b = bbuf[1:N + 1]
    assert b.data_ptr() % 16 != 0
    gemm.set_variant(kind, bn)
    y = gemm.linear_fwd(x, W, b, relu=True)
    want = torch.relu(x.float() @ W.float().t() + b.float())
    torch.testing.assert_close(y.float(), want, rtol=1.6e-2, atol=3e-2)


def test_gemm_split_k_repeated_and_epilogues(cuda_device, restore_variant):
    """Split-K (opt-in variant): the two K halves of a tile run on two CTA pairs; the flags
    reset themselves, so back-to-back calls keep giving the product."""
    g = torch.Generator().manual_seed(21)
    M, N, K = 512, 4096, 4096
    x = (torch.randn(M, K, generator=g) / 8).to(torch.bfloat16).to(cuda_device)
    W = (torch.randn(N, K, generator=g) / 8).to(torch.bfloat16).to(cuda_device)
    b = torch.randn(N, generator=g).to(torch.bfloat16).to(cuda_device)
    gemm.set_variant("splitk")
    assert gemm.plan(M, N, K) == ("splitk", 256)
    want = torch.relu(x.float() @ W.float().t() + b.float())
    for _ in range(3):
        y = gemm.linear_fwd(x, W, b, relu=True)
        torch.testing.assert_close(y.float(), want, rtol=1.6e-2, atol=5e-2)
    y32 = gemm.linear_fwd(x, W, out_dtype=torch.float32)
    torch.testing.assert_close(y32, x.float() @ W.float().t(), rtol=2e-3, atol=2e-2)
    # identical to the unsplit kernel up to fp32 summation order
    gemm.set_variant("persistent", 256)
    y_ref = gemm.linear_fwd(x, W, out_dtype=torch.float32)
    torch.testing.assert_close(y32, y_ref, rtol=1e-4, atol=1e-3)
    # on a second stream (its own workspace slot), concurrently with the first
    gemm.set_variant("splitk")
    s2 = torch.cuda.Stream()
    s2.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s2):
        y2 = gemm.linear_fwd(x, W, out_dtype=torch.float32)
    y1 = gemm.linear_fwd(x, W, out_dtype=torch.float32)
    torch.cuda.current_stream().wait_stream(s2)
    assert torch.equal(y1, y2)
