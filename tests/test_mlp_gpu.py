"""Wide MLP on tcgen05 GEMMs vs the plain PyTorch fp32 statement of the same network."""
import pytest
import torch
import torch.nn.functional as F

from csed_514_project_distributed_training_using_pytorch_b200.models.mlp import (
    WideMLP, LocalSink, MasterSGD)
from csed_514_project_distributed_training_using_pytorch_b200.utils.flat import FlatParams

pytestmark = pytest.mark.gpu


def _ref_loss_and_grads(model, x, y):
    ps = [p.detach().float().cpu().requires_grad_(True) for p in model.parameters()]
    h = x.float().cpu()
    n = len(ps) // 2
    for i in range(n):
        h = h @ ps[2 * i].t() + ps[2 * i + 1]
        if i < n - 1:
            h = torch.relu(h).to(torch.bfloat16).float()   # activations are stored in bf16
    loss = F.cross_entropy(h, y.cpu())
    loss.backward()
    return loss.detach(), [p.grad for p in ps], h.detach()


@pytest.mark.parametrize("sizes,B", [((784, 512, 256, 10), 64), ((784, 4096, 4096, 10), 128)])
def test_mlp_forward_backward(cuda_device, sizes, B):
    torch.manual_seed(0)
    model = WideMLP(sizes).to(cuda_device)
    g = torch.Generator().manual_seed(1)
    x = torch.randn(B, 784, generator=g).to(torch.bfloat16).to(cuda_device)
    y = torch.randint(0, 10, (B,), generator=g).to(cuda_device)
    want_loss, want_grads, want_logits = _ref_loss_and_grads(model, x, y)
    logits = model(x)
    torch.testing.assert_close(logits.cpu(), want_logits, rtol=3e-2, atol=3e-2)
    loss = model.loss(x, y)
    loss.backward()
    torch.testing.assert_close(loss.detach().cpu(), want_loss, rtol=2e-2, atol=2e-2)
    for p, wg in zip(model.parameters(), want_grads):
        got = p.grad.float().cpu()
        # bf16 gradients of a bf16 backward: compare in relative Frobenius norm
        err = (got - wg).norm() / (wg.norm() + 1e-12)
        assert err < 3e-2, f"relative gradient error {err:.4f} for shape {tuple(wg.shape)}"


def test_mlp_sink_and_master_sgd_learn(cuda_device):
    torch.manual_seed(0)
    model = WideMLP((784, 256, 256, 10)).to(cuda_device)
    flat = FlatParams(model, with_momentum=False)
    model.sink = LocalSink(flat)
    opt = MasterSGD(flat, lr=0.05, momentum=0.5)
    g = torch.Generator().manual_seed(2)
    x = torch.randn(256, 784, generator=g).to(torch.bfloat16).to(cuda_device)
    y = torch.randint(0, 10, (256,), generator=g).to(cuda_device)
    losses = []
    for _ in range(30):
        loss = model.loss(x, y)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    assert losses[-1] < losses[0] - 0.25 and losses[-1] < losses[10] < losses[0], losses
