"""Parent-side checks shared by the CPU (gloo) and GPU (peer-memory) DDP tests."""
import torch

from csed_514_project_distributed_training_using_pytorch_b200.models import Net, oracle
from csed_514_project_distributed_training_using_pytorch_b200.utils.flat import FlatParams
from tests import _workers


def check_ddp_generic(res, world, lr=0.05, mu=0.5):
    # T2: after construction every rank holds rank 0's parameters
    for r in range(world):
        assert torch.equal(res[r]["init"], res[0]["init"])
        assert res[r]["keys"] == ["0.weight", "0.bias", "2.weight", "2.bias", "4.weight", "4.bias"]
    # oracle: one process, mean of the per-rank gradients, SGD-momentum
    model = _workers._small_mlp(100)
    flat = FlatParams(model)
    assert torch.equal(flat.param, res[0]["init"])
    buf = torch.zeros_like(flat.param)
    for step in range(4):
        gsum = torch.zeros_like(flat.param)
        for r in range(world):
            g = torch.Generator().manual_seed(1000 * step + r)
            x = torch.randn(8, 20, generator=g)
            y = torch.randint(0, 5, (8,), generator=g)
            flat.grad.zero_()
            flat.rebind_grads()
            torch.nn.functional.cross_entropy(model(x), y).backward()
            gsum += flat.grad
        buf.mul_(mu).add_(gsum / world)
        with torch.no_grad():
            flat.param.sub_(lr * buf)
    for r in range(world):
        torch.testing.assert_close(res[r]["final"], flat.param.detach(), rtol=2e-4, atol=2e-5)
        assert torch.equal(res[r]["final"], res[0]["final"])   # replica divergence exactly 0.0
    if world > 1:  # no_sync left different local gradients on different ranks
        assert not torch.equal(res[0]["local_grad"], res[1]["local_grad"])


def check_net_ddp(res, world, lr=0.02, mu=0.5, B=16, steps=4, check_grads=True):
    for r in range(world):
        assert torch.equal(res[r]["init"], res[0]["init"])
        assert res[r]["steps"] == steps
    torch.manual_seed(1)
    net = Net()
    flat = FlatParams(net)
    assert torch.equal(flat.param, res[0]["init"])
    params = [p.detach().double().requires_grad_(True) for p in net.param_list()]
    bufs = [torch.zeros_like(p) for p in params]
    for step in range(steps):
        gsum = [torch.zeros_like(p) for p in params]
        for r in range(world):
            g = torch.Generator().manual_seed(1000 * step + r)
            x = torch.randn(B, 1, 28, 28, generator=g)
            y = torch.randint(0, 10, (B,), generator=g)
            m2, m3 = oracle.make_masks(B, g)
            for p in params:
                p.grad = None
            loss = oracle.net_loss(params, x.double(), y, m2.double(), m3.double())
            loss.backward()
            for a, p in zip(gsum, params):
                a += p.grad
            torch.testing.assert_close(res[r]["losses"][step], loss.detach().float(),
                                       rtol=1e-4, atol=1e-5)
        mean = [a / world for a in gsum]
        if check_grads:   # gradients after the fused reducer == mean of per-rank oracle grads
            for r in range(world):
                got = res[r]["grads"][step]
                for s, m in zip(flat.slots, mean):
                    torch.testing.assert_close(got[s.offset:s.offset + s.numel].view(s.shape),
                                               m.float(), rtol=2e-4, atol=2e-5)
        oracle.sgd_momentum_step(params, mean, bufs, lr, mu)
    for r in range(world):
        for s, p in zip(flat.slots, params):
            torch.testing.assert_close(res[r]["final"][s.offset:s.offset + s.numel].view(s.shape),
                                       p.detach().float(), rtol=2e-4, atol=2e-5)
        assert torch.equal(res[r]["final"], res[0]["final"])   # bit-identical replicas
