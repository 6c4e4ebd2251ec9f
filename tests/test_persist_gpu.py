"""The persistent step kernel (csrc/net_persist.cu) against the fp64 oracle, against the
two-kernel step, and against itself across its input modes and launch splits.

Reference behaviour being pinned: one loop body of src/train_dist.py:78-89 / src/train.py:71-76
(forward, nll_loss, backward, SGD-momentum step, loss read-back) per step."""
import pytest
import torch

from csed_514_project_distributed_training_using_pytorch_b200.models import Net, oracle
from csed_514_project_distributed_training_using_pytorch_b200.ops.fused import (
    FusedNetEngine, X_F32, X_U8, X_U8_GATHER, X_F32_GATHER)
from csed_514_project_distributed_training_using_pytorch_b200.ops.persist import (
    PersistentNetEngine, persist_plan)

pytestmark = pytest.mark.gpu


def _net(device, seed=1):
    torch.manual_seed(seed)
    return Net().to(device)


def _oracle_params(net):
    return [p.detach().cpu().double().requires_grad_(True) for p in net.param_list()]


def _data(B, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(B, 1, 28, 28, generator=g), torch.randint(0, 10, (B,), generator=g)


# 1..64: a CTA pair per image; 65..128: one CTA per image; above: several images per CTA
@pytest.mark.parametrize("B", [1, 8, 33, 64, 100, 200, 512])
def test_persist_step_matches_oracle(cuda_device, B):
    net = _net(cuda_device).train()
    ref = _oracle_params(net)
    eng = PersistentNetEngine(net, lr=0.01, momentum=0.5)
    bufs = [torch.zeros_like(p) for p in ref]
    for it in range(3):
        x, y = _data(B, seed=10 + it)
        m2, m3 = oracle.make_masks(B, torch.Generator().manual_seed(20 + it))
        eng.run_steps(x.to(cuda_device).view(B, 784), y.to(cuda_device), X_F32, batch=B, nsteps=1,
                      masks=(m2.to(cuda_device), m3.to(cuda_device)), write_grad=True)
        torch.cuda.synchronize()
        eng.check_errors()
        for p in ref:
            p.grad = None
        lo = oracle.net_loss(ref, x.double(), y, m2.double(), m3.double())
        lo.backward()
        for name, gv, q in zip(Net.PARAM_ORDER, eng.flat.views(eng.flat.grad), ref):
            torch.testing.assert_close(gv.cpu(), q.grad.float(), rtol=2e-4, atol=2e-5,
                                       msg=lambda m, n=name: f"step {it} grad {n}: {m}")
        oracle.sgd_momentum_step(ref, [q.grad for q in ref], bufs, 0.01, 0.5)
        for name, p, q in zip(Net.PARAM_ORDER, net.param_list(), ref):
            torch.testing.assert_close(p.detach().cpu(), q.detach().float(), rtol=1e-4, atol=1e-5,
                                       msg=lambda m, n=name: f"step {it} param {n}: {m}")
        torch.testing.assert_close(eng.last_losses(1)[0], lo.detach().float(), rtol=1e-4, atol=1e-5)
    assert eng.steps_done() == 3
    # the pads of the flat buffer never pick up garbage
    assert torch.isfinite(eng.flat.param).all() and torch.isfinite(eng.flat.momentum).all()


def test_persist_gradient_only_leaves_state_alone(cuda_device):
    net = _net(cuda_device).train()
    eng = PersistentNetEngine(net, dropout=False)
    before = eng.flat.param.clone()
    x, y = _data(16, seed=2)
    eng.run_steps(x.to(cuda_device).view(16, 784), y.to(cuda_device), X_F32, batch=16, nsteps=1,
                  write_grad=True, update=False)
    torch.cuda.synchronize()
    assert torch.equal(before, eng.flat.param)
    assert float(eng.flat.momentum.abs().max()) == 0.0
    assert float(eng.flat.grad.abs().max()) > 0.0


def _resident(cuda_device, n, seed=0):
    g = torch.Generator().manual_seed(seed)
    data = torch.randint(0, 256, (n, 784), dtype=torch.uint8, generator=g).to(cuda_device)
    labels = torch.randint(0, 10, (n,), generator=g).to(cuda_device)
    perm = torch.randperm(n, generator=g).to(torch.int32).to(cuda_device)
    return data, labels, perm


@pytest.mark.parametrize("B", [64, 160])
def test_persist_epoch_tracks_two_kernel_path(cuda_device, B):
    """40 steps with Philox dropout over a device-resident dataset: one persistent launch
    follows the two-kernel step (same masks, same data, different summation order)."""
    steps = 40
    data, labels, perm = _resident(cuda_device, B * steps)
    out = []
    for mode in ("two-kernel", "persistent"):
        cls = FusedNetEngine if mode == "two-kernel" else PersistentNetEngine
        eng = cls(_net(cuda_device).train(), lr=0.02, momentum=0.5, seed=5)
        eng.begin_epoch()
        if mode == "two-kernel":
            for _ in range(steps):
                eng.step(data, labels, X_U8_GATHER, gidx=perm, batch=B, idx_stride=B)
        else:
            eng.run_steps(data, labels, X_U8_GATHER, gidx=perm, batch=B, nsteps=steps, idx_stride=B)
            assert eng.launches == 1
        torch.cuda.synchronize()
        if mode == "persistent":
            eng.check_errors()
        assert eng.steps_done() == steps
        out.append((eng.flat.param.clone(), eng.last_losses(steps).clone()))
    torch.testing.assert_close(out[1][1], out[0][1], rtol=2e-4, atol=2e-5)
    torch.testing.assert_close(out[1][0], out[0][0], rtol=2e-3, atol=2e-5)
    assert (out[1][1] > 1.5).all() and (out[1][1] < 3.5).all()


def test_persist_launch_splits_and_modes_are_bit_equal(cuda_device):
    """The same 24 steps as one launch, as 7 + 17, and through the direct fp32 mode give
    bit-identical parameters and losses; two runs of the same launch too (deterministic)."""
    B, steps = 64, 24
    g = torch.Generator().manual_seed(3)
    xf = torch.randn(B * steps, 784, generator=g).to(cuda_device)
    yl = torch.randint(0, 10, (B * steps,), generator=g).to(cuda_device)
    ident = torch.arange(B * steps, dtype=torch.int32, device=cuda_device)
    res = []
    for mode in ("one", "one", "split", "direct"):
        eng = PersistentNetEngine(_net(cuda_device).train(), lr=0.02, momentum=0.5, seed=11)
        eng.begin_epoch()
        if mode == "one":
            eng.run_steps(xf, yl, X_F32_GATHER, gidx=ident, batch=B, nsteps=steps, idx_stride=B)
        elif mode == "split":
            eng.run_steps(xf, yl, X_F32_GATHER, gidx=ident, batch=B, nsteps=7, idx_stride=B)
            eng.run_steps(xf, yl, X_F32_GATHER, gidx=ident, batch=B, nsteps=17, idx_stride=B)
        else:
            eng.run_steps(xf, yl, X_F32, batch=B, nsteps=steps)
        torch.cuda.synchronize()
        eng.check_errors()
        assert eng.steps_done() == steps
        res.append((eng.flat.param.clone(), eng.last_losses(steps).clone()))
    for p, l in res[1:]:
        assert torch.equal(p, res[0][0])
        assert torch.equal(l, res[0][1])


def test_persist_mixes_with_two_kernel_steps(cuda_device):
    """A persistent run, a two-kernel step (the trainer's ragged last batch), another run:
    counters, loss ring and dropout streams stay in step."""
    B = 32
    data, labels, perm = _resident(cuda_device, B * 12, seed=4)
    eng = PersistentNetEngine(_net(cuda_device).train(), lr=0.02, momentum=0.5, seed=2)
    eng.begin_epoch()
    eng.run_steps(data, labels, X_U8_GATHER, gidx=perm, batch=B, nsteps=5, idx_stride=B)
    eng.step(data, labels, X_U8_GATHER, gidx=perm, batch=B, idx_stride=B)
    eng.run_steps(data, labels, X_U8_GATHER, gidx=perm, batch=B, nsteps=6, idx_stride=B)
    torch.cuda.synchronize()
    eng.check_errors()
    assert eng.steps_done() == 12
    ref = FusedNetEngine(_net(cuda_device).train(), lr=0.02, momentum=0.5, seed=2)
    ref.begin_epoch()
    for _ in range(12):
        ref.step(data, labels, X_U8_GATHER, gidx=perm, batch=B, idx_stride=B)
    torch.cuda.synchronize()
    torch.testing.assert_close(eng.last_losses(12), ref.last_losses(12), rtol=2e-4, atol=2e-5)
    torch.testing.assert_close(eng.flat.param, ref.flat.param, rtol=2e-3, atol=2e-5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.uint8])
def test_host_feeder_equals_device_run(cuda_device, dtype):
    """Pinned host -> ring -> running kernel (stream memory operations, ring wrap-around, three
    calls) is bit-identical to the same steps read from device memory, and every step's loss
    lands in the caller's pinned buffer."""
    B, steps = 64, 90          # ring of 32 slots: wraps twice, the copy stream has to wait
    g = torch.Generator().manual_seed(0)
    if dtype == torch.uint8:
        hx = torch.randint(0, 256, (steps * B, 784), dtype=torch.uint8, generator=g).pin_memory()
        mode = X_U8
    else:
        hx = torch.randn(steps * B, 784, generator=g).pin_memory()
        mode = X_F32
    hy = torch.randint(0, 10, (steps * B,), generator=g).pin_memory()
    a = PersistentNetEngine(_net(cuda_device).train(), lr=0.02, momentum=0.5, seed=7)
    a.run_steps(hx.to(cuda_device), hy.to(cuda_device), mode, batch=B, nsteps=steps)
    torch.cuda.synchronize()
    a.check_errors()
    b = PersistentNetEngine(_net(cuda_device).train(), lr=0.02, momentum=0.5, seed=7)
    f = b.feeder(B, dtype, ring_slots=32)
    hl = torch.full((steps,), -1.0).pin_memory()
    f.run(hx, hy, 5, hl[:5])
    f.run(hx[5 * B:], hy[5 * B:], 70, hl[5:75])
    f.run(hx[75 * B:], hy[75 * B:], 15, hl[75:])
    torch.cuda.synchronize()
    b.check_errors()
    assert f.launches() == 3 and b.steps_done() == steps
    assert torch.equal(a.flat.param, b.flat.param)
    assert torch.equal(a.last_losses(steps), hl)
    assert (hl > 1.5).all() and (hl < 3.5).all()


def test_trainer_host_steps_use_the_feeder(cuda_device):
    from csed_514_project_distributed_training_using_pytorch_b200.trainer import Trainer
    from csed_514_project_distributed_training_using_pytorch_b200.utils.config import TrainConfig
    B, steps = 64, 40
    g = torch.Generator().manual_seed(1)
    hx = torch.randn(steps * B, 784, generator=g).pin_memory()
    hy = torch.randint(0, 10, (steps * B,), generator=g).pin_memory()
    res = []
    for persistent in (True, False):
        tr = Trainer(TrainConfig(learning_rate=0.02), distributed_style=True)
        tr.persistent = persistent
        l = tr.train_steps_from_host(hx, hy, steps, B)
        torch.cuda.synchronize()      # asynchronous: the losses are valid only now
        if persistent:
            assert tr.engine.launches == 1        # one kernel for the whole call
        res.append((tr.engine.flat.param.clone(), l.clone(), tr.engine.steps_done()))
    assert res[0][2] == res[1][2] == steps
    torch.testing.assert_close(res[0][1], res[1][1], rtol=2e-4, atol=2e-5)
    torch.testing.assert_close(res[0][0], res[1][0], rtol=2e-3, atol=2e-5)


def test_persist_plan_host_query(native_lib):
    p = persist_plan(64)
    assert p["split"] == 1 and p["workers"] == 128 and p["grid"] == 147 and p["rows"] == 64
    assert p["service"] == 3 + p["w3_ctas"] and p["conv2_ctas"] == 0   # 2 conv1 halves + small fc + W3;
    # conv2's rows are summed by the workers between two steps
    assert p["w3_ctas"] * p["w3_cols"] >= 320 and p["w3_cols"] % 4 == 0
    p = persist_plan(100)
    assert p["split"] == 0 and p["workers"] == 100 and p["images_per_worker"] == 1
    p = persist_plan(512)
    assert p["workers"] == 128 and p["images_per_worker"] == 4 and p["grid"] <= 148


def test_hundred_step_loss_curve_follows_oracle(cuda_device):
    """100 optimizer steps on a learnable 10-template set, explicit dropout masks: the loss
    curve of the persistent kernel stays on the fp64 oracle's (the loop
    being pinned is src/train.py:66-76), and both curves go down."""
    B, steps = 64, 100
    g = torch.Generator().manual_seed(3)
    templates = torch.randn(10, 784, generator=g)
    net = _net(cuda_device, seed=4).train()
    ref = _oracle_params(net)
    eng = PersistentNetEngine(net, lr=0.02, momentum=0.5)
    bufs = [torch.zeros_like(p) for p in ref]
    ours, theirs = [], []
    for it in range(steps):
        y = torch.randint(0, 10, (B,), generator=g)
        x = (templates[y] + 0.5 * torch.randn(B, 784, generator=g)).view(B, 1, 28, 28)
        m2, m3 = oracle.make_masks(B, torch.Generator().manual_seed(100 + it))
        eng.run_steps(x.to(cuda_device).view(B, 784), y.to(cuda_device), X_F32, batch=B, nsteps=1,
                      masks=(m2.to(cuda_device), m3.to(cuda_device)))
        for p in ref:
            p.grad = None
        lo = oracle.net_loss(ref, x.double(), y, m2.double(), m3.double())
        lo.backward()
        oracle.sgd_momentum_step(ref, [q.grad for q in ref], bufs, 0.02, 0.5)
        theirs.append(float(lo))
    torch.cuda.synchronize()
    eng.check_errors()
    ours = eng.last_losses(steps).tolist()
    worst = max(abs(a - b) / max(abs(b), 1e-3) for a, b in zip(ours, theirs))
    assert worst < 2e-3, f"loss curves part by {worst:.2e}: {ours[-3:]} vs {theirs[-3:]}"
    assert sum(theirs[-10:]) / 10 < 0.5 * sum(theirs[:10]) / 10
    for name, p, q in zip(Net.PARAM_ORDER, net.param_list(), ref):
        torch.testing.assert_close(p.detach().cpu(), q.detach().float(), rtol=5e-3, atol=2e-4,
                                   msg=lambda m, n=name: f"param {n} after {steps} steps: {m}")
