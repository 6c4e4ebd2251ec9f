"""Numerics of every `Net` kernel against the plain-PyTorch fp32/fp64 oracle."""
import pytest
import torch

from csed_514_project_distributed_training_using_pytorch_b200.models import Net
from csed_514_project_distributed_training_using_pytorch_b200.models import oracle
from csed_514_project_distributed_training_using_pytorch_b200.ops import net_ops
from csed_514_project_distributed_training_using_pytorch_b200.ops.fused import (
    FusedNetEngine, X_F32, X_U8, X_U8_GATHER)

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-4, 2e-5  # fp32 SIMT path; oracle is fp64 on CPU


def _net(device, seed=1):
    torch.manual_seed(seed)
    return Net().to(device)


def _oracle_params(net, dtype=torch.float64):
    return [p.detach().cpu().to(dtype).requires_grad_(True) for p in net.param_list()]


def _data(B, seed=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, 1, 28, 28, generator=g)
    y = torch.randint(0, 10, (B,), generator=g)
    return x, y


@pytest.mark.parametrize("B", [1, 8, 16, 32, 64, 1000])
def test_layered_forward_eval(cuda_device, B):
    net = _net(cuda_device).eval()
    x, _ = _data(B)
    with torch.no_grad():
        got = net(x.to(cuda_device)).cpu()
    want = oracle.net_forward(_oracle_params(net), x.double()).float()
    torch.testing.assert_close(got, want.detach(), rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("B", [1, 32, 64, 200])
def test_layered_forward_backward_with_masks(cuda_device, B):
    net = _net(cuda_device).train()
    x, y = _data(B, seed=3)
    m2, m3 = oracle.make_masks(B, torch.Generator().manual_seed(7))
    for p in net.parameters():
        p.grad = None
    logp = net(x.to(cuda_device), masks=(m2.to(cuda_device), m3.to(cuda_device)))
    loss = net_ops.nll_loss(logp, y.to(cuda_device))
    loss.backward()

    op = _oracle_params(net)
    lo = oracle.net_loss(op, x.double(), y, m2.double(), m3.double())
    lo.backward()
    torch.testing.assert_close(loss.detach().cpu(), lo.detach().float(), rtol=RTOL, atol=ATOL)
    for name, p, q in zip(Net.PARAM_ORDER, net.param_list(), op):
        torch.testing.assert_close(p.grad.cpu(), q.grad.float(), rtol=2e-4, atol=2e-5,
                                   msg=lambda m, n=name: f"{n}: {m}")


def test_nll_sum_and_cross_entropy(cuda_device):
    net = _net(cuda_device).eval()
    x, y = _data(50)
    with torch.no_grad():
        logp = net(x.to(cuda_device))
    s = net_ops.nll_loss(logp, y.to(cuda_device), size_average=False).cpu()
    want = torch.nn.functional.nll_loss(logp.cpu(), y, reduction="sum")
    torch.testing.assert_close(s, want, rtol=1e-5, atol=1e-5)
    ce = net_ops.cross_entropy_on_log_probs(logp, y.to(cuda_device)).cpu()
    want_ce = torch.nn.CrossEntropyLoss()(logp.cpu(), y)  # src/train_dist.py:67,82
    torch.testing.assert_close(ce, want_ce, rtol=1e-5, atol=1e-6)


def test_layered_dropout_statistics(cuda_device):
    net = _net(cuda_device).train()
    B = 512
    x, _ = _data(B)
    a1 = net_ops.conv1_pool_relu(x.to(cuda_device), net.conv1.weight, net.conv1.bias)
    m2 = torch.empty(B, 20, device=cuda_device)
    a2 = net_ops.conv2_drop_pool_relu(a1, net.conv2.weight, net.conv2.bias, m2, 1, 0.5, 123, 5)
    vals = set(m2.unique().cpu().tolist())
    assert vals <= {0.0, 2.0} and len(vals) == 2          # T9: {0, 1/(1-p)}
    assert abs(m2.mean().item() - 1.0) < 0.06             # keep-probability 0.5
    dropped = (m2 == 0).view(B, 20, 1, 1).expand_as(a2)
    assert (a2[dropped] == 0).all()                       # whole channels are zeroed
    m3 = torch.empty(B, 50, device=cuda_device)
    net_ops.fc1_relu_drop(a2.view(B, 320), net.fc1.weight, net.fc1.bias, m3, 1, 0.5, 123, 5)
    assert abs(m3.mean().item() - 1.0) < 0.05
    # a different step gives a different mask, the same step the same mask
    m2b = torch.empty(B, 20, device=cuda_device)
    net_ops.conv2_drop_pool_relu(a1, net.conv2.weight, net.conv2.bias, m2b, 1, 0.5, 123, 6)
    assert not torch.equal(m2, m2b)
    m2c = torch.empty(B, 20, device=cuda_device)
    net_ops.conv2_drop_pool_relu(a1, net.conv2.weight, net.conv2.bias, m2c, 1, 0.5, 123, 5)
    assert torch.equal(m2, m2c)


def test_gather_normalize_exact(cuda_device):
    g = torch.Generator().manual_seed(0)
    data = torch.randint(0, 256, (500, 784), dtype=torch.uint8, generator=g)
    idx = torch.randint(0, 500, (64,), generator=g).to(torch.int32)
    got = net_ops.gather_normalize(data.to(cuda_device), idx.to(cuda_device)).cpu()
    want = oracle.normalize_u8(data[idx.long()])
    assert torch.equal(got, want)  # same IEEE ops in the same order
    got_all = net_ops.gather_normalize(data.to(cuda_device)).cpu()
    assert torch.equal(got_all, oracle.normalize_u8(data))


@pytest.mark.parametrize("B", [1, 8, 16, 32, 64, 200])
def test_fused_step_matches_oracle(cuda_device, B):
    """One whole step (forward, loss, backward, SGD) with injected masks == oracle."""
    net = _net(cuda_device).train()
    ref = _oracle_params(net)
    eng = FusedNetEngine(net, lr=0.01, momentum=0.5)
    bufs = [torch.zeros_like(p) for p in ref]
    for it in range(3):
        x, y = _data(B, seed=10 + it)
        m2, m3 = oracle.make_masks(B, torch.Generator().manual_seed(20 + it))
        eng.step(x.to(cuda_device), y.to(cuda_device),
                 masks=(m2.to(cuda_device), m3.to(cuda_device)), write_grad=True)
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
        got_loss = eng.last_losses(1)[0]
        torch.testing.assert_close(got_loss, lo.detach().float(), rtol=1e-4, atol=1e-5)
    assert eng.steps_done() == 3


def test_fused_input_modes_agree(cuda_device):
    g = torch.Generator().manual_seed(0)
    data = torch.randint(0, 256, (300, 784), dtype=torch.uint8, generator=g)
    labels = torch.randint(0, 10, (300,), generator=g)
    idx = torch.randperm(300, generator=g)[:64].to(torch.int32)
    xb = data[idx.long()]
    yb = labels[idx.long()]
    outs = []
    for mode in (X_F32, X_U8, X_U8_GATHER):
        net = _net(cuda_device).train()
        eng = FusedNetEngine(net, dropout=False)
        if mode == X_F32:
            eng.step(oracle.normalize_u8(xb).to(cuda_device), yb.to(cuda_device), X_F32)
        elif mode == X_U8:
            eng.step(xb.to(cuda_device), yb.to(cuda_device), X_U8)
        else:
            eng.step(data.to(cuda_device), labels.to(cuda_device), X_U8_GATHER,
                     gidx=idx.to(cuda_device))
        outs.append(eng.flat.param.clone())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])


def test_fused_graph_replay_equals_eager(cuda_device):
    B = 64
    xs = [_data(B, seed=s) for s in range(5)]
    # eager
    net_a = _net(cuda_device).train()
    eng_a = FusedNetEngine(net_a, seed=9)
    for x, y in xs:
        eng_a.step(x.to(cuda_device), y.to(cuda_device))
    # graph
    net_b = _net(cuda_device).train()
    eng_b = FusedNetEngine(net_b, seed=9)
    sx = torch.empty(B, 1, 28, 28, device=cuda_device)
    sy = torch.empty(B, dtype=torch.int64, device=cuda_device)
    sx.copy_(xs[0][0]); sy.copy_(xs[0][1])
    graph = eng_b.capture(sx, sy)
    assert eng_b.steps_done() == 0  # capture leaves the state untouched
    for x, y in xs:
        sx.copy_(x); sy.copy_(y)
        graph.replay()
    torch.cuda.synchronize()
    assert eng_b.steps_done() == 5
    assert torch.equal(eng_a.flat.param, eng_b.flat.param)  # same kernels, same masks: bit-equal
    assert torch.equal(eng_a.last_losses(5), eng_b.last_losses(5))
    # parameters are views of the flat buffer, so the module sees the update
    assert torch.equal(net_b.fc2.bias.detach(), eng_b.flat.views(eng_b.flat.param)[7])


def test_fused_is_deterministic(cuda_device):
    x, y = _data(64, seed=4)
    res = []
    for _ in range(2):
        net = _net(cuda_device).train()
        eng = FusedNetEngine(net, seed=3)
        for _ in range(4):
            eng.step(x.to(cuda_device), y.to(cuda_device))
        res.append(eng.flat.param.clone())
    assert torch.equal(res[0], res[1])


def test_eval_kernel(cuda_device):
    net = _net(cuda_device).eval()
    eng = FusedNetEngine(net)
    x, y = _data(1000, seed=5)
    logp = torch.empty(1000, 10, device=cuda_device)
    out = eng.evaluate_batch(x.to(cuda_device), y.to(cuda_device), logp_out=logp).cpu()
    want = oracle.net_forward(_oracle_params(net), x.double()).float().detach()
    torch.testing.assert_close(logp.cpu(), want, rtol=RTOL, atol=ATOL)
    want_loss = torch.nn.functional.nll_loss(want, y, reduction="sum")  # src/train.py:94
    want_correct = (want.argmax(1) == y).sum()                           # src/train.py:95-96
    torch.testing.assert_close(out[0], want_loss, rtol=1e-4, atol=1e-3)
    assert int(out[1].item()) == int(want_correct)
    out2 = eng.evaluate_batch(x.to(cuda_device), y.to(cuda_device), accumulate=True).cpu()
    assert int(out2[1].item()) == 2 * int(want_correct)


def test_sgd_kernel_matches_torch(cuda_device):
    from csed_514_project_distributed_training_using_pytorch_b200.ops.sgd import FlatSGD
    torch.manual_seed(0)
    n = 4096 + 8
    p = torch.randn(n)
    ref_p = p.clone().requires_grad_(True)
    opt = torch.optim.SGD([ref_p], lr=0.02, momentum=0.5)
    dp = p.to(cuda_device)
    dg = torch.zeros(n, device=cuda_device)
    sgd = FlatSGD(dp, dg, lr=0.02, momentum=0.5)
    for _ in range(4):
        g = torch.randn(n)
        ref_p.grad = g.clone()
        opt.step()
        dg.copy_(g)
        sgd.step()
    torch.testing.assert_close(dp.cpu(), ref_p.detach(), rtol=1e-6, atol=1e-6)


def test_native_runner_matches_python_loop(cuda_device):
    """csrc/runner.cu (staging slots, copy/compute overlap, 16-step chunk graphs with the
    copies as graph nodes, per-step submission for the remainder) == the Python step loop."""
    from csed_514_project_distributed_training_using_pytorch_b200.trainer import Trainer
    from csed_514_project_distributed_training_using_pytorch_b200.utils.config import TrainConfig
    B, steps = 64, 60
    g = torch.Generator().manual_seed(0)
    hx = torch.randn(steps * B, 784, generator=g).pin_memory()
    hy = torch.randint(0, 10, (steps * B,), generator=g).pin_memory()
    out = []
    for native in (True, False):
        tr = Trainer(TrainConfig(learning_rate=0.02), distributed_style=True)
        tr.native_runner = native
        tr.persistent = False     # this test pins the graph runner; tests/test_persist_gpu.py has the feeder
        l1 = tr.train_steps_from_host(hx, hy, 12, B)                     # per-step path only
        torch.cuda.synchronize()
        l1 = l1.clone()            # the trainer reuses one pinned loss buffer between calls
        l2 = tr.train_steps_from_host(hx[12 * B:], hy[12 * B:], 41, B)   # 2 chunks + 9 single steps
        torch.cuda.synchronize()
        l2 = l2.clone()
        l3 = tr.train_steps_from_host(hx[53 * B:], hy[53 * B:], 7, B)    # slots wrap after chunks
        torch.cuda.synchronize()
        if native:
            assert tr._graphs[("runner", B, hx.dtype)].chunks_launched() == 2
        out.append((tr.engine.flat.param.clone(), torch.cat([l1, l2, l3]).clone(),
                    tr.engine.steps_done()))
    assert out[0][2] == out[1][2] == steps
    assert torch.equal(out[0][0], out[1][0])
    assert torch.equal(out[0][1], out[1][1])
    assert (out[0][1] > 1.5).all() and (out[0][1] < 3.5).all()


@pytest.mark.parametrize("B", [1, 2, 7, 64, 200, 512, 1000])
def test_conv2_tc_implicit_gemm_matches_oracle(cuda_device, B):
    """conv2 forward as a TMA-fed implicit GEMM without im2col (csrc/conv2_tc.cu, TF32) vs the
    fp64 oracle of the layer and vs the SIMT kernel, arg-max positions included."""
    torch.manual_seed(1)
    net = Net()
    g = torch.Generator().manual_seed(B)
    a1 = torch.relu(torch.randn(B, 10, 12, 12, generator=g))
    m2 = oracle.make_masks(B, g)[0]
    conv = torch.nn.functional.conv2d(a1.double(), net.conv2.weight.double(), net.conv2.bias.double())
    want = torch.relu(torch.nn.functional.max_pool2d(conv * m2.double().view(B, 20, 1, 1), 2)).float()
    args = (a1.to(cuda_device), net.conv2.weight.to(cuda_device), net.conv2.bias.to(cuda_device),
            m2.to(cuda_device), 2)
    got = net_ops.conv2_drop_pool_relu(*args, impl="tc")
    torch.testing.assert_close(got.cpu(), want, rtol=5e-3, atol=5e-3)   # TF32 inputs
    simt = net_ops.conv2_drop_pool_relu(*args)
    torch.testing.assert_close(got, simt, rtol=5e-3, atol=5e-3)


@pytest.mark.parametrize("B", [1, 3, 64, 200, 512, 1000])
def test_conv2_tc_backward_matches_oracle(cuda_device, B):
    """conv2 data, weight and bias gradients on tcgen05 (csrc/conv2_tc.cu: implicit GEMM over
    a zero-padded dZ frame; MN-major pixel-reduction GEMM) vs fp64 autograd of the same layer
    with the same arg-max positions, and vs the SIMT kernel (autograd of src/model.py:24)."""
    torch.manual_seed(2)
    net = Net()
    g = torch.Generator().manual_seed(100 + B)
    a1 = torch.relu(torch.randn(B, 10, 12, 12, generator=g))
    m2 = oracle.make_masks(B, g)[0]
    da2 = torch.randn(B, 20, 4, 4, generator=g)
    w, b = net.conv2.weight.detach(), net.conv2.bias.detach()
    dev = cuda_device
    # forward state from the fp32-exact kernel: a2 and the arg-max position of every window
    N = net_ops.N
    a2 = torch.empty(B, 20, 4, 4, device=dev)
    pos2 = torch.empty(B, 20, 4, 4, dtype=torch.uint8, device=dev)
    a1d, wd_, bd, md = a1.to(dev), w.to(dev), b.to(dev), m2.to(dev)
    N.call("b200_conv2_fwd", N.ptr(a1d), N.ptr(wd_), N.ptr(bd), N.ptr(md), 2, 0.5, 0, 0,
           N.ptr(a2), N.ptr(pos2), B, N.stream())
    da1, dw, db = net_ops.conv2_backward_tc(a1d, wd_, a2, pos2, md, da2.to(dev))
    # fp64 reference: dense dZ from the same positions, then autograd of the convolution
    p = pos2.cpu().long()
    live = (a2.cpu() > 0).double() * m2.double().view(B, 20, 1, 1) * da2.double()
    dz = torch.zeros(B, 20, 96, dtype=torch.float64)
    dz.scatter_(2, p.view(B, 20, 16), live.view(B, 20, 16))
    dz = dz.view(B, 20, 8, 12)[..., :8]
    a1r, wr = a1.double().requires_grad_(True), w.double().requires_grad_(True)
    torch.nn.functional.conv2d(a1r, wr).backward(dz)

    def close(got, want, what):
        scale = float(want.abs().max()) + 1e-12
        err = float((got.cpu().double() - want).abs().max()) / scale
        assert err < 4e-3, f"{what}: max error {err:.2e} of the largest entry"      # TF32 operands
    close(da1, a1r.grad, "da1")
    close(dw, wr.grad, "dw")
    close(db, dz.sum(dim=(0, 2, 3)), "db")
    # and the SIMT backward from the same saved state
    ws = net_ops.workspace(dev)
    dw_s, db_s, da1_s = torch.empty_like(wd_), torch.empty(20, device=dev), torch.empty_like(a1d)
    N.call("b200_conv2_bwd", N.ptr(a1d), N.ptr(wd_), N.ptr(da2.to(dev)), N.ptr(a2), N.ptr(pos2), N.ptr(md),
           N.ptr(dw_s), N.ptr(db_s), N.ptr(da1_s), N.ptr(ws.scratch), N.ptr(ws.counter), B, N.stream())
    close(da1, da1_s.cpu().double(), "da1 vs SIMT")
    close(dw, dw_s.cpu().double(), "dw vs SIMT")


@pytest.mark.parametrize("B", [1, 2, 7, 64, 1000])
def test_conv2_tcgen05_matches_oracle(cuda_device, B):
    """conv2 + dropout2d + pool + relu on tcgen05 (TF32) vs the fp32 oracle of the layer."""
    torch.manual_seed(1)
    net = Net()
    g = torch.Generator().manual_seed(B)
    a1 = torch.relu(torch.randn(B, 10, 12, 12, generator=g))
    m2 = oracle.make_masks(B, g)[0]
    conv = torch.nn.functional.conv2d(a1.double(), net.conv2.weight.double(), net.conv2.bias.double())
    want = torch.relu(torch.nn.functional.max_pool2d(conv * m2.double().view(B, 20, 1, 1), 2)).float()
    got = net_ops.conv2_drop_pool_relu(a1.to(cuda_device), net.conv2.weight.to(cuda_device),
                                       net.conv2.bias.to(cuda_device), m2.to(cuda_device), 2,
                                       impl="tcgen05")
    torch.testing.assert_close(got.cpu(), want, rtol=5e-3, atol=5e-3)   # TF32 inputs
    # and it agrees with the SIMT kernel to TF32 accuracy, gradients flow through either
    simt = net_ops.conv2_drop_pool_relu(a1.to(cuda_device), net.conv2.weight.to(cuda_device),
                                        net.conv2.bias.to(cuda_device), m2.to(cuda_device), 2)
    torch.testing.assert_close(got, simt, rtol=5e-3, atol=5e-3)


def test_net_with_tcgen05_conv2_trains(cuda_device):
    torch.manual_seed(1)
    net = Net(conv2_impl="tcgen05").to(cuda_device).train()
    x, y = _data(64, seed=8)
    m2, m3 = oracle.make_masks(64, torch.Generator().manual_seed(9))
    logp = net(x.to(cuda_device), masks=(m2.to(cuda_device), m3.to(cuda_device)))
    loss = net_ops.nll_loss(logp, y.to(cuda_device))
    loss.backward()
    op = _oracle_params(net)
    lo = oracle.net_loss(op, x.double(), y, m2.double(), m3.double())
    lo.backward()
    torch.testing.assert_close(loss.detach().cpu(), lo.detach().float(), rtol=5e-3, atol=5e-3)
    for p_, q in zip(net.param_list(), op):
        err = (p_.grad.cpu() - q.grad.float()).norm() / (q.grad.float().norm() + 1e-12)
        assert err < 2e-2


@pytest.mark.parametrize("B", [64, 512])
def test_net_with_tc_conv2_follows_oracle_loss_curve(cuda_device, B):
    """The layered path with conv2 forward, data gradient and weight gradient all on tcgen05
    (Net(conv2_impl="tc"), TF32 operands): gradients of the first step against the fp64 oracle,
    then 100 SGD-momentum steps on a learnable 10-template set whose loss curve stays on the
    oracle's (the loop of src/train.py:66-76)."""
    torch.manual_seed(1)
    net = Net(conv2_impl="tc").to(cuda_device).train()
    ref = _oracle_params(net)
    bufs = [torch.zeros_like(p) for p in ref]
    opt = torch.optim.SGD(net.parameters(), lr=0.02, momentum=0.5)
    g = torch.Generator().manual_seed(3)
    templates = torch.randn(10, 784, generator=g)
    ours, theirs = [], []
    for it in range(100):
        y = torch.randint(0, 10, (B,), generator=g)
        x = (templates[y] + 0.5 * torch.randn(B, 784, generator=g)).view(B, 1, 28, 28)
        m2, m3 = oracle.make_masks(B, torch.Generator().manual_seed(100 + it))
        opt.zero_grad(set_to_none=True)
        loss = net_ops.nll_loss(net(x.to(cuda_device), masks=(m2.to(cuda_device), m3.to(cuda_device))),
                                y.to(cuda_device))
        loss.backward()
        for p in ref:
            p.grad = None
        lo = oracle.net_loss(ref, x.double(), y, m2.double(), m3.double())
        lo.backward()
        if it == 0:
            for name, p_, q in zip(Net.PARAM_ORDER, net.param_list(), ref):
                err = (p_.grad.cpu() - q.grad.float()).norm() / (q.grad.float().norm() + 1e-12)
                # TF32 products move a few pooling arg-max positions, which reroutes those gradients
                assert err < 3e-2, f"first-step gradient of {name}: relative error {float(err):.2e}"
        opt.step()
        oracle.sgd_momentum_step(ref, [q.grad for q in ref], bufs, 0.02, 0.5)
        ours.append(float(loss.detach()))
        theirs.append(float(lo.detach()))
    worst = max(abs(a - b) / max(abs(b), 0.05) for a, b in zip(ours, theirs))
    assert worst < 5e-2, f"loss curves part by {worst:.2e}: {ours[-3:]} vs {theirs[-3:]}"
    assert sum(ours[-10:]) / 10 < 0.5 * sum(ours[:10]) / 10


def test_conv2_auto_uses_tensor_core_backward(cuda_device):
    """conv2_impl="auto": fp32 forward below batch 2048 (bit-equal to "simt"), tcgen05 backward."""
    torch.manual_seed(3)
    net = Net()
    B = 96
    g = torch.Generator().manual_seed(5)
    a1 = torch.relu(torch.randn(B, 10, 12, 12, generator=g)).to(cuda_device)
    m2 = oracle.make_masks(B, g)[0].to(cuda_device)
    da2 = torch.randn(B, 20, 4, 4, generator=g).to(cuda_device)
    grads = {}
    for impl in ("simt", "auto"):
        a = a1.clone().requires_grad_(True)
        w = net.conv2.weight.detach().to(cuda_device).requires_grad_(True)
        b = net.conv2.bias.detach().to(cuda_device).requires_grad_(True)
        out = net_ops.conv2_drop_pool_relu(a, w, b, m2, 2, impl=impl)
        out.backward(da2)
        grads[impl] = (out.detach(), a.grad, w.grad, b.grad)
    assert torch.equal(grads["simt"][0], grads["auto"][0])
    for got, want in zip(grads["auto"][1:], grads["simt"][1:]):
        assert float((got - want).abs().max()) < 4e-3 * float(want.abs().max())
        assert not torch.equal(got, want)          # TF32 products: the tensor-core kernels ran


def test_resident_epoch_multistep_graph_equals_eager(cuda_device):
    """Device-side index stepping + 16-step graphs (PDL across steps) == one eager step at a time."""
    from csed_514_project_distributed_training_using_pytorch_b200.ops.fused import X_U8_GATHER
    g = torch.Generator().manual_seed(0)
    n, B = 64 * 40, 64
    data = torch.randint(0, 256, (n, 784), dtype=torch.uint8, generator=g).to(cuda_device)
    labels = torch.randint(0, 10, (n,), generator=g).to(cuda_device)
    perm = torch.randperm(n, generator=g).to(torch.int32).to(cuda_device)
    out = []
    for mode in ("eager", "graph16"):
        eng = FusedNetEngine(_net(cuda_device).train(), lr=0.02, momentum=0.5, seed=5)
        eng.begin_epoch()
        if mode == "eager":
            for _ in range(40):
                eng.step(data, labels, X_U8_GATHER, gidx=perm, batch=B, idx_stride=B)
        else:
            g16 = eng.capture(data, labels, X_U8_GATHER, gidx=perm, batch=B, idx_stride=B, unroll=16)
            g1 = eng.capture(data, labels, X_U8_GATHER, gidx=perm, batch=B, idx_stride=B)
            g16.replay(); g16.replay()
            for _ in range(8):
                g1.replay()
        torch.cuda.synchronize()
        assert eng.steps_done() == 40
        out.append((eng.flat.param.clone(), eng.last_losses(40).clone()))
    assert torch.equal(out[0][0], out[1][0])
    assert torch.equal(out[0][1], out[1][1])
