"""Peer-memory collectives on real GPUs (needs >= 2; run with gpurun --gpus 2)."""
import pytest
import torch

from csed_514_project_distributed_training_using_pytorch_b200.parallel import launch
from tests import _workers
from tests.test_comm_cpu import _check_collectives

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _world():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    return 8 if n >= 8 else (4 if n >= 4 else 2)


def test_p2p_demo_peer(native_lib, tmp_path):
    _world()
    launch(_workers.p2p_demo, 2, args=("b200", str(tmp_path)), timeout=300)
    assert torch.load(tmp_path / "p2p_1.pt")[0].item() == 1.0


def test_collectives_peer(native_lib, tmp_path):
    world = _world()
    launch(_workers.collectives, world, args=("b200", str(tmp_path)), timeout=600)
    res = [torch.load(tmp_path / f"coll_{r}.pt") for r in range(world)]
    _check_collectives(res, world, bf16=True)


@pytest.mark.parametrize("fuse,algo", [(False, "auto"), (True, "oneshot"), (True, "twoshot")])
def test_ddp_generic_peer(native_lib, tmp_path, fuse, algo):
    from tests._ddp_checks import check_ddp_generic
    world = _world()
    launch(_workers.ddp_generic, world, args=("b200", str(tmp_path), fuse, algo), timeout=600)
    res = [torch.load(tmp_path / f"ddp_{r}.pt") for r in range(world)]
    check_ddp_generic(res, world)


@pytest.mark.parametrize("use_graph", [False, True, "persist"])
def test_net_ddp_fused_peer(native_lib, tmp_path, use_graph):
    """Gradients == mean of the per-rank oracle gradients, parameters == oracle SGD, replicas
    bit-identical: two-kernel step (eager, graph) and the persistent kernel."""
    from tests._ddp_checks import check_net_ddp
    world = _world()
    launch(_workers.net_ddp_fused, world, args=("b200", str(tmp_path), use_graph), timeout=600)
    res = [torch.load(tmp_path / f"net_{r}.pt") for r in range(world)]
    check_net_ddp(res, world, check_grads=use_graph is not True)


def test_net_ddp_persistent_multistep_peer(native_lib, tmp_path):
    """24 steps per rank in two persistent launches: replicas bit-identical, momentum too, and
    the trajectory follows the two-kernel step (different summation order only)."""
    world = _world()
    launch(_workers.net_ddp_persist_multi, world, args=("b200", str(tmp_path)), timeout=600)
    res = [torch.load(tmp_path / f"netp_{r}.pt") for r in range(world)]
    for r in range(world):
        p, k = res[r]["persistent"], res[r]["two-kernel"]
        assert p["steps"] == k["steps"] == 24 and p["launches"] == 2
        assert torch.equal(p["final"], res[0]["persistent"]["final"])
        assert torch.equal(p["mom"], res[0]["persistent"]["mom"])
        torch.testing.assert_close(p["losses"], k["losses"], rtol=2e-4, atol=2e-5)
        torch.testing.assert_close(p["final"], k["final"], rtol=2e-3, atol=2e-5)
        assert (p["losses"] > 1.5).all() and (p["losses"] < 3.5).all()


@pytest.mark.parametrize("algo", ["oneshot", "twoshot", "nvls", "auto", "twoshot+pp", "auto+pp"])
def test_mlp_ddp_peer(native_lib, tmp_path, algo):
    """bf16 MLP: replicas stay bit-identical; result tracks an fp32 single-process oracle.
    "+pp": ping-pong parameter buffers (weight gradient and bucket kernel before the data
    gradient), an odd number of steps so that the second buffer is the live one at the end."""
    nsteps = 5 if algo.endswith("+pp") else 6
    from csed_514_project_distributed_training_using_pytorch_b200.models.mlp import WideMLP
    from csed_514_project_distributed_training_using_pytorch_b200.utils.flat import FlatParams
    world = _world()
    launch(_workers.mlp_ddp, world, args=("b200", str(tmp_path), algo), timeout=600)
    res = [torch.load(tmp_path / f"mlp_{r}.pt") for r in range(world)]
    for r in range(world):
        assert torch.equal(res[r]["init"], res[0]["init"])
        assert torch.equal(res[r]["final"], res[0]["final"])
        assert res[r]["order"] == [0, 1, 2]          # buckets complete back-to-front
    # fp32 master weights, reassembled from the ranks' shards
    got_master = res[0]["init"].clone()
    for r in range(world):
        for m, base, lo, hi in res[r]["master"]:
            n = min(m.numel(), hi - lo - base)
            if n > 0:
                got_master[lo + base:lo + base + n] = m[:n]
    # Oracle of the same schedule that rounds where the device rounds: bf16 weights from the
    # fp32 master, bf16 hidden activations, bf16 dlogits / dx / dW / db, fp32 accumulation of
    # the ranks' bf16 gradients, SGD-momentum on the fp32 master.
    bf = lambda t: t.to(torch.bfloat16).float()
    torch.manual_seed(50)
    model = WideMLP((784, 256, 128, 10))
    flat = FlatParams(model, with_momentum=False)
    master = flat.param.float().clone()
    assert torch.equal(master, res[0]["init"])
    ps = [v.clone() for v in flat.views(master)]
    bufs = [torch.zeros_like(p) for p in ps]
    B = 32
    for step in range(nsteps):
        gs = [torch.zeros_like(p) for p in ps]
        Wq = [bf(p) for p in ps]
        for r in range(world):
            g = torch.Generator().manual_seed(1000 * step + r)
            x = bf(torch.randn(B, 784, generator=g))
            y = torch.randint(0, 10, (B,), generator=g)
            h0 = x
            h1 = bf(torch.relu(h0 @ Wq[0].t() + Wq[1]))
            h2 = bf(torch.relu(h1 @ Wq[2].t() + Wq[3]))
            logits = h2 @ Wq[4].t() + Wq[5]
            d3 = torch.softmax(logits, 1)
            d3[torch.arange(B), y] -= 1.0
            d3 = bf(d3 / B)
            d2 = bf((d3 @ Wq[4]) * (h2 > 0))
            d1 = bf((d2 @ Wq[2]) * (h1 > 0))
            for i, (d, h) in enumerate(((d1, h0), (d2, h1), (d3, h2))):
                gs[2 * i] += bf(d.t() @ h)
                gs[2 * i + 1] += bf(d.sum(0))
        for p, a, b in zip(ps, gs, bufs):
            b.mul_(0.5).add_(a / world)
            p.sub_(0.1 * b)
    errs = {}
    for s, p in zip(flat.slots, ps):
        init = res[0]["init"][s.offset:s.offset + s.numel].view(s.shape)
        delta_ref = p - init
        delta_got = got_master[s.offset:s.offset + s.numel].view(s.shape) - init
        errs[s.name] = float((delta_got - delta_ref).norm() / (delta_ref.norm() + 1e-12))
        # the bf16 replica is the fp32 master rounded once
        torch.testing.assert_close(res[0]["final"][s.offset:s.offset + s.numel].view(s.shape),
                                   bf(got_master[s.offset:s.offset + s.numel].view(s.shape)),
                                   rtol=0, atol=0)
    print("mlp update errors vs the rounding oracle:", {k: round(v, 4) for k, v in errs.items()})
    # What is left is summation order inside the GEMMs and the all-reduce and the bf16 rounding
    # flips it causes downstream (measured on 2 GPUs: 0.3 % for the last layer, 2.4-3.1 % for the
    # two below it, the in-switch NVLS reduction being the largest). A sign error shows as an
    # error above 1, a dropped rank as 1/world >= 12 %.
    for name, err in errs.items():
        assert err < 0.05, f"update mismatch {err:.4f} for {name}: {errs}"


@pytest.mark.parametrize("kind,bn", [("cta", 128), ("persistent", 256), ("persistent", 128)])
def test_gemm_reduce_scatter_peer(native_lib, tmp_path, kind, bn):
    """GEMM with the reduce-scatter epilogue + the bucket kernel == mean of the per-rank products."""
    world = _world()
    launch(_workers.gemm_rs, world, args=("b200", str(tmp_path), kind, bn), timeout=600)
    res = [torch.load(tmp_path / f"rs_{r}.pt") for r in range(world)]
    want = sum(r["local"].to(torch.bfloat16).float() for r in res) / world
    for r in range(world):
        assert torch.equal(res[r]["got"], res[0]["got"])
        torch.testing.assert_close(res[r]["got"], want, rtol=1.6e-2, atol=0.3)


@pytest.mark.parametrize("what", ["oneshot", "twoshot", "persist"])
def test_late_rank_is_named(native_lib, tmp_path, what):
    """SURVEY §5 failure detection (reference site: src/train_dist.py:146 runs with library
    defaults and would hang): a rank that stays away longer than the timeout surfaces on every
    other rank as an exception naming it, after about the timeout, and the group tears down."""
    world = _world()
    launch(_workers.late_rank, world, args=("b200", str(tmp_path), what), timeout=300)
    res = [torch.load(tmp_path / f"late_{r}.pt") for r in range(world)]
    late = world - 1
    for r in range(world - 1):
        assert res[r]["err"] is not None, f"rank {r} did not notice the missing rank"
        assert f"rank {late}" in res[r]["err"] or f"peer {late}" in res[r]["err"], res[r]["err"]
        assert "timed out" in res[r]["err"]
        assert 0.9 < res[r]["waited"] < 3.4, res[r]["waited"]      # gave up at the timeout, not at the sleep


def test_protocols_under_random_delays(native_lib, tmp_path):
    """Tagged-word and flag protocols with every rank arriving at a random time, 240 collectives
    in a row on the same staging buffers (double-buffer reuse, sequence wrap of the slot bit)."""
    world = _world()
    launch(_workers.protocol_stress, world, args=("b200", str(tmp_path), 240), timeout=600)
    res = [torch.load(tmp_path / f"stress_{r}.pt") for r in range(world)]
    for r in range(world):
        assert res[r]["iters"] == 240 and res[r]["bad"] == [], res[r]["bad"][:5]


def test_trainer_distributed_peer(native_lib, tmp_path):
    world = 2 if torch.cuda.device_count() >= 2 else 0
    if not world:
        pytest.skip("needs 2 GPUs")
    launch(_workers.trainer_dist, world, args=("b200", str(tmp_path)), timeout=600)
    res = [torch.load(tmp_path / f"tr_{r}.pt") for r in range(world)]
    assert torch.equal(res[0]["param"], res[1]["param"])            # replicas never diverge
    assert res[0]["test_losses"][-1] < res[0]["test_losses"][0]     # and the model learns
    # per-rank batch int(64/2) = 32 over a shard of 1036 samples: 33 batches per epoch
    assert res[0]["n_train_losses"] == 2 * 33
    assert (tmp_path / "results" / "model.pt").exists()             # rank 0 only


def test_nvls_allreduce_peer(native_lib, tmp_path):
    """NVLink SHARP: multimem.ld_reduce / multimem.st all-reduce vs the expected sums."""
    world = _world()
    launch(_workers.nvls_allreduce, world, args=("b200", str(tmp_path)), timeout=600)
    res = [torch.load(tmp_path / f"nvls_{r}.pt") for r in range(world)]
    if not res[0]["supported"]:
        pytest.skip("multicast (NVLS) is not offered by this driver / fabric")
    for name, dtype in (("f32", torch.float32), ("bf16", torch.bfloat16)):
        for n in (8, 4096 + 8, 1 << 20):
            xs = [torch.randn(n, generator=torch.Generator().manual_seed(100 + r)).to(dtype).float()
                  for r in range(world)]
            want = torch.stack(xs).sum(0)
            tol = dict(rtol=1e-5, atol=1e-5) if name == "f32" else dict(rtol=2e-2, atol=2e-2)
            for r in range(world):
                torch.testing.assert_close(res[r][f"{name}_{n}_sum"], want, **tol)
                torch.testing.assert_close(res[r][f"{name}_{n}_mean"], want / world, **tol)
                assert torch.equal(res[r][f"{name}_{n}_sum"], res[0][f"{name}_{n}_sum"])
