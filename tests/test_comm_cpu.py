"""Process-group, launcher and collective *logic* on gloo/CPU with world_size 2
(the no-GPU CI mode)."""
import torch

from csed_514_project_distributed_training_using_pytorch_b200.parallel import launch
from tests import _workers


def test_p2p_demo_gloo(tmp_path):
    launch(_workers.p2p_demo, 2, args=("gloo", str(tmp_path)), timeout=120)
    r0 = torch.load(tmp_path / "p2p_0.pt")
    r1 = torch.load(tmp_path / "p2p_1.pt")
    assert r0[0].item() == 1.0 and r1[0].item() == 1.0  # "Rank r has data 1." (src/run1.py:17)


def test_collectives_gloo(tmp_path):
    world = 2
    launch(_workers.collectives, world, args=("gloo", str(tmp_path)), timeout=180)
    res = [torch.load(tmp_path / f"coll_{r}.pt") for r in range(world)]
    _check_collectives(res, world)


def _check_collectives(res, world, bf16=False):
    for r in range(world):
        assert torch.equal(res[r]["bc_arena"], torch.arange(4000, dtype=torch.float32))
        assert torch.equal(res[r]["bc_plain"], torch.full((37,), float(world - 1)))
        assert res[r]["ring_ok"]
    assert res[1]["oneway_ok"]
    names = ["f32"] + (["bf16"] if bf16 else [])
    for name in names:
        dtype = torch.float32 if name == "f32" else torch.bfloat16
        for algo in ("oneshot", "twoshot"):
            for n in (8, 4096 + 8, 300000):
                xs = [torch.randn(n, generator=torch.Generator().manual_seed(100 + r)).to(dtype).float()
                      for r in range(world)]
                want = torch.stack(xs).sum(0)
                tol = dict(rtol=1e-5, atol=1e-5) if name == "f32" else dict(rtol=2e-2, atol=2e-2)
                for r in range(world):
                    torch.testing.assert_close(res[r][f"ar_{name}_{algo}_{n}"], want / world, **tol)
                    torch.testing.assert_close(res[r][f"ar2_{name}_{algo}_{n}"], want, **tol)
                    # every rank holds bit-identical results (fixed rank-order sums)
                    assert torch.equal(res[r][f"ar_{name}_{algo}_{n}"], res[0][f"ar_{name}_{algo}_{n}"])


def test_ddp_generic_gloo(tmp_path):
    from tests._ddp_checks import check_ddp_generic
    for fuse in (False, True):
        d = tmp_path / f"fuse{int(fuse)}"
        d.mkdir()
        launch(_workers.ddp_generic, 2, args=("gloo", str(d), fuse, "auto"), timeout=180)
        res = [torch.load(d / f"ddp_{r}.pt") for r in range(2)]
        check_ddp_generic(res, 2)


def test_trainer_distributed_gloo(tmp_path):
    launch(_workers.trainer_dist, 2, args=("gloo", str(tmp_path)), timeout=300)
    res = [torch.load(tmp_path / f"tr_{r}.pt") for r in range(2)]
    assert torch.equal(res[0]["param"], res[1]["param"])
    assert res[0]["test_losses"][-1] < res[0]["test_losses"][0]
    assert res[0]["n_train_losses"] == 2 * 33
    assert (tmp_path / "results" / "model.pt").exists()
