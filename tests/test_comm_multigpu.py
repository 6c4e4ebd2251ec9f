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


@pytest.mark.parametrize("use_graph", [False, True])
def test_net_ddp_fused_peer(native_lib, tmp_path, use_graph):
    from tests._ddp_checks import check_net_ddp
    world = _world()
    launch(_workers.net_ddp_fused, world, args=("b200", str(tmp_path), use_graph), timeout=600)
    res = [torch.load(tmp_path / f"net_{r}.pt") for r in range(world)]
    check_net_ddp(res, world, check_grads=not use_graph)
