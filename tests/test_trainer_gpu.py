"""Trainer on a real GPU: the reference's single-process script body (src/train.py) with the
fused engine, device-resident data, CUDA graphs, checkpoints, evaluation and resume."""
import os

import pytest
import torch

from csed_514_project_distributed_training_using_pytorch_b200.trainer import Trainer
from csed_514_project_distributed_training_using_pytorch_b200.utils.config import TrainConfig
from csed_514_project_distributed_training_using_pytorch_b200.utils.data import synthetic_mnist
from csed_514_project_distributed_training_using_pytorch_b200.models import Net, oracle

pytestmark = pytest.mark.gpu


def test_trainer_fit_gpu_learns_and_matches_cpu_eval(cuda_device, tmp_path, capsys):
    cfg = TrainConfig(epochs=2, train_size=4000 + 37, test_size=1500, batch_size_test=1000,
                      learning_rate=0.05, results_dir=str(tmp_path / "results"),
                      images_dir=str(tmp_path / "images"))
    tr = Trainer(cfg, distributed_style=False)
    tr.set_data(synthetic_mnist(cfg.train_size, seed=0), synthetic_mnist(cfg.test_size, seed=1))
    log = tr.fit()
    out = capsys.readouterr().out
    assert "Train Epoch: 1 [0/4037 (0%)]\tLoss:" in out and "Test set: Avg. loss:" in out
    assert len(log.test_losses) == 3
    assert log.test_losses[-1] < 0.5 * log.test_losses[0]          # it learns the synthetic digits
    # 4037 = 63 full batches of 64 + a tail of 5: every batch was logged exactly once
    assert tr.engine.steps_done() == 2 * 64
    assert os.path.exists(tmp_path / "results" / "model.pth")
    assert os.path.exists(tmp_path / "images" / "train_test_curve.csv")
    # GPU evaluation == PyTorch fp32 evaluation of the same weights on the CPU
    nll, correct, n, _ = tr.evaluate()
    sd = {k: v.detach().cpu() for k, v in tr.model.state_dict().items()}
    ref = Net(); ref.load_state_dict(sd); ref.eval()
    x = oracle.normalize_u8(tr.test_set.images)
    with torch.no_grad():
        lp = ref(x)
    want_nll = torch.nn.functional.nll_loss(lp, tr.test_set.labels, reduction="sum").item()
    want_correct = int((lp.argmax(1) == tr.test_set.labels).sum())
    assert abs(nll - want_nll) / max(1.0, abs(want_nll)) < 1e-3
    assert abs(correct - want_correct) <= 1
    # resume restores the state (weights + momentum) into the flat buffers
    tr.save()
    tr2 = Trainer(TrainConfig(resume=str(tmp_path / "results" / "model.pth")), distributed_style=False)
    tr2.resume(str(tmp_path / "results" / "model.pth"), str(tmp_path / "results" / "optimizer.pth"))
    assert torch.equal(tr2.flat.param, tr.flat.param)
    for a, b in zip(tr2.flat.views(tr2.flat.momentum), tr.flat.views(tr.flat.momentum)):
        assert torch.equal(a, b)


def test_trainer_layered_path_matches_fused_first_steps(cuda_device):
    """fused=False runs the per-layer autograd kernels + FlatSGD; without dropout it follows the
    same trajectory as the fused engine (different kernels, same math)."""
    data = synthetic_mnist(640, seed=3)
    res = []
    for fused in (True, False):
        cfg = TrainConfig(epochs=1, train_size=640, test_size=100, dropout=False, fused=fused,
                          checkpoint_every_log=False, learning_rate=0.02)
        tr = Trainer(cfg, distributed_style=False)
        if not fused:
            tr.model.eval()   # layered path: dropout off the same way
        tr.set_data(data, synthetic_mnist(100, seed=4))
        if fused:
            tr.train_epoch(1)
        else:
            tr.model.train = lambda mode=True: tr.model   # keep eval-mode dropout for parity
            tr.train_epoch(1)
        torch.cuda.synchronize()
        res.append(tr.flat.param.clone())
    torch.testing.assert_close(res[0], res[1], rtol=2e-3, atol=2e-4)


def test_trainer_host_feed_matches_resident(cuda_device):
    """feed="host" (pinned dataset, native runner, H2D per step) follows exactly the same
    trajectory as the device-resident feed: same kernels, same batches, same dropout keys."""
    data = synthetic_mnist(64 * 20 + 9, seed=5)
    res = []
    for feed in ("resident", "host"):
        cfg = TrainConfig(epochs=1, train_size=len(data), test_size=100, feed=feed,
                          checkpoint_every_log=False, learning_rate=0.02)
        tr = Trainer(cfg, distributed_style=False)
        tr.set_data(data, synthetic_mnist(100, seed=6))
        tr.log.quiet = True
        tr.train_epoch(1)
        torch.cuda.synchronize()
        res.append((tr.flat.param.clone(), list(tr.log.train_losses)))
    assert torch.equal(res[0][0], res[1][0])
    assert res[0][1] == res[1][1]
