"""The command-line entry points on CPU (reference: python train.py / train_dist.py / run1.py)."""
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
PKG = "csed_514_project_distributed_training_using_pytorch_b200"


def _run(args, cwd, timeout=300):
    env = dict(os.environ, PYTHONPATH=str(ROOT))
    return subprocess.run([sys.executable, "-m"] + args, cwd=cwd, env=env, capture_output=True,
                          text=True, timeout=timeout)


def test_train_cli_cpu(tmp_path):
    r = _run([f"{PKG}.cli.train", "--device", "cpu", "--epochs", "1", "--train_size", "640",
              "--test_size", "200", "--batch_size_test", "100", "--metrics_dir", str(tmp_path / "m")],
             cwd=tmp_path)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Train Epoch: 1 [0/640 (0%)]" in r.stdout and "Test set: Avg. loss:" in r.stdout
    assert (tmp_path / "results" / "model.pth").exists()           # src/train.py:84
    assert (tmp_path / "results" / "optimizer.pth").exists()       # src/train.py:85
    assert (tmp_path / "images" / "train_test_curve.csv").exists()
    assert (tmp_path / "m" / "metrics.jsonl").exists()


def test_p2p_demo_cli_gloo(tmp_path):
    r = _run([f"{PKG}.cli.p2p_demo", "--spawn", "--backend", "gloo", "--master_port", "29641"],
             cwd=tmp_path)
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.count("has data") == 2 and "tensor(1.)" in r.stdout   # src/run1.py:17


def test_launch_cli_distributed_cpu(tmp_path):
    r = _run([f"{PKG}.cli.launch", "--nproc", "2", "--", "--device", "cpu", "--epochs", "1",
              "--train_size", "512", "--test_size", "200", "--batch_size_test", "100"],
             cwd=tmp_path)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Epoch=0, train_loss=" in r.stdout and "accuracy=" in r.stdout   # src/train_dist.py:113
    assert (tmp_path / "results" / "model.pt").exists()                     # src/train_dist.py:164
