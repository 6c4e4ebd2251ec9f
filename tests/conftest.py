import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 GPUs (skipped when fewer are visible)")


@pytest.fixture(scope="session")
def native_lib():
    """Build (if needed) and load the native library; GPU tests must never run on a fallback."""
    from csed_514_project_distributed_training_using_pytorch_b200 import _native as N
    if not N.available():
        import build
        build.build()
    return N.lib()


@pytest.fixture(scope="session")
def cuda_device(native_lib):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device visible")
    return torch.device("cuda", 0)
