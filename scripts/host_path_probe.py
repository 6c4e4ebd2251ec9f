#!/usr/bin/env python
"""Host-side cost of one Trainer.train_steps_from_host call, piece by piece (perf_counter, us)."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from csed_514_project_distributed_training_using_pytorch_b200 import _native as N
from csed_514_project_distributed_training_using_pytorch_b200.trainer import Trainer
from csed_514_project_distributed_training_using_pytorch_b200.utils.config import TrainConfig
from csed_514_project_distributed_training_using_pytorch_b200.ops import persist

B, K = 64, 20
tr = Trainer(TrainConfig(batch_size_train=B, per_rank_batch=B), distributed_style=True)
hx = torch.randn(K * B, 784).pin_memory()
hy = torch.randint(0, 10, (K * B,)).pin_memory()
hl = torch.empty(64).pin_memory()
for _ in range(3):
    tr.train_steps_from_host(hx, hy, K, B, loss_out=hl[:K])
torch.cuda.synchronize()


def med(fn, n=200, sync=False):
    ts = []
    for _ in range(n):
        if sync:
            torch.cuda.synchronize()
        t0 = time.perf_counter(); fn(); ts.append((time.perf_counter() - t0) * 1e6)
    return round(sorted(ts)[n // 2], 2)


e = torch.cuda.Event(enable_timing=True)
print("event.record()", med(e.record))
print("N.stream()", med(N.stream))
print("loss slice", med(lambda: hl[:K]))
print("full call (async, incl. C)", med(lambda: tr.train_steps_from_host(hx, hy, K, B, loss_out=hl[:K]), n=50, sync=True))
real = persist.HostFeeder.run
feeder = tr.engine.feeder(B, hx.dtype)
h = feeder._h
print("C call alone", med(lambda: N.lib().b200_feeder_run(h, hx.data_ptr(), hy.data_ptr(), K, hl.data_ptr(), N.stream()), n=50, sync=True))
persist.HostFeeder.run = lambda self, *a, **k: None
print("python path with the C call stubbed", med(lambda: tr.train_steps_from_host(hx, hy, K, B, loss_out=hl[:K])))
persist.HostFeeder.run = real
lib = N.lib()
fn = lib.b200_feeder_run
saved = fn
class Stub:
    def b200_feeder_run(self, *a): return 0
    def __getattr__(self, k): return getattr(lib, k)
N_lib = N.lib
N.lib = lambda: Stub()
print("python path incl. HostFeeder.run checks, C stubbed", med(lambda: tr.train_steps_from_host(hx, hy, K, B, loss_out=hl[:K])))
N.lib = N_lib
torch.cuda.synchronize()
