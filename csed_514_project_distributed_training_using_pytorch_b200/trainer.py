"""Training loops: the public entry point a user of the reference switches to.

Reference loops being reproduced:
* single process, src/train.py:69-109: `test(); for epoch in 1..n: train(epoch); test()`,
  a log line + loss-curve point + checkpoint every `log_interval` batches;
* distributed, src/train_dist.py:58-116: per epoch `sampler.set_epoch(i)`, the DDP step
  for every batch, then validation of the full test set on every rank and one line
  `Epoch=, train_loss=, val_loss=, accuracy=, time_elapsed=`.

Execution on a GPU (default): the dataset is uploaded to HBM once as uint8 (47 MB), the
epoch's sample order is uploaded once per epoch (240 KB), and every optimizer step is ONE
CUDA-graph replay of two kernels (ops/fused.py); the kernel gathers + normalises its own
batch from the step counter, so the hot loop has no host->device traffic and no host sync.
Losses are read back from a device ring once per log interval (K18). `feed="host"` keeps
the dataset in pinned host memory instead: the epoch's shuffled order is gathered on the host
into a pinned buffer and the native runner copies one batch up per step (the role of the
reference's DataLoader(pin_memory=True), src/train_dist.py:40-45).

Execution on CPU (`device="cpu"`, BASELINE config 1): the PyTorch oracle path with the same
sampler, logging and checkpoints; multi-process through gloo.
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from .models import Net
from .ops import net_ops
from .ops.fused import FusedNetEngine, X_F32, X_U8, X_U8_GATHER
from .ops.persist import PersistentNetEngine, persistent_enabled
from .ops.sgd import FlatSGD
from .utils import checkpoint as ckpt
from .utils.config import TrainConfig
from .utils.data import MNISTArrays, load_mnist
from .utils.flat import FlatParams
from .utils.metrics import MetricsLogger
from .utils.sampler import ShardSampler
from .utils import nvtx


class Trainer:
    def __init__(self, cfg: Optional[TrainConfig] = None, comm=None, model: Optional[Net] = None,
                 logger: Optional[MetricsLogger] = None, distributed_style: Optional[bool] = None):
        self.cfg = cfg = cfg or TrainConfig()
        self.comm = comm
        self.rank = comm.rank if comm is not None else 0
        self.world = comm.world_size if comm is not None else 1
        self.distributed_style = (self.world > 1 or comm is not None) \
            if distributed_style is None else distributed_style
        if comm is not None:
            # the per-rank batch is int(batch / world) of the REAL world (src/train_dist.py:134),
            # whatever the config object was built with
            cfg.world_size, cfg.rank = self.world, self.rank
        use_cuda = cfg.device != "cpu" and torch.cuda.is_available()
        if cfg.device != "cpu" and not use_cuda:
            raise RuntimeError("device='cuda' requested but no GPU is visible; pass device='cpu'")
        if use_cuda:
            self.device = comm.device if comm is not None and comm.device.type == "cuda" \
                else torch.device("cuda", torch.cuda.current_device())
        else:
            self.device = torch.device("cpu")
        torch.manual_seed(cfg.random_seed)               # src/train.py:21, src/train_dist.py:137
        self.model = (model or Net(conv2_impl=cfg.conv2_impl)).to(self.device)
        self.log = logger or MetricsLogger(rank=self.rank)
        self.engine: Optional[FusedNetEngine] = None
        self.ddp = None
        self.opt = None
        self.flat = None
        if self.device.type == "cuda" and cfg.fused:
            if comm is not None and self.world > 1:
                from .parallel.data_parallel import DataParallel
                self.ddp = DataParallel(self.model, comm)
                self.engine = self.ddp.fused_engine(cfg.learning_rate, cfg.momentum,
                                                    seed=cfg.random_seed, dropout=cfg.dropout)
            else:
                self.engine = PersistentNetEngine(self.model, cfg.learning_rate, cfg.momentum,
                                                  seed=cfg.random_seed, dropout=cfg.dropout)
            self.flat = self.engine.flat
        else:
            if comm is not None and self.world > 1:
                from .parallel.data_parallel import DataParallel
                self.ddp = DataParallel(self.model, comm)
                self.flat = self.ddp.flat
            else:
                alloc = comm.alloc if comm is not None else None
                self.flat = FlatParams(self.model, alloc=alloc)
            self.opt = FlatSGD(self.flat.param, self.flat.grad, cfg.learning_rate, cfg.momentum,
                               momentum_buffer=self.flat.momentum)
        self.train_set: Optional[MNISTArrays] = None
        self.test_set: Optional[MNISTArrays] = None
        self._dev_train = self._dev_test = None
        self._graphs = {}
        self._perm_dev = None
        self.examples_seen = 0
        self.native_runner = True   # host-streaming steps go through native code, not Python
        # persistent step kernel (csrc/net_persist.cu) for whole runs of steps; the two-kernel
        # graphs remain for B200DP_PERSIST=0 and for cuda_graph=False debugging
        self.persistent = persistent_enabled() and isinstance(self.engine, PersistentNetEngine)
        if cfg.resume:
            # momentum lives in optimizer.pth beside model.pth (src/train.py:84-85)
            self.resume(cfg.resume, os.path.join(os.path.dirname(cfg.resume) or ".", "optimizer.pth"))

    # ------------------------------------------------------------------ data
    def set_data(self, train: MNISTArrays, test: Optional[MNISTArrays] = None):
        self.train_set, self.test_set = train, test
        self._dev_train = self._dev_test = None
        self._graphs = {}

    def load_data(self):
        c = self.cfg
        self.set_data(load_mnist(c.data_dir, True, c.train_size, seed=c.random_seed),
                      load_mnist(c.data_dir, False, c.test_size, seed=c.random_seed))

    def _resident(self, which: str) -> MNISTArrays:
        if which == "train":
            if self._dev_train is None:
                self._dev_train = self.train_set.to(self.device)
            return self._dev_train
        if self._dev_test is None:
            self._dev_test = self.test_set.to(self.device)
        return self._dev_test

    # ------------------------------------------------------------ checkpoints
    def optimizer_state(self) -> dict:
        c = self.cfg
        return ckpt.sgd_state_dict(self.flat.views(self.flat.momentum), c.learning_rate, c.momentum)

    def save(self, model_path: Optional[str] = None, optim_path: Optional[str] = None):
        """results/model.pth + results/optimizer.pth (src/train.py:84-85), rank 0 only."""
        c = self.cfg
        model_path = model_path or os.path.join(c.results_dir, "model.pth")
        optim_path = optim_path or os.path.join(c.results_dir, "optimizer.pth")
        ckpt.save_checkpoint(model_path, self.model, self.rank)
        ckpt.save_optimizer(optim_path, self.optimizer_state(), self.rank)

    def resume(self, model_path: str, optim_path: Optional[str] = None):
        """Load + re-broadcast (the reference has no load path; SURVEY §5 checkpoint row)."""
        ckpt.load_checkpoint(model_path, self.model)
        if optim_path and os.path.exists(optim_path):
            ckpt.load_momentum(optim_path, self.flat.views(self.flat.momentum))
        if self.comm is not None and self.world > 1:
            self.comm.broadcast(self.flat.param, root=0)
            self.comm.broadcast(self.flat.momentum, root=0)

    # ------------------------------------------------------------------ train
    def _sampler(self) -> ShardSampler:
        c = self.cfg
        if self.distributed_style:
            return ShardSampler(len(self.train_set), self.world, self.rank, shuffle=True,
                                seed=c.sampler_seed)          # src/train_dist.py:33-37
        # single process: DataLoader(shuffle=True) (src/train.py:25-32)
        return ShardSampler(len(self.train_set), 1, 0, shuffle=True, seed=c.random_seed)

    def train_epoch(self, epoch: int) -> float:
        """One epoch over this rank's shard. Returns the reference's `epoch_loss`
        (sum over batches of batch-mean-loss / batch-size, src/train_dist.py:86)."""
        self.model.train()
        sampler = self._sampler()
        sampler.set_epoch(epoch if self.distributed_style else epoch - 1)
        with nvtx.range(f"train_epoch {epoch}"):
            if self.device.type == "cuda" and self.engine is not None:
                if self.cfg.feed == "host":
                    return self._train_epoch_host_feed(epoch, sampler)
                return self._train_epoch_fused(epoch, sampler)
            return self._train_epoch_autograd(epoch, sampler)

    def _train_epoch_host_feed(self, epoch: int, sampler: ShardSampler) -> float:
        """Dataset stays on the host: gather this epoch's order into pinned memory once, then
        stream it through the native runner (H2D per step, copy/compute overlap)."""
        c = self.cfg
        B = c.rank_batch() if self.distributed_style else c.batch_size_train
        idx = sampler.indices()
        n = idx.numel()
        if getattr(self, "_host_x", None) is None or self._host_x.shape[0] != n:
            self._host_x = torch.empty(n, 784, dtype=torch.uint8).pin_memory()
            self._host_y = torch.empty(n, dtype=torch.int64).pin_memory()
        torch.index_select(self.train_set.images, 0, idx, out=self._host_x)
        torch.index_select(self.train_set.labels, 0, idx, out=self._host_y)
        full, tail = n // B, n % B
        n_dataset = len(self.train_set)
        batches = sampler.batches(B)
        losses = []
        if getattr(self, "_epoch_loss_pinned", None) is None or self._epoch_loss_pinned.numel() < full + 1:
            self._epoch_loss_pinned = torch.empty(full + 1, dtype=torch.float32).pin_memory()
        if full:
            losses.append(self.train_steps_from_host(self._host_x, self._host_y, full, B,
                                                     loss_out=self._epoch_loss_pinned[:full]))
        if tail:
            losses.append(self.train_steps_from_host(self._host_x[full * B:], self._host_y[full * B:],
                                                     1, tail,
                                                     loss_out=self._epoch_loss_pinned[full:full + 1]))
        torch.cuda.synchronize(self.device)
        if self.persistent:
            self.engine.check_errors()
        elif self.comm is not None:
            self.comm.check_errors()
        epoch_loss = 0.0
        flat = torch.cat(losses).tolist() if losses else []
        for bi, l in enumerate(flat):
            blen = batches[bi][1]
            epoch_loss += l / blen
            if self.distributed_style:
                self.log.train_losses.append(l)
                self.log.train_counter.append(bi * 64 + epoch * n_dataset)
            elif bi % c.log_interval == 0:
                self.log.train_log(epoch, bi, blen, n_dataset, len(batches), l,
                                   bi * 64 + (epoch - 1) * n_dataset)
        if not self.distributed_style and c.checkpoint_every_log:
            self.save()
        self.examples_seen += n
        return epoch_loss

    UNROLL = 16   # steps per captured graph in the resident path (PDL spans step boundaries)

    def _graph_for(self, B: int, data: MNISTArrays, unroll: int = 1):
        g = self._graphs.get((B, unroll))
        if g is None:
            g = self.engine.capture(data.images, data.labels, X_U8_GATHER, gidx=self._perm_dev,
                                    batch=B, idx_stride=self._stride, unroll=unroll)
            self._graphs[(B, unroll)] = g
        return g

    def _train_epoch_fused(self, epoch: int, sampler: ShardSampler) -> float:
        c, eng = self.cfg, self.engine
        B = c.rank_batch() if self.distributed_style else c.batch_size_train
        data = self._resident("train")
        idx = sampler.indices().to(torch.int32)
        n = idx.numel()
        if self._perm_dev is None or self._perm_dev.numel() != n:
            self._perm_dev = torch.empty(n, dtype=torch.int32, device=self.device)
            self._graphs = {}
        self._perm_dev.copy_(idx, non_blocking=True)
        self._stride = B
        batches = sampler.batches(B)
        n_batches = len(batches)
        full = n // B
        tail = n - full * B
        if not self.persistent:
            eng.begin_epoch()
        epoch_open = False
        use_graph = c.cuda_graph and not self.persistent
        g_full = self._graph_for(B, data) if (use_graph and full > 0) else None
        g_multi = self._graph_for(B, data, self.UNROLL) if (use_graph and full >= self.UNROLL) else None
        epoch_loss = 0.0
        li = c.log_interval
        n_dataset = len(self.train_set)
        pending = 0   # steps whose losses are still only in the device ring

        def flush(upto_batch: int):
            """Read the ring once; emit the reference's per-log-interval lines."""
            nonlocal pending, epoch_loss
            if pending == 0:
                return
            losses = eng.last_losses(pending).tolist()
            first = upto_batch - pending
            for k, l in enumerate(losses):
                bi = first + k
                blen = batches[bi][1]
                epoch_loss += l / blen
                if self.distributed_style:
                    # src/train_dist.py:88-89 records every batch
                    self.log.train_losses.append(l)
                    self.log.train_counter.append(bi * 64 + epoch * n_dataset)
                elif bi % li == 0:
                    self.log.train_log(epoch, bi, blen, n_dataset, n_batches, l,
                                       bi * 64 + (epoch - 1) * n_dataset)
            if self.distributed_style and losses:
                self.log.progress(upto_batch, n_batches, losses[-1], final=upto_batch >= n_batches)
            pending = 0

        chunk = min(eng.ring - self.UNROLL, max(li, 256))
        bi = 0
        while bi < full:
            if self.persistent:
                # one launch runs every step up to the next loss flush
                adv = min(full - bi, chunk - pending)
                eng.run_steps(data.images, data.labels, X_U8_GATHER, gidx=self._perm_dev, batch=B,
                              nsteps=adv, idx_stride=B, begin_epoch=not epoch_open)
                epoch_open = True
            elif g_multi is not None and full - bi >= self.UNROLL:
                g_multi.replay()
                adv = self.UNROLL
            elif g_full is not None:
                g_full.replay()
                adv = 1
            else:
                eng.step(data.images, data.labels, X_U8_GATHER, gidx=self._perm_dev, batch=B,
                         idx_stride=B)
                adv = 1
            bi += adv
            pending += adv
            if pending >= chunk:
                flush(bi)
                if not self.distributed_style and c.checkpoint_every_log:
                    self.save()
        if tail:
            # the short last batch (60000 % 64 = 32): own launch, indices at an explicit offset
            if self.persistent:
                eng.run_steps(data.images, data.labels, X_U8_GATHER, gidx=self._perm_dev[full * B:],
                              batch=tail, nsteps=1)
            else:
                eng.step(data.images, data.labels, X_U8_GATHER, gidx=self._perm_dev[full * B:],
                         batch=tail)
            pending += 1
        flush(n_batches)
        if not self.distributed_style and c.checkpoint_every_log:
            self.save()
        self.examples_seen += n
        if self.persistent:
            eng.check_errors()
        elif self.comm is not None:
            self.comm.check_errors()
        return epoch_loss

    def _train_epoch_autograd(self, epoch: int, sampler: ShardSampler) -> float:
        """Layered path: module forward, nll_loss, backward, (DDP all-reduce), SGD.
        On CPU this is the reference's own arithmetic (BASELINE config 1)."""
        c = self.cfg
        B = c.rank_batch() if self.distributed_style else c.batch_size_train
        data = self._resident("train")
        idx = sampler.indices().to(self.device)
        n_dataset = len(self.train_set)
        batches = sampler.batches(B)
        fwd = self.ddp if self.ddp is not None else self.model
        epoch_loss = 0.0
        for bi, (start, blen) in enumerate(batches):
            sel = idx[start:start + blen]
            x = net_ops.gather_normalize(data.images, sel.to(torch.int32))
            y = data.labels[sel]
            self.flat.grad.zero_()
            self.flat.rebind_grads()
            out = fwd(x)
            loss = net_ops.nll_loss(out, y)
            loss.backward()
            if self.ddp is not None:
                # launches any bucket a parameter without gradient kept from filling
                self.ddp.finish_backward()
            if self.ddp is not None and self.device.type == "cuda":
                torch.cuda.current_stream().wait_stream(self.ddp.reducer.stream)
            self.opt.step()
            if self.distributed_style:
                l = loss.item()                           # src/train_dist.py:85
                epoch_loss += l / blen
                self.log.train_losses.append(l)
                self.log.train_counter.append(bi * 64 + epoch * n_dataset)
            elif bi % c.log_interval == 0:
                l = loss.item()
                self.log.train_log(epoch, bi, blen, n_dataset, len(batches), l,
                                   bi * 64 + (epoch - 1) * n_dataset)
                if c.checkpoint_every_log:
                    self.save()
        self.examples_seen += idx.numel()
        return epoch_loss

    # ------------------------------------------------------------------- eval
    @torch.no_grad()
    def evaluate(self):
        """Full test set on this rank (every rank evaluates everything, as in
        src/train_dist.py:21-24,92-104). Returns (nll_sum, correct, n, sum_of_batch_means)."""
        self.model.eval()
        c = self.cfg
        nvtx.push("evaluate")
        data = self._resident("test")
        n = len(data)
        Bt = c.batch_size_test
        if self.device.type == "cuda" and self.engine is not None:
            tot = torch.zeros(3, dtype=torch.float32, device=self.device)
            for s in range(0, n, Bt):
                b = min(Bt, n - s)
                out = self.engine.evaluate_batch(data.images[s:s + b], data.labels[s:s + b], X_U8,
                                                 accumulate=False, batch=b)
                tot[:2] += out
                tot[2] += out[0] / b
            vals = tot.cpu().tolist()
            nll_sum, correct, sum_means = vals[0], int(round(vals[1])), vals[2]
        else:
            nll_sum, correct, sum_means = 0.0, 0, 0.0
            for s in range(0, n, Bt):
                x = net_ops.gather_normalize(data.images[s:s + Bt])
                y = data.labels[s:s + Bt]
                out = self.model(x)
                l = net_ops.nll_loss(out, y, reduction="sum").item()
                nll_sum += l
                sum_means += l / x.shape[0]
                correct += int((out.argmax(1) == y).sum().item())
        self.model.train()
        nvtx.pop()
        return nll_sum, correct, n, sum_means

    # -------------------------------------------------------------------- fit
    def fit(self, epochs: Optional[int] = None):
        """Run the reference's whole script body; returns the metrics logger."""
        c = self.cfg
        epochs = epochs or c.epochs
        if self.train_set is None:
            self.load_data()
        n_train = len(self.train_set)
        if self.distributed_style:
            self.log.test_counter = [i * n_train for i in range(epochs)]   # src/train_dist.py:153
            for i in range(epochs):
                train_loss = self.train_epoch(i)
                nll_sum, correct, n, sum_means = self.evaluate()
                # src/train_dist.py:101,108: sum of batch-MEAN losses / dataset size (kept as is)
                val_loss = sum_means / n
                self.log.test_losses.append(val_loss)
                self.log.epoch_log(i, train_loss, val_loss, 100.0 * correct / n)
            if self.rank == 0:
                ckpt.save_checkpoint("model.pt" if not c.results_dir else
                                     os.path.join(c.results_dir, "model.pt"), self.model, 0)
            self.log.save_curves(c.images_dir, "train_test_curve_dist")
        else:
            self.log.test_counter = [i * n_train for i in range(epochs + 1)]  # src/train.py:67
            if self.test_set is not None and len(self.test_set) >= 6:         # src/train.py:43-57
                self.log.save_sample_digits(c.images_dir, self.test_set.images[:6].cpu().numpy(),
                                            self.test_set.labels[:6].tolist())
            nll_sum, correct, n, _ = self.evaluate()                           # src/train.py:106
            self.log.test_log(nll_sum / n, correct, n)
            for epoch in range(1, epochs + 1):
                self.train_epoch(epoch)
                nll_sum, correct, n, _ = self.evaluate()
                self.log.test_log(nll_sum / n, correct, n)
            self.log.save_curves(c.images_dir, "train_test_curve")
        return self.log

    # ------------------------------------------------- host-streaming (e2e) path
    def train_steps_from_host(self, hx: torch.Tensor, hy: torch.Tensor, steps: int, batch: int,
                              loss_out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Feed `steps` batches from PINNED host memory: per step one H2D copy of the
        inputs (images fp32 [batch,784] or uint8, labels int64), one fused step, one D2H
        copy of that step's loss into pinned host memory. Fully asynchronous; the caller
        synchronises. hx: [steps*batch, 784], hy: [steps*batch]."""
        eng = self.engine
        if eng is None:
            raise RuntimeError("train_steps_from_host needs the fused CUDA engine")
        if loss_out is None:
            # one pinned buffer per trainer, grown on demand: pinning memory is a blocking
            # driver call that must not sit in the step loop. The returned view is
            # overwritten by the next call; pass loss_out to keep several results.
            buf = getattr(self, "_loss_pinned", None)
            if buf is None or buf.numel() < steps:
                buf = self._loss_pinned = torch.empty(max(steps, 1024), dtype=torch.float32).pin_memory()
            loss_out = buf[:steps]
        if self.persistent and self.native_runner:
            # ONE persistent launch for the whole call; the copy stream feeds it step by step
            eng.feeder(batch, hx.dtype).run(hx, hy, steps, loss_out)
            return loss_out
        if self.cfg.cuda_graph and self.native_runner:
            # native loop (csrc/runner.cu): per-step copies, 16-step kernel graphs, no Python
            key = ("runner", batch, hx.dtype)
            r = self._graphs.get(key)
            if r is None:
                from .ops.runner import NativeRunner
                r = self._graphs[key] = NativeRunner(eng, batch, hx.dtype)
            r.run(hx, hy, steps, loss_out)
            return loss_out
        mode = X_U8 if hx.dtype == torch.uint8 else X_F32
        key = ("host", batch, hx.dtype)
        st = self._graphs.get(key)
        if st is None:
            sx = torch.empty(batch, 784, dtype=hx.dtype, device=self.device)
            sy = torch.empty(batch, dtype=torch.int64, device=self.device)
            sx.copy_(hx[:batch]); sy.copy_(hy[:batch])
            g = eng.capture(sx, sy, mode, batch=batch) if self.cfg.cuda_graph else None
            st = self._graphs[key] = (sx, sy, g)
        sx, sy, g = st
        last = eng.last_loss
        for i in range(steps):
            sx.copy_(hx[i * batch:(i + 1) * batch], non_blocking=True)
            sy.copy_(hy[i * batch:(i + 1) * batch], non_blocking=True)
            if g is not None:
                g.replay()
            else:
                eng.step(sx, sy, mode, batch=batch)
            loss_out[i:i + 1].copy_(last, non_blocking=True)
        return loss_out
