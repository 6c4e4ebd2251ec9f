"""Logging / curves (SURVEY §5 "Metrics / logging"): the reference's printed fields
(src/train.py:77-80,100-104; src/train_dist.py:113-114) plus JSONL and CSV output.
matplotlib is not on the image, so curves are written as CSV/JSON and a PNG is rendered
only if matplotlib imports (reference: src/train.py:111-117, src/train_dist.py:49-56).
"""
from __future__ import annotations

import csv
import json
import os
import time
from typing import List, Optional


class MetricsLogger:
    def __init__(self, out_dir: str = "", rank: int = 0, quiet: bool = False):
        self.rank, self.quiet = rank, quiet
        self.out_dir = out_dir
        self.t0 = time.time()
        self.train_losses: List[float] = []
        self.train_counter: List[int] = []
        self.test_losses: List[float] = []
        self.test_counter: List[int] = []
        self._jsonl = None
        if out_dir and rank == 0:
            os.makedirs(out_dir, exist_ok=True)
            self._jsonl = open(os.path.join(out_dir, "metrics.jsonl"), "a")

    def elapsed(self) -> float:
        return time.time() - self.t0

    def _emit(self, rec: dict):
        if self._jsonl is not None:
            rec = dict(rec, t=round(self.elapsed(), 4))
            self._jsonl.write(json.dumps(rec) + "\n")
            self._jsonl.flush()

    def say(self, line: str):
        if self.rank == 0 and not self.quiet:
            print(line, flush=True)

    def train_log(self, epoch: int, batch_idx: int, batch_len: int, n_dataset: int,
                  n_batches: int, loss: float, counter: int):
        """src/train.py:78-83 line and curve point."""
        self.say("Train Epoch: {} [{}/{} ({:.0f}%)]\tLoss: {:.6f}".format(
            epoch, batch_idx * batch_len, n_dataset, 100. * batch_idx / max(1, n_batches), loss))
        self.train_losses.append(loss)
        self.train_counter.append(counter)
        self._emit({"kind": "train", "epoch": epoch, "batch": batch_idx, "loss": loss})

    def test_log(self, test_loss: float, correct: int, n: int, counter: Optional[int] = None):
        """src/train.py:100-104 summary."""
        self.say("\nTest set: Avg. loss: {:.4f}, Accuracy: {}/{} ({:.0f}%), time_elapsed={:.4f}\n"
                 .format(test_loss, correct, n, 100. * correct / max(1, n), self.elapsed()))
        self.test_losses.append(test_loss)
        if counter is not None:
            self.test_counter.append(counter)
        self._emit({"kind": "test", "loss": test_loss, "correct": correct, "n": n})

    def epoch_log(self, epoch: int, train_loss: float, val_loss: float, accuracy: float):
        """src/train_dist.py:113-114 per-epoch line."""
        self.say(f"Epoch={epoch}, train_loss={train_loss:.4f}, val_loss={val_loss:.4f}, "
                 f"accuracy={accuracy:.2f}, time_elapsed={self.elapsed():.4f}")
        self._emit({"kind": "epoch", "epoch": epoch, "train_loss": train_loss,
                    "val_loss": val_loss, "accuracy": accuracy})

    def progress(self, done: int, total: int, loss: float, final: bool = False):
        """The reference's tqdm bar text, `training batch_loss=...` (src/train_dist.py:76,87).
        Losses reach the host once per flush of the device ring, so the line advances per
        flush, not per batch. Written to stderr like tqdm; silent when quiet or rank != 0."""
        if self.rank != 0 or self.quiet:
            return
        import sys
        pct = 100.0 * done / max(1, total)
        rate = done / max(1e-9, self.elapsed())
        sys.stderr.write(f"\rtraining batch_loss={loss:.4f}: {pct:3.0f}%| {done}/{total} "
                         f"[{self.elapsed():.1f}s, {rate:.1f}it/s]" + ("\n" if final else ""))
        sys.stderr.flush()

    def save_sample_digits(self, images_dir: str, images_u8, labels, name: str = "train_images"):
        """The reference's 2x3 figure of six test digits with their ground truth
        (src/train.py:43-57). Without matplotlib the same content is written as
        `<name>.npy` (uint8 [6,28,28]), `<name>_labels.csv` and a 2x3 montage `<name>.pgm`
        (plain binary PGM, viewable anywhere); the PNG with titles is added when matplotlib
        imports."""
        if self.rank != 0:
            return None
        import numpy as np
        os.makedirs(images_dir, exist_ok=True)
        imgs = np.asarray(images_u8[:6], dtype=np.uint8).reshape(-1, 28, 28)
        labs = [int(x) for x in list(labels[:6])]
        np.save(os.path.join(images_dir, name + ".npy"), imgs)
        with open(os.path.join(images_dir, name + "_labels.csv"), "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(["panel", "ground_truth"])
            for i, l in enumerate(labs):
                w.writerow([i, l])
        n = imgs.shape[0]
        rows = (n + 2) // 3
        canvas = np.zeros((rows * 30 + 2, 3 * 30 + 2), dtype=np.uint8)
        for i in range(n):
            r, c = divmod(i, 3)
            canvas[2 + r * 30:2 + r * 30 + 28, 2 + c * 30:2 + c * 30 + 28] = imgs[i]
        path = os.path.join(images_dir, name + ".pgm")
        with open(path, "wb") as f:
            f.write(b"P5\n%d %d\n255\n" % (canvas.shape[1], canvas.shape[0]))
            f.write(canvas.tobytes())
        try:
            import matplotlib
            matplotlib.use("Agg")
            import matplotlib.pyplot as plt
            fig = plt.figure()
            for i in range(n):
                plt.subplot(2, 3, i + 1)
                plt.tight_layout()
                plt.imshow(imgs[i], cmap="gray", interpolation="none")
                plt.title("Ground Truth: {}".format(labs[i]))
                plt.xticks([])
                plt.yticks([])
            fig.savefig(os.path.join(images_dir, name + ".png"))
            plt.close(fig)
        except Exception:
            pass
        return path

    def save_curves(self, images_dir: str, name: str = "train_test_curve"):
        """CSV always; PNG if matplotlib is importable."""
        if self.rank != 0:
            return None
        os.makedirs(images_dir, exist_ok=True)
        path = os.path.join(images_dir, name + ".csv")
        with open(path, "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(["series", "examples_seen", "nll"])
            for c, l in zip(self.train_counter, self.train_losses):
                w.writerow(["train", c, l])
            for c, l in zip(self.test_counter, self.test_losses):
                w.writerow(["test", c, l])
        try:
            import matplotlib
            matplotlib.use("Agg")
            import matplotlib.pyplot as plt
            fig = plt.figure()
            plt.plot(self.train_counter, self.train_losses, color="blue")
            plt.scatter(self.test_counter[:len(self.test_losses)], self.test_losses, color="red")
            plt.legend(["Train Loss", "Test Loss"], loc="upper right")
            plt.xlabel("number of training examples seen")
            plt.ylabel("negative log likelihood loss")
            fig.savefig(os.path.join(images_dir, name + ".png"))
            plt.close(fig)
        except Exception:
            pass
        return path

    def close(self):
        if self._jsonl is not None:
            self._jsonl.close()
            self._jsonl = None
