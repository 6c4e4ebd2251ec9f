"""MNIST data: local IDX files when given, synthetic otherwise.

The reference downloads MNIST through torchvision (src/train.py:25-41,
src/train_dist.py:16-30); the GPU box has no network, so this loader reads the four IDX
files from `data_dir` (optionally .gz) and otherwise generates a synthetic stand-in of
the same shape and dtype: uint8 [N,28,28] images, labels in 0..9. Images stay uint8;
normalisation ((x/255 - 0.1307)/0.3081, T12) is fused into the kernels that consume them.
"""
from __future__ import annotations

import gzip
import os
import struct
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

_FILES = {
    True: ("train-images-idx3-ubyte", "train-labels-idx1-ubyte"),
    False: ("t10k-images-idx3-ubyte", "t10k-labels-idx1-ubyte"),
}


def _open(path):
    if os.path.exists(path):
        return open(path, "rb")
    if os.path.exists(path + ".gz"):
        return gzip.open(path + ".gz", "rb")
    return None


def read_idx(path: str) -> Optional[np.ndarray]:
    f = _open(path)
    if f is None:
        return None
    with f:
        magic = struct.unpack(">I", f.read(4))[0]
        ndim = magic & 0xFF
        if (magic >> 8) != 0x08:
            raise ValueError(f"{path}: not a uint8 IDX file (magic {magic:#x})")
        dims = struct.unpack(">" + "I" * ndim, f.read(4 * ndim))
        data = np.frombuffer(f.read(), dtype=np.uint8)
        return data.reshape(dims)


@dataclass
class MNISTArrays:
    images: torch.Tensor  # uint8 [N, 784]
    labels: torch.Tensor  # int64 [N]
    synthetic: bool

    def __len__(self):
        return self.images.shape[0]

    def to(self, device, pin: bool = False) -> "MNISTArrays":
        if pin and torch.cuda.is_available():
            return MNISTArrays(self.images.pin_memory(), self.labels.pin_memory(), self.synthetic)
        return MNISTArrays(self.images.to(device), self.labels.to(device), self.synthetic)


def synthetic_mnist(n: int, seed: int = 0, separable: bool = True) -> MNISTArrays:
    """Synthetic 28x28 uint8 digits. separable=True draws each image as a class template
    (a fixed random blob per class) plus noise, so a model can actually learn it;
    separable=False is pure noise with random labels (loss stays at ln 10)."""
    g = torch.Generator().manual_seed(seed)
    labels = torch.randint(0, 10, (n,), generator=g)
    if separable:
        tg = torch.Generator().manual_seed(1234)
        templates = (torch.rand(10, 7, 7, generator=tg) > 0.55).float()
        templates = templates.repeat_interleave(4, 1).repeat_interleave(4, 2)  # [10,28,28]
        img = templates[labels] * 200.0 + torch.rand(n, 28, 28, generator=g) * 55.0
        flip = torch.rand(n, 28, 28, generator=g) < 0.08
        img = torch.where(flip, 255.0 - img, img)
    else:
        img = torch.rand(n, 28, 28, generator=g) * 255.0
    return MNISTArrays(img.clamp(0, 255).to(torch.uint8).view(n, 784).contiguous(), labels, True)


def load_mnist(data_dir: str = "", train: bool = True, size: Optional[int] = None,
               seed: int = 0) -> MNISTArrays:
    """IDX files under data_dir (also data_dir/MNIST/raw, torchvision's layout) or synthetic."""
    if data_dir:
        for sub in ("", os.path.join("MNIST", "raw")):
            base = os.path.join(data_dir, sub)
            imgs = read_idx(os.path.join(base, _FILES[train][0]))
            labs = read_idx(os.path.join(base, _FILES[train][1]))
            if imgs is not None and labs is not None:
                x = torch.from_numpy(imgs.copy()).view(imgs.shape[0], 784)
                y = torch.from_numpy(labs.copy()).long()
                if size:
                    x, y = x[:size], y[:size]
                return MNISTArrays(x.contiguous(), y.contiguous(), False)
    n = size or (60000 if train else 10000)
    return synthetic_mnist(n, seed=seed + (0 if train else 1))


def normalized_float(ds: MNISTArrays) -> torch.Tensor:
    """fp32 [N,1,28,28] normalised like the reference's transform (src/train_dist.py:16-19)."""
    from ..models.oracle import normalize_u8
    return normalize_u8(ds.images)
