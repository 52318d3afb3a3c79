"""Data sources.

The reference downloads MNIST through torchvision, keeps the first tenth (6000 train /
1000 test images) and iterates it un-shuffled in batches of 60 on rank 0 only
(/root/reference/simple_distributed.py:87-95).  There is no network here, so the default
is a deterministic synthetic MNIST-shaped set (class-dependent stroke pattern + noise,
learnable to ~100 %); real MNIST idx files are used when present under ``data_root``.

Every rank can build the same dataset from the same seed, so the first pipeline stage
reads the images and the last stage reads the labels -- no label transport.
"""
from __future__ import annotations

import gzip
import os
import struct
from dataclasses import dataclass
from typing import Iterator, Optional, Tuple

import numpy as np
import torch


@dataclass
class ArrayDataset:
    """In-memory dataset of (images float32 [N,1,28,28] in [0,1], labels int64 [N])."""
    images: torch.Tensor
    labels: torch.Tensor

    def __len__(self) -> int:
        return self.images.shape[0]

    def subset(self, n: int) -> "ArrayDataset":
        return ArrayDataset(self.images[:n], self.labels[:n])


def synthetic_mnist(n: int, seed: int = 0, train: bool = True) -> ArrayDataset:
    """Deterministic MNIST-shaped data: 10 classes, each a fixed random low-frequency
    pattern, plus per-sample uniform noise.  Same (n, seed, train) -> same tensors on
    every rank and every run."""
    g = torch.Generator().manual_seed(1234567 + seed * 2 + (0 if train else 1))
    proto_g = torch.Generator().manual_seed(987654321)
    # class prototypes: smooth blobs (7x7 noise upsampled to 28x28)
    proto = torch.rand(10, 1, 7, 7, generator=proto_g)
    proto = torch.nn.functional.interpolate(proto, size=(28, 28), mode="bilinear", align_corners=False)
    proto = (proto - proto.amin(dim=(2, 3), keepdim=True)) / (
        proto.amax(dim=(2, 3), keepdim=True) - proto.amin(dim=(2, 3), keepdim=True) + 1e-6)
    labels = torch.randint(0, 10, (n,), generator=g)
    noise = torch.rand(n, 1, 28, 28, generator=g)
    images = (0.7 * proto[labels] + 0.3 * noise).clamp_(0.0, 1.0)
    return ArrayDataset(images.contiguous(), labels.contiguous())


def _read_idx(path: str) -> np.ndarray:
    opener = gzip.open if path.endswith(".gz") else open
    with opener(path, "rb") as f:
        data = f.read()
    magic, = struct.unpack(">I", data[:4])
    ndim = magic & 0xFF
    dims = struct.unpack(">" + "I" * ndim, data[4:4 + 4 * ndim])
    return np.frombuffer(data, dtype=np.uint8, offset=4 + 4 * ndim).reshape(dims)


def real_mnist(root: str, train: bool) -> Optional[ArrayDataset]:
    """Load MNIST idx files (torchvision layout ``<root>/MNIST/raw``) if they exist."""
    prefix = "train" if train else "t10k"
    for base in (os.path.join(root, "MNIST", "raw"), root):
        for ext in ("", ".gz"):
            img = os.path.join(base, f"{prefix}-images-idx3-ubyte{ext}")
            lab = os.path.join(base, f"{prefix}-labels-idx1-ubyte{ext}")
            if os.path.exists(img) and os.path.exists(lab):
                images = torch.from_numpy(_read_idx(img).astype(np.float32) / 255.0).unsqueeze(1)
                labels = torch.from_numpy(_read_idx(lab).astype(np.int64))
                return ArrayDataset(images.contiguous(), labels.contiguous())
    return None


def load_mnist_like(root: str, train: bool, n: int, synthetic: bool, seed: int) -> ArrayDataset:
    if not synthetic:
        ds = real_mnist(root, train)
        if ds is not None:
            return ds.subset(n)   # ref:91-92: first tenth only
    return synthetic_mnist(n, seed=seed, train=train)


def synthetic_mlp(n: int, in_dim: int, classes: int, seed: int = 0) -> ArrayDataset:
    """Wide-MLP inputs: unit-variance Gaussian features, labels from a fixed random
    linear teacher restricted to the first 64 features (cheap, learnable)."""
    g = torch.Generator().manual_seed(4242 + seed)
    x = torch.randn(n, in_dim, generator=g)
    k = min(64, in_dim)
    teacher = torch.randn(k, min(classes, 64), generator=torch.Generator().manual_seed(7))
    y = (x[:, :k] @ teacher).argmax(dim=1).to(torch.int64)
    return ArrayDataset(x.contiguous(), y.contiguous())


class BatchLoader:
    """Sequential, un-shuffled batches like ``DataLoader(ds, batch_size)`` (ref:94-95);
    the last batch may be short (100 train batches of 60; 17 test batches, last = 40)."""

    def __init__(self, ds: ArrayDataset, batch_size: int, drop_last: bool = False):
        self.ds = ds
        self.batch_size = batch_size
        self.drop_last = drop_last

    @property
    def dataset(self) -> ArrayDataset:
        return self.ds

    def __len__(self) -> int:
        n = len(self.ds)
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        n = len(self.ds)
        for i in range(len(self)):
            lo, hi = i * self.batch_size, min(n, (i + 1) * self.batch_size)
            yield self.ds.images[lo:hi], self.ds.labels[lo:hi]


def split_microbatches(n: int, m: int) -> list:
    """Sizes of ``m`` micro-batches covering ``n`` samples (ragged allowed: 60 -> 8,8,8,8,7,7,7,7).
    Zero-sized micro-batches are dropped, so fewer than ``m`` may be returned."""
    base, rem = divmod(n, m)
    sizes = [base + (1 if i < rem else 0) for i in range(m)]
    return [s for s in sizes if s > 0]
