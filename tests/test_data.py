"""Data sources (replace ref:87-95): deterministic synthetic sets, the real-MNIST idx reader,
the reference's batching (100 train batches of 60; 17 test batches, the last one 40)."""
import gzip
import struct

import numpy as np
import torch

from simple_distributed_machine_learning_b200.utils.data import (BatchLoader, load_mnist_like, real_mnist,
                                                                 split_microbatches, synthetic_mlp,
                                                                 synthetic_mnist)


def test_synthetic_sets_are_deterministic_and_shaped_like_mnist():
    a, b = synthetic_mnist(256, seed=3), synthetic_mnist(256, seed=3)
    assert torch.equal(a.images, b.images) and torch.equal(a.labels, b.labels)
    assert a.images.shape == (256, 1, 28, 28) and a.images.dtype == torch.float32
    assert 0.0 <= float(a.images.min()) and float(a.images.max()) <= 1.0
    assert a.labels.dtype == torch.int64 and set(a.labels.tolist()) <= set(range(10))
    assert not torch.equal(a.images, synthetic_mnist(256, seed=4).images)
    assert not torch.equal(a.images, synthetic_mnist(256, seed=3, train=False).images)
    m = synthetic_mlp(128, 512, 512, seed=1)
    assert m.images.shape == (128, 512) and torch.equal(m.labels, synthetic_mlp(128, 512, 512, seed=1).labels)


def test_reference_batching():
    train = BatchLoader(synthetic_mnist(6000), 60)
    test = BatchLoader(synthetic_mnist(1000, train=False), 60)
    assert len(train) == 100 and len(test) == 17                      # ref:91-95
    sizes = [d.shape[0] for d, _ in test]
    assert sizes[:-1] == [60] * 16 and sizes[-1] == 40
    first = next(iter(train))
    assert torch.equal(first[0], train.dataset.images[:60])          # shuffle=False in the reference
    assert len(BatchLoader(synthetic_mnist(1000), 60, drop_last=True)) == 16


def test_split_microbatches_ragged():
    assert split_microbatches(60, 8) == [8, 8, 8, 8, 7, 7, 7, 7]
    assert split_microbatches(3, 8) == [1, 1, 1]
    assert sum(split_microbatches(16384, 16)) == 16384


def _write_idx(path, arr, gz):
    header = struct.pack(">I", (0x08 << 8) | arr.ndim) + struct.pack(">" + "I" * arr.ndim, *arr.shape)
    opener = gzip.open if gz else open
    with opener(path, "wb") as f:
        f.write(header + arr.astype(np.uint8).tobytes())


def test_real_mnist_idx_files_are_used_when_present(tmp_path):
    rng = np.random.default_rng(0)
    raw = tmp_path / "MNIST" / "raw"
    raw.mkdir(parents=True)
    imgs = rng.integers(0, 256, size=(50, 28, 28), dtype=np.uint8)
    labs = rng.integers(0, 10, size=(50,), dtype=np.uint8)
    _write_idx(str(raw / "train-images-idx3-ubyte"), imgs, gz=False)
    _write_idx(str(raw / "train-labels-idx1-ubyte"), labs, gz=False)
    _write_idx(str(raw / "t10k-images-idx3-ubyte.gz"), imgs[:20], gz=True)
    _write_idx(str(raw / "t10k-labels-idx1-ubyte.gz"), labs[:20], gz=True)
    ds = real_mnist(str(tmp_path), train=True)
    assert ds.images.shape == (50, 1, 28, 28) and ds.labels.tolist() == labs.tolist()
    torch.testing.assert_close(ds.images[:, 0], torch.from_numpy(imgs.astype(np.float32) / 255.0))   # ToTensor()
    assert len(real_mnist(str(tmp_path), train=False)) == 20
    # the loader keeps the first n samples (ref:91-92 keeps the first tenth) and falls back to synthetic data
    assert len(load_mnist_like(str(tmp_path), True, 10, synthetic=False, seed=0)) == 10
    assert len(load_mnist_like(str(tmp_path / "nothing_here"), True, 10, synthetic=False, seed=0)) == 10
