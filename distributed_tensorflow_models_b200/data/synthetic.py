"""Synthetic datasets of the reference's shapes (BASELINE.json mandates synthetic data for benchmarks: there
is no network to fetch CIFAR-10 / ImageNet — the role of ``maybe_download_and_extract``,
cnn/cifar10.py:394-411, cannot be played here).

``SyntheticBatches`` yields pinned host batches ``(images NHWC float32, labels int64)`` already in the
post-augmentation format the trainers consume; ``write_cifar10_binary`` / ``write_imagenet_tfrecords`` create
small on-disk datasets in the *real* file formats so the real readers can be exercised offline.
"""
from __future__ import annotations

import os
from typing import Iterator, Tuple

import numpy as np
import torch


class SyntheticBatches:
    """Endless iterator of random batches; a small pool is pre-generated and cycled."""

    def __init__(self, batch_size: int, image_size: int, num_classes: int, channels: int = 3, seed: int = 0,
                 pool: int = 8, pin: bool = True):
        g = torch.Generator().manual_seed(seed)
        self.pool = []
        for _ in range(pool):
            x = torch.randn(batch_size, image_size, image_size, channels, generator=g)
            y = torch.randint(0, num_classes, (batch_size,), generator=g)
            if pin and torch.cuda.is_available():
                x, y = x.pin_memory(), y.pin_memory()
            self.pool.append((x, y))
        self.i = 0

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        return self

    def __next__(self):
        b = self.pool[self.i % len(self.pool)]
        self.i += 1
        return b


def write_cifar10_binary(data_dir: str, train_per_file: int = 64, test: int = 64, seed: int = 0) -> str:
    """Write ``<data_dir>/cifar-10-batches-bin/{data_batch_1..5,test_batch}.bin`` with random records in the
    CIFAR-10 binary format (1 label byte + 3072 CHW bytes; cnn/cifar10_input.py:68-95)."""
    rng = np.random.RandomState(seed)
    d = os.path.join(data_dir, "cifar-10-batches-bin")
    os.makedirs(d, exist_ok=True)
    def dump(path, n):
        rec = np.empty((n, 1 + 3072), dtype=np.uint8)
        rec[:, 0] = rng.randint(0, 10, size=n)
        rec[:, 1:] = rng.randint(0, 256, size=(n, 3072))
        rec.tofile(path)
    for i in range(1, 6):
        dump(os.path.join(d, f"data_batch_{i}.bin"), train_per_file)
    dump(os.path.join(d, "test_batch.bin"), test)
    return d


class SyntheticCifarBatches:
    """Random 32x32x3 uint8 records pushed through the real CIFAR training distortion
    (``cifar10.distort_batch``): what a worker sees when no dataset is on disk."""

    def __init__(self, batch_size: int, image_size: int, seed: int = 0, pool: int = 8):
        from . import cifar10
        g = torch.Generator().manual_seed(seed)
        self.pool = []
        for _ in range(pool):
            raw = torch.randint(0, 256, (batch_size, 32, 32, 3), generator=g, dtype=torch.uint8)
            x = cifar10.distort_batch(raw, image_size, g)
            y = torch.randint(0, 10, (batch_size,), generator=g)
            if torch.cuda.is_available():
                x, y = x.pin_memory(), y.pin_memory()
            self.pool.append((x, y))
        self.i = 0

    def __iter__(self):
        return self

    def __next__(self):
        b = self.pool[self.i % len(self.pool)]
        self.i += 1
        return b


def write_imagenet_tfrecords(data_dir: str, subset: str = "train", num_images: int = 16, num_shards: int = 2,
                             size: int = 64, num_classes: int = 1000, seed: int = 0):
    """Tiny random-JPEG ImageNet look-alike in the real TFRecord/Example format (for offline tests)."""
    import io

    from PIL import Image

    from .build_image_data import convert_to_example
    from .tfrecord import TFRecordWriter
    rng = np.random.RandomState(seed)
    os.makedirs(data_dir, exist_ok=True)
    paths = []
    per = (num_images + num_shards - 1) // num_shards
    k = 0
    for s in range(num_shards):
        p = os.path.join(data_dir, "%s-%.5d-of-%.5d" % (subset, s, num_shards))
        with TFRecordWriter(p) as w:
            for _ in range(per):
                if k >= num_images:
                    break
                arr = rng.randint(0, 256, size=(size, size + 8, 3)).astype(np.uint8)
                buf = io.BytesIO()
                Image.fromarray(arr).save(buf, format="JPEG")
                label = int(rng.randint(1, num_classes + 1))
                w.write(convert_to_example("img_%d.JPEG" % k, buf.getvalue(), label, "synthetic", size, size + 8,
                                           [[0.1, 0.1, 0.9, 0.9]]))
                k += 1
        paths.append(p)
    return paths
