"""Image sources held as uint8 NHWC arrays: synthetic CIFAR/ImageNet-shaped data and CIFAR-10
from disk.

The reference reads torchvision's ``CIFAR10(root='./cifar10_data', download=False)``
(ref:data.py:21-28) and decodes/augments per sample with PIL in 4 worker processes per loader
(ref:data.py:41-53).  Here a source is one contiguous ``uint8 [N, H, W, 3]`` tensor plus
``int64 [N]`` labels (pinned when CUDA is present) so that a per-rank batch is a single gather and
a single H2D copy, and augmentation runs as one fused GPU kernel (csrc/data_aug.cu) or as a
vectorised torch expression on CPU (data/loader.py).
"""
from __future__ import annotations

import os
import pickle
from dataclasses import dataclass
from typing import Tuple

import numpy as np
import torch


@dataclass
class ImageSource:
    images: torch.Tensor      # uint8 [N, H, W, 3]
    labels: torch.Tensor      # int64 [N]
    num_classes: int
    name: str

    def __len__(self) -> int:
        return int(self.images.shape[0])

    @property
    def hw(self) -> int:
        return int(self.images.shape[1])

    def pin(self) -> 'ImageSource':
        if torch.cuda.is_available() and not self.images.is_pinned():
            self.images = self.images.pin_memory()
            self.labels = self.labels.pin_memory()
        return self


def synthetic_source(n: int, hw: int = 32, num_classes: int = 10, seed: int = 0,
                     learnable: bool = True) -> ImageSource:
    """Random images. With ``learnable`` the class shifts the mean colour of the image a little,
    so a loss curve on synthetic data actually goes down (useful for end-to-end tests)."""
    rng = np.random.default_rng(seed)
    labels = rng.integers(0, num_classes, size=n, dtype=np.int64)
    imgs = rng.integers(0, 256, size=(n, hw, hw, 3), dtype=np.uint8)
    if learnable:
        tint = rng.integers(-48, 49, size=(num_classes, 1, 1, 3))
        imgs = np.clip(imgs.astype(np.int16) // 2 + 64 + tint[labels], 0, 255).astype(np.uint8)
    return ImageSource(torch.from_numpy(imgs), torch.from_numpy(labels), num_classes,
                       f'synthetic-{hw}x{hw}')


def ref_fake_source(n: int, hw: int = 32, train: bool = True) -> ImageSource:
    """The exact tensors of baseline/shim/ref_fake_cifar.py (the synthetic CIFAR10 stand-in the reference baseline
    harness trains on): same generator, same call order -- for loss-curve parity runs (tools/parity_vs_reference.py)."""
    rng = np.random.default_rng(0 if train else 1)
    imgs = rng.integers(0, 256, size=(n, hw, hw, 3), dtype=np.uint8)
    labels = np.asarray(rng.integers(0, 10, size=n), dtype=np.int64)
    return ImageSource(torch.from_numpy(imgs), torch.from_numpy(labels), 10, f'ref-fake-{hw}x{hw}')


def cifar10_source(root: str = './cifar10_data', train: bool = True) -> ImageSource:
    """CIFAR-10 python pickles under ``root/cifar-10-batches-py`` (the on-disk layout torchvision's
    ``CIFAR10(download=False)`` expects, ref:data.py:21-28). No torchvision dependency."""
    base = os.path.join(root, 'cifar-10-batches-py')
    names = [f'data_batch_{i}' for i in range(1, 6)] if train else ['test_batch']
    xs, ys = [], []
    for nm in names:
        path = os.path.join(base, nm)
        if not os.path.exists(path):
            raise FileNotFoundError(
                f'{path} not found: CIFAR-10 must already be on disk (the reference uses '
                f'download=False too); use --data synthetic when it is not')
        with open(path, 'rb') as f:
            d = pickle.load(f, encoding='latin1')
        xs.append(np.asarray(d['data'], dtype=np.uint8).reshape(-1, 3, 32, 32))
        ys.extend(d['labels'] if 'labels' in d else d['fine_labels'])
    x = np.ascontiguousarray(np.concatenate(xs).transpose(0, 2, 3, 1))
    return ImageSource(torch.from_numpy(x), torch.tensor(ys, dtype=torch.int64), 10,
                       'cifar10-train' if train else 'cifar10-test')


def make_sources(kind: str, root: str, hw: int, train_size: int, test_size: int,
                 num_classes: int = 10, seed: int = 0) -> Tuple[ImageSource, ImageSource]:
    if kind == 'cifar10':
        return cifar10_source(root, True), cifar10_source(root, False)
    if kind == 'ref-fake':
        return ref_fake_source(train_size, hw, True), ref_fake_source(test_size, hw, False)
    if kind == 'synthetic':
        return (synthetic_source(train_size, hw, num_classes, seed),
                synthetic_source(test_size, hw, num_classes, seed + 1))
    if kind == 'auto':
        try:
            return make_sources('cifar10', root, hw, train_size, test_size, num_classes, seed)
        except FileNotFoundError:
            return make_sources('synthetic', root, hw, train_size, test_size, num_classes, seed)
    raise ValueError(f"unknown data source '{kind}'")
