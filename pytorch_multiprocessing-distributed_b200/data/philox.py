"""Host reproduction of the in-kernel crop/flip draws (csrc/data_aug.cu: philox4x32_10, gather_params).

With a device-resident dataset the training kernel draws a sample's augmentation from
Philox4x32-10(key = (seed, stream), counter = (dataset index, epoch, 0, 0)):

    x offset = r0 % (2*pad + 1),  y offset = r1 % (2*pad + 1),  flip = r2 & 1

i.e. a pure function of (seed, epoch, dataset index) -- seeded and different per epoch, unlike the
reference's unseeded, rank-identical torchvision transforms (ref:data.py:11-15), and
independent of which rank owns the sample.  This module computes the same numbers with numpy so the
CPU path (``augment_cpu``) is a bit-exact oracle for the GPU path and CPU-only runs see the same data.
"""
from __future__ import annotations

import numpy as np
import torch

M0, M1 = 0xD2511F53, 0xCD9E8D57
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = 0xFFFFFFFF
AUG_STREAM = 0x5EED0A06        # key word 1: names the augmentation stream


def philox4x32_10(counter: np.ndarray, key: tuple) -> np.ndarray:
    """counter: uint32 [..., 4]; key: (k0, k1).  Returns uint32 [..., 4]."""
    c = [counter[..., i].astype(np.uint64) for i in range(4)]
    k0, k1 = int(key[0]) & MASK, int(key[1]) & MASK
    for _ in range(10):
        p0 = np.uint64(M0) * c[0]
        p1 = np.uint64(M1) * c[2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & np.uint64(MASK)
        hi1, lo1 = p1 >> np.uint64(32), p1 & np.uint64(MASK)
        c = [hi1 ^ c[1] ^ np.uint64(k0), lo1, hi0 ^ c[3] ^ np.uint64(k1), lo0]
        k0, k1 = (k0 + W0) & MASK, (k1 + W1) & MASK
    return np.stack(c, axis=-1).astype(np.uint32)


def aug_params(indices, seed: int, epoch: int, train: bool, pad: int = 8) -> torch.Tensor:
    """int32 [n, 3] = (x offset, y offset, flip) for dataset indices ``indices`` -- what the kernel draws."""
    idx = np.asarray(indices, dtype=np.int64).reshape(-1)
    out = np.zeros((idx.size, 3), dtype=np.int32)
    if not train:
        out[:, 0] = pad
        out[:, 1] = pad
        return torch.from_numpy(out)
    ctr = np.zeros((idx.size, 4), dtype=np.uint32)
    ctr[:, 0] = idx.astype(np.uint32)
    ctr[:, 1] = np.uint32(int(epoch) & MASK)
    r = philox4x32_10(ctr, (int(seed) & MASK, AUG_STREAM))
    out[:, 0] = (r[:, 0] % np.uint32(2 * pad + 1)).astype(np.int32)
    out[:, 1] = (r[:, 1] % np.uint32(2 * pad + 1)).astype(np.int32)
    out[:, 2] = (r[:, 2] & np.uint32(1)).astype(np.int32)
    return torch.from_numpy(out)
