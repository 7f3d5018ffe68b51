"""Per-rank batch iterator + the augmentation math shared by the CPU and GPU paths.

Parity target: ``get_loader`` (ref:data.py:6-59).  Train transform = ``RandomCrop(32, padding=8)``
-> ``RandomHorizontalFlip`` -> ``ToTensor`` -> ``Normalize(0.5, 0.5)`` (ref:data.py:11-15); test
transform = ``ToTensor`` -> ``Normalize`` (ref:data.py:17-19); per-rank batch =
``int(batch_size / world_size)`` (ref:data.py:39); no ``drop_last``.

Design: a batch is (uint8 NHWC images, labels, valid mask, crop/flip parameters).  The random
crop offsets and flips are drawn on the host from a generator seeded by (seed, epoch, rank) --
the reference leaves them unseeded and identical across ranks -- and the pixel
work is done either by ``augment_cpu`` (torch, oracle + CPU trainer) or by the fused sm_100a
kernel (csrc/data_aug.cu: gather + zero-pad crop + flip + normalise + NHWC bf16, channel-padded).
Both consume the same parameters, so the GPU kernel is testable bit-for-bit against the CPU path.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Iterator, Optional

import torch

from .sampler import ShardSampler
from .sources import ImageSource

CROP_PAD = 8          # ref:data.py:12
MEAN, STD = 0.5, 0.5  # ref:data.py:7-8


@dataclass
class Batch:
    u8: torch.Tensor        # uint8 [B, H, W, 3] (pinned host memory when CUDA is present)
    labels: torch.Tensor    # int64 [B]
    valid: torch.Tensor     # bool  [B]  False for sampler wrap-around duplicates
    aug: torch.Tensor       # int32 [B, 3] = (x offset, y offset, flip); offsets in [0, 2*pad]
    train: bool

    @property
    def n(self) -> int:
        return int(self.labels.shape[0])


def draw_aug_params(n: int, train: bool, gen: Optional[torch.Generator], pad: int = CROP_PAD) -> torch.Tensor:
    if not train:
        p = torch.zeros(n, 3, dtype=torch.int32)
        p[:, 0] = pad
        p[:, 1] = pad
        return p
    ox = torch.randint(0, 2 * pad + 1, (n,), generator=gen, dtype=torch.int32)
    oy = torch.randint(0, 2 * pad + 1, (n,), generator=gen, dtype=torch.int32)
    fl = torch.randint(0, 2, (n,), generator=gen, dtype=torch.int32)
    return torch.stack([ox, oy, fl], dim=1).contiguous()


def augment_cpu(u8: torch.Tensor, aug: torch.Tensor, pad: int = CROP_PAD) -> torch.Tensor:
    """uint8 [B,H,W,3] + params -> float32 NCHW normalised to [-1, 1]; pixels that fall in the
    zero padding become (0/255 - 0.5)/0.5 = -1, exactly as torchvision pads *before* ToTensor."""
    b, h, w, _ = u8.shape
    ys = torch.arange(h).view(1, h, 1) + aug[:, 1].view(b, 1, 1).long() - pad        # source row
    xs_out = torch.arange(w).view(1, 1, w).expand(b, 1, w)
    flip = aug[:, 2].view(b, 1, 1).bool()
    xs_crop = torch.where(flip, (w - 1) - xs_out, xs_out)                            # undo the flip
    xs = xs_crop + aug[:, 0].view(b, 1, 1).long() - pad                              # source col
    inside = (ys >= 0) & (ys < h) & (xs >= 0) & (xs < w)                             # [B,H,W]
    ysc = ys.clamp(0, h - 1).expand(b, h, w)
    xsc = xs.clamp(0, w - 1).expand(b, h, w)
    bi = torch.arange(b).view(b, 1, 1).expand(b, h, w)
    px = u8[bi, ysc, xsc].to(torch.float32)                                          # [B,H,W,3]
    px = px * inside.unsqueeze(-1)
    out = (px / 255.0 - MEAN) / STD
    return out.permute(0, 3, 1, 2).contiguous()


@dataclass
class IndexBatch:
    """A batch of a DEVICE-RESIDENT dataset: only the indices travel (``packed`` = int32 [4 + B] in pinned memory:
    {seed, stream, epoch, train} + dataset indices, ~i for an uncounted wrap-around duplicate).  Images and labels
    are gathered, cropped / flipped (in-kernel Philox) and normalised on the GPU (csrc/data_aug.cu)."""
    packed: torch.Tensor
    n: int
    train: bool
    seed: int
    epoch: int
    draw: bool = True          # crop / flip are drawn (training with augmentation)

    @property
    def indices(self) -> torch.Tensor:
        raw = self.packed[4:4 + self.n]
        return torch.where(raw < 0, ~raw, raw).long()

    @property
    def valid(self) -> torch.Tensor:
        return self.packed[4:4 + self.n] >= 0

    def to_host_batch(self, source: ImageSource) -> 'Batch':
        """The same batch in host form with the crop/flip the kernel will draw (oracle / CPU path)."""
        from .philox import aug_params
        idx = self.indices
        return Batch(source.images.index_select(0, idx), source.labels.index_select(0, idx), self.valid.clone(),
                     aug_params(idx.numpy(), self.seed, self.epoch, self.train and self.draw), self.train)


class DeviceBatchLoader:
    """Index-only iterator over a dataset that lives in GPU memory (``NativeEngine.attach_dataset``): replaces the
    reference's worker processes + pin thread + blocking ``.cuda()`` (ref:data.py:41-53, ref:main.py:101).  Per
    step the host writes B+4 int32 into a pinned slot; the sampler's shard for the epoch is the only host work."""

    SLOTS = 4      # pinned slots in rotation: a slot is rewritten SLOTS steps after its H2D copy was enqueued

    def __init__(self, source: ImageSource, sampler: ShardSampler, batch_size: int, train: bool, seed: int = 0,
                 augment: bool = True) -> None:
        if batch_size <= 0:
            raise ValueError('per-rank batch size must be positive (global batch smaller than world size?)')
        self.source, self.sampler, self.batch_size, self.train, self.seed = source, sampler, batch_size, train, seed
        self.augment = augment
        self.epoch = 0
        slots = torch.empty((self.SLOTS, 4 + batch_size), dtype=torch.int32)
        self._slots = slots.pin_memory() if torch.cuda.is_available() else slots
        self._next = 0

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)
        self.sampler.set_epoch(epoch)

    def __len__(self) -> int:
        return (len(self.sampler) + self.batch_size - 1) // self.batch_size

    @property
    def dataset_len(self) -> int:
        return len(self.source)

    def __iter__(self) -> Iterator[IndexBatch]:
        from .philox import AUG_STREAM
        idx, valid = self.sampler.indices_and_mask()
        enc = torch.where(valid, idx, ~idx).to(torch.int32)
        draw = self.train and self.augment           # header word 3: draw crop / flip (else centre crop, no flip)
        hdr = torch.tensor([self.seed & 0x7FFFFFFF, AUG_STREAM, self.epoch, int(draw)], dtype=torch.int32)
        for s in range(0, enc.numel(), self.batch_size):
            n = min(self.batch_size, enc.numel() - s)
            slot = self._slots[self._next % self.SLOTS]
            self._next += 1
            slot[:4] = hdr
            slot[4:4 + n] = enc[s:s + n]
            yield IndexBatch(slot, n, self.train, self.seed & 0x7FFFFFFF, self.epoch, draw)


class BatchLoader:
    def __init__(self, source: ImageSource, sampler: ShardSampler, batch_size: int, train: bool,
                 seed: int = 0, pin: bool = False, augment: bool = True) -> None:
        if batch_size <= 0:
            raise ValueError('per-rank batch size must be positive '
                             '(global batch smaller than world size?)')
        self.augment = augment          # False: training batches with the centre crop, no flip (parity runs)
        self.source = source
        self.sampler = sampler
        self.batch_size = batch_size
        self.train = train
        self.seed = seed
        self.pin = pin and torch.cuda.is_available()
        self.epoch = 0
        self._stage_u8 = None
        self._stage_lb = None

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)
        self.sampler.set_epoch(epoch)

    def __len__(self) -> int:
        return (len(self.sampler) + self.batch_size - 1) // self.batch_size

    @property
    def dataset_len(self) -> int:
        return len(self.source)

    def __iter__(self) -> Iterator[Batch]:
        idx, valid = self.sampler.indices_and_mask()
        gen = torch.Generator()
        gen.manual_seed((self.seed * 1000003 + self.epoch) * 1009 + self.sampler.rank + 17)
        for s in range(0, idx.numel(), self.batch_size):
            sel = idx[s:s + self.batch_size]
            n = sel.numel()
            if self.pin:
                # two staging buffers so the copy of batch i+1 can be filled while i is in flight
                if self._stage_u8 is None:
                    shp = (2, self.batch_size) + tuple(self.source.images.shape[1:])
                    self._stage_u8 = torch.empty(shp, dtype=torch.uint8).pin_memory()
                    self._stage_lb = torch.empty((2, self.batch_size), dtype=torch.int64).pin_memory()
                slot = (s // self.batch_size) & 1
                u8 = self._stage_u8[slot, :n]
                lb = self._stage_lb[slot, :n]
                torch.index_select(self.source.images, 0, sel, out=u8)
                torch.index_select(self.source.labels, 0, sel, out=lb)
            else:
                u8 = self.source.images.index_select(0, sel)
                lb = self.source.labels.index_select(0, sel)
            yield Batch(u8, lb, valid[s:s + n], draw_aug_params(n, self.train and self.augment, gen), self.train)


def make_loaders(train_src: ImageSource, test_src: ImageSource, global_batch: int, world_size: int,
                 rank: int, seed: int = 0, pin: bool = False, compat_eval_shuffle: bool = False, augment: bool = True):
    """(train_loader, test_loader) for one rank (ref:data.py:31-53)."""
    from .sampler import per_rank_batch
    bs = per_rank_batch(global_batch, world_size)
    tr = BatchLoader(train_src, ShardSampler(len(train_src), world_size, rank, True, seed), bs, True, seed, pin, augment)
    te = BatchLoader(test_src, ShardSampler(len(test_src), world_size, rank, compat_eval_shuffle, seed),
                     bs, False, seed, pin)
    return tr, te


def make_device_loaders(train_src: ImageSource, test_src: ImageSource, global_batch: int, world_size: int,
                        rank: int, seed: int = 0, compat_eval_shuffle: bool = False, augment: bool = True):
    """Index-only loaders for datasets resident in GPU memory (same sharding as make_loaders)."""
    from .sampler import per_rank_batch
    bs = per_rank_batch(global_batch, world_size)
    tr = DeviceBatchLoader(train_src, ShardSampler(len(train_src), world_size, rank, True, seed), bs, True, seed, augment)
    te = DeviceBatchLoader(test_src, ShardSampler(len(test_src), world_size, rank, compat_eval_shuffle, seed),
                           bs, False, seed)
    return tr, te
