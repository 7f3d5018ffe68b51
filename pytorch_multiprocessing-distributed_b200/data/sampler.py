"""Per-rank index sharding, index-exact with ``torch.utils.data.DistributedSampler``.

The reference shards both train and test sets with ``DistributedSampler(shuffle=True)``
(ref:data.py:31-37) and never calls ``set_epoch``, so it replays one permutation
forever.  This sampler keeps torch's index math (torch:utils/data/distributed.py:98-134:
``randperm(n, generator=seed+epoch)``, wrap-around padding to ``ceil(n/W)*W``, stride-W
subsample) so a run can be compared sample-for-sample against the reference, and adds what the
reference lacks: a per-epoch reshuffle, ``shuffle=False`` for evaluation and a mask that marks
the wrap-around duplicates so metrics can ignore them.
"""
from __future__ import annotations

import math
from typing import Tuple

import torch


class ShardSampler:
    def __init__(self, dataset_len: int, world_size: int, rank: int, shuffle: bool = True,
                 seed: int = 0, drop_last: bool = False) -> None:
        if not (0 <= rank < world_size):
            raise ValueError(f'rank {rank} outside [0, {world_size})')
        self.n = int(dataset_len)
        self.world_size = world_size
        self.rank = rank
        self.shuffle = shuffle
        self.seed = seed
        self.drop_last = drop_last
        self.epoch = 0
        if drop_last and self.n % world_size != 0:
            self.num_samples = math.ceil((self.n - world_size) / world_size)
        else:
            self.num_samples = math.ceil(self.n / world_size)
        self.total_size = self.num_samples * world_size

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)

    def __len__(self) -> int:
        return self.num_samples

    def indices_and_mask(self) -> Tuple[torch.Tensor, torch.Tensor]:
        """(indices [num_samples] int64, valid [num_samples] bool). ``valid`` is False for the
        entries that exist only because the index list was padded by wrapping around."""
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            order = torch.randperm(self.n, generator=g)
        else:
            order = torch.arange(self.n)
        valid = torch.ones(self.n, dtype=torch.bool)
        if not self.drop_last:
            pad = self.total_size - self.n
            if pad > 0:
                reps = math.ceil(pad / max(self.n, 1))
                order = torch.cat([order, order.repeat(reps)[:pad]])
                valid = torch.cat([valid, torch.zeros(pad, dtype=torch.bool)])
        else:
            order = order[:self.total_size]
            valid = valid[:self.total_size]
        sel = slice(self.rank, self.total_size, self.world_size)
        return order[sel].contiguous(), valid[sel].contiguous()

    def indices(self) -> torch.Tensor:
        return self.indices_and_mask()[0]

    def __iter__(self):
        return iter(self.indices().tolist())


def per_rank_batch(global_batch: int, world_size: int) -> int:
    """``int(batch_size / world_size)`` exactly as ref:data.py:39 (the remainder is dropped)."""
    return int(global_batch / world_size)
