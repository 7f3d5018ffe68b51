"""Learning-rate table and optimiser hyper-parameters.

The reference builds ``SGD(lr=0.1, momentum=0.9, weight_decay=1e-4, nesterov=True)``
(ref:main.py:51-55) and ``MultiStepLR(milestones=[60, 80], gamma=0.1)`` (ref:main.py:57-59) but
steps the scheduler on rank 0 only and before the epoch's training (ref:main.py:69-70), so rank 0
trains epochs 1-59 at 0.1, 60-79 at 0.01, 80+ at 0.001 while other ranks never decay (SURVEY.md
D3).  Here every rank evaluates the same closed-form table, with the reference's epoch indexing.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Tuple


@dataclass(frozen=True)
class SGDConfig:
    lr: float = 0.1
    momentum: float = 0.9
    weight_decay: float = 1e-4
    nesterov: bool = True
    milestones: Tuple[int, ...] = (60, 80)
    gamma: float = 0.1

    def lr_at(self, epoch: int) -> float:
        """LR used while training 1-based ``epoch`` (scheduler stepped once per epoch, before
        training: after e steps ``last_epoch == e``; a milestone m takes effect when e >= m)."""
        n = sum(1 for m in self.milestones if epoch >= m)
        return self.lr * (self.gamma ** n)


def lr_for_epoch(epoch: int, base_lr: float = 0.1, milestones=(60, 80), gamma: float = 0.1) -> float:
    return SGDConfig(lr=base_lr, milestones=tuple(milestones), gamma=gamma).lr_at(epoch)
