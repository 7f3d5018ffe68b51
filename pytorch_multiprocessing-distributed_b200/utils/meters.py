"""Running-average meter, text logger and top-k accuracy.

Behavioural parity targets (formats are byte-exact, see tests/test_utils.py):
  * AverageMeter  -- ref:utils.py:3-17  (sample-weighted running mean)
  * Logger        -- ref:utils.py:19-62 (``0001 2.302585 10.000000`` lines, append mode)
  * accuracy      -- ref:utils.py:64-77 (top-1 percent + per-sample correct mask)

Deliberate differences from the reference:
  * ``collections.Iterable`` (removed in Python 3.10) is replaced by ``collections.abc``.
  * ``accuracy(topk=(1, 5))`` works (the reference crashes on a non-contiguous view,
    SURVEY.md D6) and ``accuracy_topk`` returns every requested k.
  * ``Logger.write`` accepts 0-dim tensors / numpy scalars by converting them to Python
    numbers first (the reference raises ``Not supported type.``); pass ``strict=True`` to get
    the reference's type rules back.
"""
from __future__ import annotations

from collections.abc import Iterable
from typing import List, Sequence, Tuple

import torch


class AverageMeter:
    """val / sum / count / avg with sample weights (ref:utils.py:3-17)."""

    def __init__(self) -> None:
        self.reset()

    def reset(self) -> None:
        self.val = 0
        self.avg = 0
        self.sum = 0
        self.count = 0

    def update(self, val, n: int = 1) -> None:
        self.val = val
        self.sum += val * n
        self.count += n
        self.avg = self.sum / self.count


class Logger:
    """Space-separated text log, one row per ``write`` (ref:utils.py:19-62).

    ``int`` -> ``{:04d}``, ``float`` -> ``{:.6f}``, ``str`` -> verbatim; the first call fixes
    the column count; the file is opened in append mode on every write.
    """

    def __init__(self, path: str, int_form: str = ':04d', float_form: str = ':.6f',
                 strict: bool = False) -> None:
        self.path = path
        self.int_form = int_form
        self.float_form = float_form
        self.width = 0
        self.strict = strict

    def __len__(self) -> int:
        try:
            return len(self.read())
        except Exception:
            return 0

    @staticmethod
    def _to_python(v):
        if isinstance(v, (int, float, str)):
            return v
        if isinstance(v, torch.Tensor) and v.numel() == 1:
            v = v.item()
            return v
        if hasattr(v, 'item') and getattr(v, 'shape', None) == ():
            return v.item()
        return v

    def write(self, values) -> None:
        if not isinstance(values, Iterable) or isinstance(values, str):
            values = [values]
        values = list(values)
        if not self.strict:
            values = [self._to_python(v) for v in values]
        if self.width == 0:
            self.width = len(values)
        assert self.width == len(values), 'Inconsistent number of items.'
        cells = []
        for v in values:
            if isinstance(v, int):
                cells.append(('{' + self.int_form + '}').format(v))
            elif isinstance(v, float):
                cells.append(('{' + self.float_form + '}').format(v))
            elif isinstance(v, str):
                cells.append(v)
            else:
                raise Exception('Not supported type.')
        with open(self.path, 'a') as f:
            f.write(' '.join(cells) + '\n')

    def read(self) -> List[list]:
        rows = []
        with open(self.path, 'r') as f:
            for line in f:
                row = []
                for tok in line.split(' '):
                    try:
                        tok = float(tok)
                    except ValueError:
                        pass
                    row.append(tok)
                rows.append(row)
        return rows

    def truncate(self) -> None:
        open(self.path, 'w').close()
        self.width = 0


def accuracy_topk(output: torch.Tensor, target: torch.Tensor,
                  topk: Sequence[int] = (1, 5)) -> Tuple[List[torch.Tensor], torch.Tensor]:
    """Percent correct for every k in ``topk`` plus the ``[maxk, B]`` hit matrix."""
    maxk = min(max(topk), output.size(1))
    batch = target.size(0)
    _, pred = output.topk(maxk, 1, True, True)
    pred = pred.t()
    correct = pred.eq(target.view(1, -1).expand_as(pred))
    res = []
    for k in topk:
        k = min(k, maxk)
        res.append(correct[:k].reshape(-1).float().sum(0) * (100.0 / batch))
    return res, correct


def accuracy(output: torch.Tensor, target: torch.Tensor, topk: Sequence[int] = (1,)):
    """Reference-shaped return: ``(top-1 percent as 0-dim tensor, correct.squeeze())``
    (ref:utils.py:64-77)."""
    res, correct = accuracy_topk(output, target, topk)
    return res[0], correct.squeeze()


def topk_counts(output: torch.Tensor, target: torch.Tensor) -> Tuple[int, int]:
    """Number of top-1 and top-5 hits (integers), used by the CPU trainer path."""
    _, correct = accuracy_topk(output, target, (1, 5))
    c1 = int(correct[:1].sum().item())
    c5 = int(correct[:5].sum().item())
    return c1, c5
