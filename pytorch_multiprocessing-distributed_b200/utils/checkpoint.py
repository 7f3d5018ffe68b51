"""Checkpoint write (reference-compatible) and full-state resume.

Parity target: ``torch.save(model.state_dict(), save_path/'model_{epoch}.pth')`` on rank 0 at the
last epoch (ref:main.py:75-77).  Because the reference saves the *DDP wrapper*, every key carries
a ``module.`` prefix (74 entries for ResNet18: 38 parameters + 36 BN buffers, fp32 + int64
``num_batches_tracked``; SURVEY.md probe:4).  ``save_model`` writes exactly that key set, as
contiguous CPU tensors, so the file is interchangeable with the reference's.

The reference has no load/resume code at all.  Added here: a second file
``state_{epoch}.pth`` with momentum, epoch, step count and the flag values, ``--save-every`` for
periodic checkpoints, and ``load_model`` that accepts keys with or without the prefix.
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Dict, Optional, Tuple

import torch

PREFIX = 'module.'


def to_reference_keys(sd: Dict[str, torch.Tensor]) -> 'OrderedDict[str, torch.Tensor]':
    out = OrderedDict()
    for k, v in sd.items():
        out[k if k.startswith(PREFIX) else PREFIX + k] = v.detach().to('cpu').contiguous().clone()
    return out


def strip_prefix(sd: Dict[str, torch.Tensor]) -> 'OrderedDict[str, torch.Tensor]':
    return OrderedDict((k[len(PREFIX):] if k.startswith(PREFIX) else k, v) for k, v in sd.items())


def model_path(save_path: str, epoch: int) -> str:
    return os.path.join(save_path, 'model_{0}.pth'.format(epoch))


def state_path(save_path: str, epoch: int) -> str:
    return os.path.join(save_path, 'state_{0}.pth'.format(epoch))


def save_model(save_path: str, epoch: int, model_state: Dict[str, torch.Tensor]) -> str:
    path = model_path(save_path, epoch)
    torch.save(to_reference_keys(model_state), path)
    return path


def _plain(v):
    """Only tensors, numbers, strings, None and lists / dicts of those go into a state file, so that it loads with
    torch's restricted unpickler (``weights_only=True``): a checkpoint directory is data, not code."""
    if isinstance(v, torch.Tensor) or v is None or isinstance(v, (bool, int, float, str)):
        return v
    if isinstance(v, dict):
        return {str(k): _plain(x) for k, x in v.items()}
    if isinstance(v, (list, tuple)):
        return [_plain(x) for x in v]
    return repr(v)


def save_train_state(save_path: str, epoch: int, optimizer_state: dict, extra: Optional[dict] = None) -> str:
    path = state_path(save_path, epoch)
    torch.save(_plain({'epoch': epoch, 'optimizer': optimizer_state, 'extra': extra or {}}), path)
    return path


def load_train_state(path: str) -> dict:
    return torch.load(path, map_location='cpu', weights_only=True)


def load_model(path: str) -> 'OrderedDict[str, torch.Tensor]':
    sd = torch.load(path, map_location='cpu', weights_only=True)
    return strip_prefix(sd)


def find_resume(path: str) -> Tuple[str, Optional[str], int]:
    """``path`` is a ``model_E.pth`` file or a run directory (latest epoch wins).
    Returns (model file, state file or None, epoch)."""
    if os.path.isdir(path):
        epochs = []
        for f in os.listdir(path):
            if f.startswith('model_') and f.endswith('.pth'):
                try:
                    epochs.append(int(f[len('model_'):-len('.pth')]))
                except ValueError:
                    pass
        if not epochs:
            raise FileNotFoundError(f'no model_*.pth in {path}')
        e = max(epochs)
        mp = model_path(path, e)
    else:
        mp = path
        name = os.path.basename(path)
        if not (name.startswith('model_') and name.endswith('.pth') and name[len('model_'):-len('.pth')].isdigit()):
            raise ValueError(f"--resume expects a run directory or a file named model_<epoch>.pth, got '{path}'")
        e = int(name[len('model_'):-len('.pth')])
    sp = state_path(os.path.dirname(mp) or '.', e)
    return mp, (sp if os.path.exists(sp) else None), e
