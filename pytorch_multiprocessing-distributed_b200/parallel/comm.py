"""Process-group bring-up and the small collective surface the trainer needs.

Parity target: ``init_process`` (ref:main.py:190-193: MASTER_ADDR 127.0.0.1, fixed port 20080,
``init_process_group('nccl')``) and ``destroy_process_group`` (ref:main.py:84).

Two transports implement the same ``Comm`` interface:
  * ``TorchComm``  -- torch.distributed (gloo on CPU, NCCL on GPU).  This is the plumbing /
    baseline / oracle transport (BASELINE.json: "a path that only calls NCCL ... is the baseline").
  * ``PeerComm``   -- parallel/peer.py: the product transport, symmetric peer-mapped buffers over
    NVLink 5 / NVSwitch driven by hand-written kernels (csrc/comm.cu, csrc/sym.cuh; host side csrc/runtime.cpp).
The endpoint comes from the launcher's environment (unique per run) instead of a fixed port.
"""
from __future__ import annotations

import datetime
import hashlib
import os
from typing import List, Sequence

import torch
import torch.distributed as dist


class Comm:
    rank: int
    world_size: int
    kind: str

    def broadcast_(self, t: torch.Tensor, src: int = 0) -> None: raise NotImplementedError
    def all_reduce_sum_(self, t: torch.Tensor, async_op: bool = False): raise NotImplementedError
    def barrier(self) -> None: raise NotImplementedError
    def all_gather_bytes(self, payload: bytes) -> List[bytes]: raise NotImplementedError
    def close(self) -> None: pass

    # -- derived helpers -------------------------------------------------------------------
    def verify_same(self, what: str, payload: bytes) -> None:
        """Replacement for DDP's ``_verify_param_shape_across_processes``
        (torch:nn/parallel/distributed.py:881): every rank must hold the same digest."""
        digest = hashlib.sha256(payload).digest()
        got = self.all_gather_bytes(digest)
        bad = [r for r, d in enumerate(got) if d != got[0]]
        if bad:
            raise RuntimeError(f'{what} differs across ranks: ranks {bad} disagree with rank 0')

    def reduce_metrics(self, values: Sequence[float]) -> List[float]:
        """Global SUM of a short float vector -- the live version of the reference's dead
        ``reduce_tensor`` (ref:main.py:173-177)."""
        t = torch.tensor(list(values), dtype=torch.float64)
        if self.world_size > 1:
            t = self._metric_tensor(t)
            self.all_reduce_sum_(t)
        return t.cpu().tolist()

    def _metric_tensor(self, t: torch.Tensor) -> torch.Tensor:
        return t


class SingleComm(Comm):
    """world_size == 1: every collective is the identity."""
    kind = 'single'

    def __init__(self) -> None:
        self.rank, self.world_size = 0, 1

    def broadcast_(self, t, src=0): return None
    def all_reduce_sum_(self, t, async_op=False): return None
    def barrier(self): return None
    def all_gather_bytes(self, payload): return [payload]


class TorchComm(Comm):
    def __init__(self, rank: int, world_size: int, backend: str = 'gloo', device=None,
                 timeout_s: float = 600.0) -> None:
        self.rank, self.world_size, self.kind = rank, world_size, backend
        self.device = device
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        os.environ.setdefault('MASTER_PORT', '29533')
        self._owns = not dist.is_initialized()
        if self._owns:
            kw = {}
            if backend == 'nccl' and device is not None:
                kw['device_id'] = device
            dist.init_process_group(backend, rank=rank, world_size=world_size,
                                    timeout=datetime.timedelta(seconds=timeout_s), **kw)

    def broadcast_(self, t, src=0):
        dist.broadcast(t, src)

    def all_reduce_sum_(self, t, async_op=False):
        return dist.all_reduce(t, op=dist.ReduceOp.SUM, async_op=async_op)

    def barrier(self):
        dist.barrier()

    def all_gather_bytes(self, payload: bytes) -> List[bytes]:
        out: List[bytes] = [b''] * self.world_size
        dist.all_gather_object(out, payload)
        return out

    def _metric_tensor(self, t):
        return t.to(self.device) if self.kind == 'nccl' else t

    def close(self):
        if self._owns and dist.is_initialized():
            dist.destroy_process_group()


def make_comm(rank: int, world_size: int, backend: str, device=None) -> Comm:
    if world_size == 1 and backend in ('single', 'auto'):
        return SingleComm()
    if backend in ('gloo', 'nccl'):
        return TorchComm(rank, world_size, backend, device)
    raise ValueError(f"unknown comm backend '{backend}'")
