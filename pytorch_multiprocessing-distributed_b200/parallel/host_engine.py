"""Torch-autograd data-parallel engine: the CPU/gloo plumbing path and the NCCL baseline/oracle.

Parity target: what ``DDP(model, device_ids=[rank])`` + ``optim.SGD(...).step()`` do for the
reference (ref:main.py:44, 51-55, 104-110; torch Reducer, reducer.hpp:74-524): verify structure
across ranks, broadcast rank 0's parameters and buffers, then per iteration average gradients
bucket by bucket while backward is still running and apply SGD-Nesterov.

Structure (mirrors the native engine so both are driven by the same trainer):
  * parameters, gradients and momentum live in three flat fp32 slabs laid out in gradient-arrival
    order (parallel/buckets.py); ``p.data`` and ``p.grad`` are views, so there is no
    copy-in/copy-out around the collective (torch's Reducer copies both ways);
  * a post-accumulate hook per parameter counts down its bucket and fires an async all-reduce on
    the bucket's slab range when the last gradient lands;
  * after backward, one fused slab-wide update does x(1/W), weight decay, momentum, Nesterov
    (math = torch:optim/sgd.py:343-380; zero-initialised momentum == torch's first-step ``buf=g``).
The per-forward buffer broadcast torch DDP performs (C2b) is dropped: with synchronised BN every
rank computes identical running statistics.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..data.loader import Batch, augment_cpu
from ..utils.schedule import SGDConfig
from .buckets import (TORCH_BUCKET_CAP_BYTES, TORCH_FIRST_BUCKET_BYTES, Bucket, params_of, plan_buckets)
from .comm import Comm
from .sync_bn import convert_sync_batchnorm


class HostEngine:
    kind = 'host'

    def __init__(self, model: nn.Module, comm: Comm, sgd: SGDConfig, device: torch.device,
                 sync_bn: bool = True, first_bucket_bytes: int = TORCH_FIRST_BUCKET_BYTES,
                 bucket_bytes: int = TORCH_BUCKET_CAP_BYTES, autocast_bf16: bool = False) -> None:
        self.comm, self.sgd, self.device = comm, sgd, device
        self.autocast_bf16 = autocast_bf16
        model = model.to(device)
        if sync_bn and comm.world_size > 1:
            model = convert_sync_batchnorm(model, comm)
        self.model = model
        infos = params_of(model)
        comm.verify_same('model structure',
                         repr([(n, tuple(p.shape), str(p.dtype)) for n, p in model.named_parameters()]).encode())
        self.buckets, total = plan_buckets(infos, first_bucket_bytes, bucket_bytes)
        self.flat_p = torch.zeros(total, dtype=torch.float32, device=device)
        self.flat_g = torch.zeros(total, dtype=torch.float32, device=device)
        self.flat_m = torch.zeros(total, dtype=torch.float32, device=device)
        params = dict(model.named_parameters())
        self._bucket_of: Dict[str, int] = {}
        for bi, b in enumerate(self.buckets):
            for e in b.entries:
                p = params[e.name]
                view = self.flat_p[e.offset:e.offset + e.numel].view_as(p)
                view.copy_(p.data)
                p.data = view
                p.grad = self.flat_g[e.offset:e.offset + e.numel].view_as(p)
                self._bucket_of[e.name] = bi
        self._pending = [0] * len(self.buckets)
        self._works: List[Tuple[int, object]] = []
        self.bucket_launch_order: List[int] = []
        if comm.world_size > 1:
            for name, p in params.items():
                p.register_post_accumulate_grad_hook(self._make_hook(self._bucket_of[name]))
        self._broadcast_state()
        self.step_count = 0
        self._m = torch.zeros(4, dtype=torch.float64)     # loss_sum, correct1, correct5, count

    # ---------------------------------------------------------------- initial state (C1, C2)
    def _broadcast_state(self) -> None:
        if self.comm.world_size == 1:
            return
        self.comm.broadcast_(self.flat_p, 0)
        for _, buf in self.model.named_buffers():
            self.comm.broadcast_(buf, 0)

    # ---------------------------------------------------------------- gradient path (N4, C5)
    def _make_hook(self, bi: int):
        def hook(_param):
            self._pending[bi] -= 1
            if self._pending[bi] == 0:
                b: Bucket = self.buckets[bi]
                w = self.comm.all_reduce_sum_(self.flat_g[b.start:b.start + b.numel], async_op=True)
                self._works.append((bi, w))
                self.bucket_launch_order.append(bi)
        return hook

    def _arm(self) -> None:
        self._pending = [len(b.entries) for b in self.buckets]
        self._works.clear()
        self.bucket_launch_order.clear()
        self.flat_g.zero_()

    def _sgd(self, lr: float) -> None:
        for _, w in self._works:
            if w is not None:
                w.wait()
        g, p, m, c = self.flat_g, self.flat_p, self.flat_m, self.sgd
        if self.comm.world_size > 1:
            g.mul_(1.0 / self.comm.world_size)
        if c.weight_decay:
            g.add_(p, alpha=c.weight_decay)
        if c.momentum:
            m.mul_(c.momentum).add_(g)
            if c.nesterov:
                g.add_(m, alpha=c.momentum)
            else:
                g.copy_(m)
        p.add_(g, alpha=-lr)

    # ---------------------------------------------------------------- steps
    def _inputs(self, batch: Batch) -> Tuple[torch.Tensor, torch.Tensor]:
        x = augment_cpu(batch.u8, batch.aug)
        return (x.to(self.device, non_blocking=True), batch.labels.to(self.device, non_blocking=True))

    def _forward(self, x):
        if self.autocast_bf16:
            with torch.autocast(self.device.type, dtype=torch.bfloat16):
                return self.model(x).float()
        return self.model(x)

    def _account(self, logits, y, loss_sum, valid: Optional[torch.Tensor]) -> None:
        k = min(5, logits.shape[1])
        top = logits.topk(k, 1).indices.eq(y.view(-1, 1))
        if valid is not None:
            top = top & valid.view(-1, 1).to(top.device)
            n = float(valid.sum())
        else:
            n = float(y.numel())
        self._m += torch.tensor([float(loss_sum), float(top[:, :1].sum()), float(top.sum()), n],
                                dtype=torch.float64)

    def train_step(self, batch: Batch, lr: float) -> None:
        self.model.train()
        x, y = self._inputs(batch)
        self._arm()
        logits = self._forward(x)
        loss = F.cross_entropy(logits, y)
        loss.backward()
        self._sgd(lr)
        self.step_count += 1
        self._account(logits.detach(), y, loss.detach() * y.numel(), None)

    @torch.no_grad()
    def eval_step(self, batch: Batch) -> None:
        self.model.eval()
        x, y = self._inputs(batch)
        logits = self._forward(x)
        valid = batch.valid.to(self.device)
        losses = F.cross_entropy(logits, y, reduction='none')
        self._account(logits, y, (losses * valid).sum(), valid)

    def read_metrics(self, reset: bool = True) -> List[float]:
        out = self._m.tolist()
        if reset:
            self._m.zero_()
        return out

    def synchronize(self) -> None:
        if self.device.type == 'cuda':
            torch.cuda.synchronize(self.device)

    # ---------------------------------------------------------------- state
    def model_state(self) -> Dict[str, torch.Tensor]:
        return {k: v.detach().clone() for k, v in self.model.state_dict().items()}

    def load_model_state(self, sd: Dict[str, torch.Tensor]) -> None:
        own = self.model.state_dict()
        for k, v in sd.items():
            own[k].copy_(v)

    def optimizer_state(self) -> Dict[str, torch.Tensor]:
        names = [e.name for b in self.buckets for e in b.entries]
        out = {}
        for b in self.buckets:
            for e in b.entries:
                out[e.name] = self.flat_m[e.offset:e.offset + e.numel].detach().clone().cpu()
        return {'momentum': out, 'step': self.step_count, 'order': names}

    def load_optimizer_state(self, st) -> None:
        for b in self.buckets:
            for e in b.entries:
                self.flat_m[e.offset:e.offset + e.numel].copy_(st['momentum'][e.name].view(-1))
        self.step_count = int(st.get('step', 0))

    def weights_fingerprint(self) -> float:
        return float(self.flat_p.double().sum())

    def close(self) -> None:
        pass
