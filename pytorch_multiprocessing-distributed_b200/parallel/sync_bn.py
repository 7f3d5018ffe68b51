"""Cross-replica BatchNorm for the torch (CPU / oracle) path.

Parity target: ``SyncBatchNorm.convert_sync_batchnorm(model)`` (ref:main.py:43) and the autograd
function behind it (torch:nn/modules/_functions.py:39-179).  torch's SyncBatchNorm refuses CPU
tensors (torch:nn/parallel/distributed.py:2259-2266), which is why the reference cannot run its
own plumbing config without a GPU; this version is transport-agnostic (any ``Comm``), so the
2-process gloo configuration keeps synchronised statistics, and it doubles as the multi-rank
oracle for the fused sm_100a kernels (csrc/bn.cu + the push-exchange kernels of csrc/comm.cu).

Math (identical to torch): forward all-reduces [sum(x), sum(x^2), n] per channel -> global mean,
biased var for normalisation, unbiased var for ``running_var`` with momentum 0.1; backward
all-reduces [sum(dy), sum(dy * (x - mean))] and applies
``dx = (dy - sum_dy/N - (x - mean) * invstd^2 * sum_dy_xmu/N) * invstd * gamma``.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .comm import Comm


class _SyncBNFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, eps, momentum, comm: Comm):
        c = x.shape[1]
        red = [0] + list(range(2, x.dim()))
        n_local = x.numel() // c
        xf = x.float()
        stats = torch.cat([xf.sum(red), (xf * xf).sum(red),
                           torch.tensor([float(n_local)], device=x.device)])
        comm.all_reduce_sum_(stats)
        n = stats[-1]
        mean = stats[:c] / n
        var = (stats[c:2 * c] / n - mean * mean).clamp_min_(0.0)
        invstd = torch.rsqrt(var + eps)
        if running_mean is not None:
            with torch.no_grad():
                unbiased = var * (n / (n - 1).clamp_min(1.0))
                running_mean.mul_(1 - momentum).add_(mean.to(running_mean.dtype), alpha=momentum)
                running_var.mul_(1 - momentum).add_(unbiased.to(running_var.dtype), alpha=momentum)
        shape = [1, c] + [1] * (x.dim() - 2)
        xhat = (xf - mean.view(shape)) * invstd.view(shape)
        y = xhat * weight.float().view(shape) + bias.float().view(shape)
        ctx.save_for_backward(xhat, weight, invstd)
        ctx.comm, ctx.n_total = comm, n
        return y.to(x.dtype)

    @staticmethod
    def backward(ctx, dy):
        xhat, weight, invstd = ctx.saved_tensors
        c = dy.shape[1]
        red = [0] + list(range(2, dy.dim()))
        shape = [1, c] + [1] * (dy.dim() - 2)
        dyf = dy.float()
        sum_dy = dyf.sum(red)
        sum_dy_xhat = (dyf * xhat).sum(red)
        dweight, dbias = sum_dy_xhat.clone(), sum_dy.clone()      # local, DDP averages them later
        packed = torch.cat([sum_dy, sum_dy_xhat])
        ctx.comm.all_reduce_sum_(packed)
        n = ctx.n_total
        mean_dy = (packed[:c] / n).view(shape)
        mean_dy_xhat = (packed[c:] / n).view(shape)
        dx = (dyf - mean_dy - xhat * mean_dy_xhat) * (invstd * weight.float()).view(shape)
        return dx.to(dy.dtype), dweight.to(weight.dtype), dbias.to(weight.dtype), None, None, None, None, None


class CommSyncBatchNorm(nn.modules.batchnorm._BatchNorm):
    """Drop-in for nn.BatchNorm2d with statistics shared over ``comm`` while training."""

    def __init__(self, num_features: int, comm: Comm, eps: float = 1e-5, momentum: float = 0.1) -> None:
        super().__init__(num_features, eps, momentum, True, True)
        self.comm = comm

    def _check_input_dim(self, x):
        if x.dim() < 2:
            raise ValueError('expected at least 2D input')

    def forward(self, x):
        if not self.training or self.comm.world_size == 1:
            return super().forward(x)
        if self.num_batches_tracked is not None:
            self.num_batches_tracked.add_(1)
        return _SyncBNFn.apply(x, self.weight, self.bias, self.running_mean, self.running_var,
                               self.eps, self.momentum, self.comm)


def convert_sync_batchnorm(module: nn.Module, comm: Comm) -> nn.Module:
    """Recursive swap that shares parameters and buffers with the originals, like
    ``torch.nn.SyncBatchNorm.convert_sync_batchnorm`` (torch:nn/modules/batchnorm.py:845-905);
    state_dict keys are unchanged."""
    out = module
    if isinstance(module, nn.modules.batchnorm._BatchNorm) and not isinstance(module, CommSyncBatchNorm):
        out = CommSyncBatchNorm(module.num_features, comm, module.eps, module.momentum)
        with torch.no_grad():
            out.weight, out.bias = module.weight, module.bias
        out.running_mean, out.running_var = module.running_mean, module.running_var
        out.num_batches_tracked = module.num_batches_tracked
        out.training = module.training
    for name, child in module.named_children():
        out.add_module(name, convert_sync_batchnorm(child, comm))
    return out
