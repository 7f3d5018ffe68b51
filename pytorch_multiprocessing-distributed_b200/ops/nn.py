"""Python face of the fused BatchNorm / head / loss / optimiser / augmentation kernels
(csrc/bn.cu, head.cu, optim.cu, data_aug.cu).  For tests and interactive use; the engine calls the
same kernels from C++."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr

_F = C.c_float
_LL = C.c_longlong


def _ptr_table(tensors: Sequence[torch.Tensor]) -> torch.Tensor:
    return torch.tensor([t.data_ptr() for t in tensors], dtype=torch.int64, device=tensors[0].device)


def bn_stats(y: torch.Tensor, stats: Optional[torch.Tensor] = None) -> torch.Tensor:
    Cc = y.shape[-1]
    rows = y.numel() // Cc
    if stats is None:
        stats = torch.zeros(2 * Cc, dtype=torch.float32, device=y.device)
    check(_lib.lib().pmd_bn_stats(ptr(y), _LL(rows), C.c_int(Cc), ptr(stats), stream_ptr()))
    return stats


def bn_finalize(stats_ranks: Sequence[torch.Tensor], count_per_rank: float, gamma, beta, running_mean, running_var,
                num_batches, momentum: float = 0.1, eps: float = 1e-5):
    Cc = gamma.numel()
    dev = gamma.device
    scale, shift, mean, invstd = (torch.empty(Cc, dtype=torch.float32, device=dev) for _ in range(4))
    table = _ptr_table(stats_ranks)
    check(_lib.lib().pmd_bn_finalize(ptr(table), C.c_int(len(stats_ranks)), _F(count_per_rank), C.c_int(Cc), ptr(gamma),
                                     ptr(beta), ptr(running_mean), ptr(running_var), ptr(num_batches), _F(momentum),
                                     _F(eps), ptr(scale), ptr(shift), ptr(mean), ptr(invstd), stream_ptr()))
    return scale, shift, mean, invstd


def bn_eval_affine(gamma, beta, running_mean, running_var, eps: float = 1e-5):
    Cc = gamma.numel()
    scale, shift = (torch.empty(Cc, dtype=torch.float32, device=gamma.device) for _ in range(2))
    check(_lib.lib().pmd_bn_eval_affine(C.c_int(Cc), ptr(gamma), ptr(beta), ptr(running_mean), ptr(running_var), _F(eps),
                                        ptr(scale), ptr(shift), stream_ptr()))
    return scale, shift


def bn_apply(y, scale, shift, relu: bool, res=None, res_scale=None, res_shift=None) -> torch.Tensor:
    Cc = y.shape[-1]
    z = torch.empty_like(y)
    check(_lib.lib().pmd_bn_apply(ptr(y), ptr(scale), ptr(shift), ptr(res), ptr(res_scale), ptr(res_shift),
                                  C.c_int(int(relu)), ptr(z), _LL(y.numel() // Cc), C.c_int(Cc), stream_ptr()))
    return z


def bn_bwd_reduce(dz, z, y, mean, invstd, relu: bool, sums: Optional[torch.Tensor] = None) -> torch.Tensor:
    Cc = y.shape[-1]
    if sums is None:
        sums = torch.zeros(2 * Cc, dtype=torch.float32, device=y.device)
    check(_lib.lib().pmd_bn_bwd_reduce(ptr(dz), ptr(z), ptr(y), ptr(mean), ptr(invstd), C.c_int(int(relu)), ptr(sums),
                                       _LL(y.numel() // Cc), C.c_int(Cc), stream_ptr()))
    return sums


def bn_bwd_apply(dz, z, y, mean, invstd, gamma, sums_ranks: Sequence[torch.Tensor], my_rank: int, count_total: float,
                 relu: bool, dgamma=None, dbeta=None, g_out=None) -> torch.Tensor:
    Cc = y.shape[-1]
    dy = torch.empty_like(y)
    table = _ptr_table(sums_ranks)
    check(_lib.lib().pmd_bn_bwd_apply(ptr(dz), ptr(z), ptr(y), ptr(mean), ptr(invstd), ptr(gamma), ptr(table),
                                      C.c_int(len(sums_ranks)), ptr(sums_ranks[my_rank]), _F(count_total), C.c_int(int(relu)),
                                      ptr(dy), ptr(dgamma), ptr(dbeta), ptr(g_out), _LL(y.numel() // Cc), C.c_int(Cc),
                                      stream_ptr()))
    return dy


def pool_fc_fwd(z: torch.Tensor, w: torch.Tensor, b: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    N, H, W, Cc = z.shape
    classes = w.shape[0]
    feat = torch.empty((N, Cc), dtype=torch.float32, device=z.device)
    logits = torch.empty((N, classes), dtype=torch.float32, device=z.device)
    check(_lib.lib().pmd_pool_fc_fwd(ptr(z), C.c_int(N), C.c_int(H * W), C.c_int(Cc), ptr(w), ptr(b), C.c_int(classes),
                                     ptr(feat), ptr(logits), stream_ptr()))
    return feat, logits


def ce_fwd_bwd(logits, labels, valid=None, grad_scale: float = 1.0, want_grad: bool = True, metrics=None):
    N, classes = logits.shape
    if metrics is None:
        metrics = torch.zeros(4, dtype=torch.float64, device=logits.device)
    dl = torch.empty_like(logits) if want_grad else None
    check(_lib.lib().pmd_ce_fwd_bwd(ptr(logits), ptr(labels), ptr(valid), C.c_int(N), C.c_int(classes), _F(grad_scale),
                                    ptr(dl), ptr(metrics), stream_ptr()))
    return dl, metrics


def pool_fc_bwd(dlogits, feat, w, hw: Tuple[int, int], dw=None, db=None):
    N, classes = dlogits.shape
    Cc = feat.shape[1]
    if dw is None:
        dw = torch.zeros_like(w)
    if db is None:
        db = torch.zeros(classes, dtype=torch.float32, device=w.device)
    dz = torch.empty((N, hw[0], hw[1], Cc), dtype=torch.bfloat16, device=w.device)
    check(_lib.lib().pmd_pool_fc_bwd(ptr(dlogits), ptr(feat), ptr(w), C.c_int(N), C.c_int(hw[0] * hw[1]), C.c_int(Cc),
                                     C.c_int(classes), ptr(dw), ptr(db), ptr(dz), stream_ptr()))
    return dw, db, dz


def sgd_step(param, grad, mom, param_bf16, lr, momentum=0.9, weight_decay=1e-4, nesterov=True, grad_scale=1.0):
    check(_lib.lib().pmd_sgd_step(ptr(param), ptr(grad), ptr(mom), ptr(param_bf16), _LL(param.numel()), _F(lr),
                                  _F(momentum), _F(weight_decay), C.c_int(int(nesterov)), _F(grad_scale), stream_ptr()))


def augment(u8: torch.Tensor, aug: torch.Tensor, pad: int = 8, c_out: int = 4) -> torch.Tensor:
    N, H, W, _ = u8.shape
    out = torch.empty((N, H, W, c_out), dtype=torch.bfloat16, device=u8.device)
    check(_lib.lib().pmd_augment(ptr(u8), ptr(aug), C.c_int(N), C.c_int(H), C.c_int(W), C.c_int(pad), C.c_int(c_out),
                                 ptr(out), stream_ptr()))
    return out
