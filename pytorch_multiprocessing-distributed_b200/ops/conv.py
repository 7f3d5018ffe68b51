"""Python face of the convolution kernels (csrc/conv_tc.cu, csrc/conv_direct.cu).

Tensors are NHWC bf16 CUDA tensors; weights KRSC (fprop / wgrad) or RSCK (dgrad) bf16; dW is fp32
KRSC.  These wrappers exist for tests, probes and interactive use -- the training engine drives the
same kernels from C++ (csrc/engine.cu) without going through Python.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import PmdConvShape, check, ptr, stream_ptr


def out_dim(n: int, R: int, stride: int) -> int:
    return (n + 2 * (R // 2) - R) // stride + 1


def _shape(x_shape, K: int, R: int, stride: int) -> PmdConvShape:
    N, H, W, Cc = x_shape
    return PmdConvShape(N, H, W, Cc, K, R, stride)


def krsc_from_torch(w: torch.Tensor) -> torch.Tensor:
    """[K, C, R, S] (torch layout) -> contiguous [K, R, S, C]."""
    return w.permute(0, 2, 3, 1).contiguous()


def rsck_from_torch(w: torch.Tensor) -> torch.Tensor:
    """[K, C, R, S] -> contiguous [R, S, C, K] (dgrad operand)."""
    return w.permute(2, 3, 1, 0).contiguous()


def conv_fprop(x: torch.Tensor, w_krsc: torch.Tensor, stride: int = 1, stats: Optional[torch.Tensor] = None,
               direct: bool = False, x_channels: Optional[int] = None) -> torch.Tensor:
    N, H, W, Cs = x.shape
    K, R = w_krsc.shape[0], w_krsc.shape[1]
    Cc = x_channels if x_channels is not None else Cs
    y = torch.empty((N, out_dim(H, R, stride), out_dim(W, R, stride), K), dtype=torch.bfloat16, device=x.device)
    s = PmdConvShape(N, H, W, Cc, K, R, stride)
    if direct:
        check(_lib.lib().pmd_conv_fprop_direct(C.byref(s), C.c_int(Cs), ptr(x), ptr(w_krsc), ptr(y), stream_ptr()))
    else:
        check(_lib.lib().pmd_conv_fprop(C.byref(s), ptr(x), ptr(w_krsc), ptr(y), ptr(stats), stream_ptr()))
    return y


def conv_dgrad(dy: torch.Tensor, w: torch.Tensor, x_shape: Tuple[int, int, int, int], R: int, stride: int = 1,
               second: Optional[Tuple[torch.Tensor, torch.Tensor, int]] = None,
               addend: Optional[torch.Tensor] = None, direct: bool = False, krsc: bool = False) -> torch.Tensor:
    """``w`` is RSCK for the tensor-core path (KRSC with ``krsc=True``: MN-major B operand, no transposed copy) and
    KRSC for ``direct``.  ``second`` = (dy2, w2, R2) fuses a second convolution over the same input (projection
    shortcut); its weight uses the same layout as ``w``."""
    K = dy.shape[-1]
    dx = torch.empty(x_shape, dtype=torch.bfloat16, device=dy.device)
    s = _shape(x_shape, K, R, stride)
    if direct:
        check(_lib.lib().pmd_conv_dgrad_direct(C.byref(s), ptr(dy), ptr(w), ptr(dx), stream_ptr()))
        return dx
    fn = _lib.lib().pmd_conv_dgrad_krsc if krsc else _lib.lib().pmd_conv_dgrad
    if second is not None:
        dy2, w2, R2 = second
        s2 = _shape(x_shape, dy2.shape[-1], R2, stride)
        check(fn(C.byref(s), ptr(dy), ptr(w), C.byref(s2), ptr(dy2), ptr(w2), ptr(addend),
                                        ptr(dx), stream_ptr()))
    else:
        check(fn(C.byref(s), ptr(dy), ptr(w), None, None, None, ptr(addend), ptr(dx), stream_ptr()))
    return dx


def conv_dgrad_bn_reduce(dy: torch.Tensor, w_krsc: torch.Tensor, y_prev: torch.Tensor, scale: torch.Tensor,
                         shift: torch.Tensor, mean: torch.Tensor, invstd: torch.Tensor,
                         sums: torch.Tensor, addend: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Stride-1 dgrad whose epilogue applies the ReLU mask of the layer below and accumulates that layer's
    BatchNorm-backward sums (``pmd_conv_dgrad_bn_reduce``).  ``y_prev`` is that layer's pre-BN output [N,H,W,C]."""
    K, R = w_krsc.shape[0], w_krsc.shape[1]
    dz = torch.empty_like(y_prev)
    s = _shape(tuple(y_prev.shape), K, R, 1)
    check(_lib.lib().pmd_conv_dgrad_bn_reduce_addend(C.byref(s), ptr(dy), ptr(w_krsc), ptr(addend), ptr(y_prev), ptr(scale),
                                                     ptr(shift), ptr(mean), ptr(invstd), ptr(sums), ptr(dz), stream_ptr()))
    return dz


def conv_wgrad(x: torch.Tensor, dy: torch.Tensor, R: int, stride: int = 1, split_k: int = 0,
               out: Optional[torch.Tensor] = None, direct: bool = False,
               x_channels: Optional[int] = None) -> torch.Tensor:
    N, H, W, Cs = x.shape
    Cc = x_channels if x_channels is not None else Cs
    K = dy.shape[-1]
    if out is None:
        out = torch.zeros((K, R, R, Cc), dtype=torch.float32, device=x.device)
    s = PmdConvShape(N, H, W, Cc, K, R, stride)
    if direct:
        check(_lib.lib().pmd_conv_wgrad_direct(C.byref(s), C.c_int(Cs), ptr(x), ptr(dy), ptr(out), stream_ptr()))
    else:
        check(_lib.lib().pmd_conv_wgrad(C.byref(s), ptr(x), ptr(dy), ptr(out), C.c_int(split_k), stream_ptr()))
    return out
