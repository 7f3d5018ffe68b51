"""GPU numerics cases shared by tests/test_gpu_*.py and tools/gpu_probe.py.

Every case compares one native sm_100a kernel against a plain PyTorch fp32 reference of the same op
on bf16-rounded inputs and returns a dict with error statistics and a pass flag.
"""
from __future__ import annotations

import json
import math
from typing import Callable, Dict

import torch
import torch.nn.functional as F

CASES: Dict[str, Callable[[], dict]] = {}
# the oracles are plain fp32 PyTorch: keep cuDNN / cuBLAS from silently computing them in TF32
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def case(name):
    def deco(fn):
        CASES[name] = fn
        return fn
    return deco


def _cmp(got: torch.Tensor, ref: torch.Tensor, rel_tol: float) -> dict:
    got, ref = got.float(), ref.float()
    scale = ref.abs().max().item() + 1e-12
    err = (got - ref).abs()
    out = {'max_abs_err': err.max().item(), 'ref_max': scale, 'rel_to_max': err.max().item() / scale,
           'mean_abs_err': err.mean().item(), 'nan': bool(torch.isnan(got).any().item())}
    out['ok'] = (not out['nan']) and out['rel_to_max'] <= rel_tol
    return out


def _rand_bf16(*shape, scale=1.0, seed=0):
    g = torch.Generator(device='cuda').manual_seed(seed)
    return (torch.randn(*shape, device='cuda', generator=g) * scale).to(torch.bfloat16)


def _nchw(x_nhwc):
    return x_nhwc.float().permute(0, 3, 1, 2).contiguous()


def _conv_ref(x_nhwc, w_kcrs, stride):
    R = w_kcrs.shape[-1]
    y = F.conv2d(_nchw(x_nhwc), w_kcrs.float(), stride=stride, padding=R // 2)
    return y.permute(0, 2, 3, 1).contiguous()


# ------------------------------------------------------------------ fprop
def _fprop_case(N, H, C, K, R, stride, stats=True, direct=False, W=None):
    from pmdist.ops import conv as cv
    W = W or H
    x = _rand_bf16(N, H, W, C, seed=1)
    w = _rand_bf16(K, C, R, R, scale=1.0 / math.sqrt(C * R * R), seed=2)
    st = torch.zeros(2 * K, dtype=torch.float32, device='cuda') if (stats and not direct) else None
    y = cv.conv_fprop(x, cv.krsc_from_torch(w), stride, st, direct=direct)
    torch.cuda.synchronize()
    ref = _conv_ref(x, w, stride)
    res = _cmp(y, ref, 2e-2)
    if st is not None:
        yb = y.float()
        s1, s2 = yb.sum((0, 1, 2)), (yb * yb).sum((0, 1, 2))
        r1 = _cmp(st[:K], s1, 1e-3)
        r2 = _cmp(st[K:], s2, 1e-3)
        res['stats_sum_rel'] = r1['rel_to_max']
        res['stats_sq_rel'] = r2['rel_to_max']
        res['ok'] = res['ok'] and r1['ok'] and r2['ok']
    return res


for _name, _args in {
    'fprop_3x3_s1_c64_k64_h32': (4, 32, 64, 64, 3, 1),
    'fprop_3x3_s1_c128_k128_h16': (4, 16, 128, 128, 3, 1),
    'fprop_3x3_s2_c64_k128_h32': (4, 32, 64, 128, 3, 2),
    'fprop_1x1_s2_c64_k128_h32': (4, 32, 64, 128, 1, 2),
    'fprop_1x1_s1_c256_k64_h8': (8, 8, 256, 64, 1, 1),
    'fprop_3x3_s1_c512_k512_h4': (16, 4, 512, 512, 3, 1),
    'fprop_3x3_s1_c64_k64_h32_n3_tail': (3, 32, 64, 64, 3, 1),
    'fprop_3x3_s2_c256_k512_h8_n3': (3, 8, 256, 512, 3, 2),
    'fprop_3x3_s1_c64_k64_h28_imagenetish': (2, 28, 64, 64, 3, 1),
    # weight-resident persistent kernel with several tiles per CTA (96 images -> 768 tiles on 148 SMs), 64x64 maps
    'fprop_3x3_s1_c64_k64_h32_n96_resident': (96, 32, 64, 64, 3, 1),
    'fprop_3x3_s1_c64_k64_h64_n5_resident': (5, 64, 64, 64, 3, 1),
}.items():
    CASES[_name] = (lambda a=_args: _fprop_case(*a))
CASES['fprop_direct_stem_c3'] = lambda: _fprop_case(4, 32, 3, 64, 3, 1, direct=True)
CASES['fprop_direct_c16_s2'] = lambda: _fprop_case(2, 8, 16, 24, 3, 2, direct=True)


# ------------------------------------------------------------------ dgrad
def _dgrad_ref(dy_nhwc, w_kcrs, x_shape_nhwc, stride):
    N, H, W, C = x_shape_nhwc
    R = w_kcrs.shape[-1]
    dx = torch.nn.grad.conv2d_input((N, C, H, W), w_kcrs.float(), _nchw(dy_nhwc), stride=stride, padding=R // 2)
    return dx.permute(0, 2, 3, 1).contiguous()


def _dgrad_case(N, H, C, K, R, stride, shortcut_K=0, addend=False, direct=False, krsc=False):
    from pmdist.ops import conv as cv
    Ho = cv.out_dim(H, R, stride)
    dy = _rand_bf16(N, Ho, Ho, K, seed=3)
    w = _rand_bf16(K, C, R, R, scale=1.0 / math.sqrt(K * R * R), seed=4)
    ref = _dgrad_ref(dy, w, (N, H, H, C), stride)
    second = None
    if shortcut_K:
        dy2 = _rand_bf16(N, Ho, Ho, shortcut_K, seed=5)
        w2 = _rand_bf16(shortcut_K, C, 1, 1, scale=1.0 / math.sqrt(shortcut_K), seed=6)
        ref = ref + _dgrad_ref(dy2, w2, (N, H, H, C), stride)
        second = (dy2, cv.krsc_from_torch(w2) if krsc else cv.rsck_from_torch(w2), 1)
    add = None
    if addend:
        add = _rand_bf16(N, H, H, C, seed=7)
        ref = ref + add.float()
    if direct:
        dx = cv.conv_dgrad(dy, cv.krsc_from_torch(w), (N, H, H, C), R, stride, direct=True)
    elif krsc:
        dx = cv.conv_dgrad(dy, cv.krsc_from_torch(w), (N, H, H, C), R, stride, second=second, addend=add, krsc=True)
    else:
        dx = cv.conv_dgrad(dy, cv.rsck_from_torch(w), (N, H, H, C), R, stride, second=second, addend=add)
    torch.cuda.synchronize()
    return _cmp(dx, ref, 2e-2)


for _name, _args in {
    'dgrad_3x3_s1_c64_k64_h32': (4, 32, 64, 64, 3, 1),
    'dgrad_3x3_s1_c128_k128_h16_addend': (4, 16, 128, 128, 3, 1, 0, True),
    'dgrad_3x3_s2_c64_k128_h32': (4, 32, 64, 128, 3, 2),
    'dgrad_3x3_s2_c64_k128_h32_shortcut': (4, 32, 64, 128, 3, 2, 128),
    'dgrad_1x1_s2_c64_k128_h32': (4, 32, 64, 128, 1, 2),
    'dgrad_1x1_s1_c256_k64_h8_shortcut': (8, 8, 256, 64, 1, 1, 256),
    'dgrad_3x3_s1_c512_k512_h4': (16, 4, 512, 512, 3, 1),
    'dgrad_3x3_s2_c256_k512_h8_n3_shortcut': (3, 8, 256, 512, 3, 2, 512),
}.items():
    CASES[_name] = (lambda a=_args: _dgrad_case(*a))
CASES['dgrad_direct_c16_s2'] = lambda: _dgrad_case(2, 8, 16, 24, 3, 2, direct=True)
# dgrad straight from the KRSC weights (MN-major B operand): the path the engine uses
for _name, _args in {
    'dgrad_krsc_3x3_s1_c64_k64_h32': (4, 32, 64, 64, 3, 1),
    'dgrad_krsc_3x3_s1_c128_k128_h16_addend': (4, 16, 128, 128, 3, 1, 0, True),
    'dgrad_krsc_3x3_s2_c64_k128_h32_shortcut': (4, 32, 64, 128, 3, 2, 128),
    'dgrad_krsc_1x1_s2_c64_k128_h32': (4, 32, 64, 128, 1, 2),
    'dgrad_krsc_1x1_s1_c256_k64_h8_shortcut': (8, 8, 256, 64, 1, 1, 256),
    'dgrad_krsc_3x3_s1_c512_k512_h4': (16, 4, 512, 512, 3, 1),
    'dgrad_krsc_3x3_s2_c256_k512_h8_n3_shortcut': (3, 8, 256, 512, 3, 2, 512),
    'dgrad_krsc_3x3_s1_c64_k64_h32_n96_addend_resident': (96, 32, 64, 64, 3, 1, 0, True),
    'dgrad_krsc_3x3_s1_c64_k64_h16_n7_resident': (7, 16, 64, 64, 3, 1),
}.items():
    CASES[_name] = (lambda a=_args: _dgrad_case(*a[:6], shortcut_K=(a[6] if len(a) > 6 else 0),
                                                addend=(a[7] if len(a) > 7 else False), krsc=True))


def _dgrad_bn_reduce_case(N, H, C, K, R, addend=False):
    """dgrad + BatchNorm-backward reduction in one kernel vs plain PyTorch fp32: masked dz and both channel sums."""
    from pmdist.ops import conv as cv
    dy = _rand_bf16(N, H, H, K, seed=11)
    w = _rand_bf16(K, C, R, R, scale=1.0 / math.sqrt(K * R * R), seed=12)
    y_prev = _rand_bf16(N, H, H, C, seed=13)
    g = torch.Generator(device='cuda').manual_seed(14)
    scale = torch.rand(C, device='cuda', generator=g) + 0.5
    shift = torch.randn(C, device='cuda', generator=g) * 0.3
    mean = torch.randn(C, device='cuda', generator=g) * 0.1
    invstd = torch.rand(C, device='cuda', generator=g) + 0.5
    sums = torch.zeros(2 * C, device='cuda')
    add = _rand_bf16(N, H, H, C, seed=15) if addend else None
    dz = cv.conv_dgrad_bn_reduce(dy, cv.krsc_from_torch(w), y_prev, scale, shift, mean, invstd, sums, addend=add)
    torch.cuda.synchronize()
    ref = _dgrad_ref(dy, w, (N, H, H, C), 1)
    if addend:
        ref = ref + add.float()
    yf = y_prev.float()
    mask = (yf * scale + shift) > 0
    ref = torch.where(mask, ref, torch.zeros_like(ref))
    res = _cmp(dz, ref, 2e-2)
    dzf = dz.float()                      # the sums are defined over the bf16 tensor that was written
    s1 = dzf.sum((0, 1, 2))
    s2 = (dzf * (yf - mean) * invstd).sum((0, 1, 2))
    r1, r2 = _cmp(sums[:C], s1, 2e-3), _cmp(sums[C:], s2, 2e-3)
    res['sum_g_rel'], res['sum_gx_rel'] = r1['rel_to_max'], r2['rel_to_max']
    res['ok'] = res['ok'] and r1['ok'] and r2['ok']
    return res


for _name, _args in {
    'dgrad_bn_reduce_3x3_c64_k64_h32': (4, 32, 64, 64, 3),
    'dgrad_bn_reduce_3x3_c128_k128_h16_n3': (3, 16, 128, 128, 3),
    'dgrad_bn_reduce_3x3_c256_k256_h8': (8, 8, 256, 256, 3),
    'dgrad_bn_reduce_1x1_c64_k256_h8': (8, 8, 64, 256, 1),
    'dgrad_bn_reduce_3x3_c512_k512_h4': (16, 4, 512, 512, 3),
}.items():
    CASES[_name] = (lambda a=_args: _dgrad_bn_reduce_case(*a))


def _dgrad_block_bn_reduce_case(N, H, C, K, R, stride, shortcut_K=0, addend=False, nb=1):
    """Block dgrad (+ fused projection dgrad / identity addend) that also masks its output with the previous block's
    ReLU bits and accumulates that block's BatchNorm-backward sums (1 or 2 branches) -- vs plain PyTorch fp32."""
    import ctypes as C_
    from pmdist.ops import _lib, conv as cv
    from pmdist.ops._lib import PmdConvShape, check, ptr, stream_ptr
    Ho = cv.out_dim(H, R, stride)
    dy = _rand_bf16(N, Ho, Ho, K, seed=21)
    w = _rand_bf16(K, C, R, R, scale=1.0 / math.sqrt(K * R * R), seed=22)
    ref = _dgrad_ref(dy, w, (N, H, H, C), stride)
    s2 = dy2 = w2k = None
    if shortcut_K:
        dy2 = _rand_bf16(N, Ho, Ho, shortcut_K, seed=23)
        w2 = _rand_bf16(shortcut_K, C, 1, 1, scale=1.0 / math.sqrt(shortcut_K), seed=24)
        ref = ref + _dgrad_ref(dy2, w2, (N, H, H, C), stride)
        w2k = cv.krsc_from_torch(w2)
        s2 = PmdConvShape(N, H, H, C, shortcut_K, 1, stride)
    add = None
    if addend:
        add = _rand_bf16(N, H, H, C, seed=25)
        ref = ref + add.float()
    g = torch.Generator(device='cuda').manual_seed(26)
    keep = torch.rand(N, H, H, C, device='cuda', generator=g) > 0.4                      # the block output's ReLU mask
    bits = ((keep.view(-1, 8).to(torch.int32) << torch.arange(8, device='cuda', dtype=torch.int32)).sum(1)).to(torch.uint8)
    ys = [_rand_bf16(N, H, H, C, seed=27 + i) for i in range(nb)]
    means = [torch.randn(C, device='cuda', generator=g) * 0.1 for _ in range(nb)]
    invs = [torch.rand(C, device='cuda', generator=g) + 0.5 for _ in range(nb)]
    sums = [torch.zeros(2 * C, device='cuda') for _ in range(nb)]
    dx = torch.empty(N, H, H, C, dtype=torch.bfloat16, device='cuda')
    s = PmdConvShape(N, H, H, C, K, R, stride)
    wk = cv.krsc_from_torch(w)
    check(_lib.lib().pmd_conv_dgrad_block_bn_reduce(
        C_.byref(s), ptr(dy), ptr(wk), C_.byref(s2) if s2 is not None else None, ptr(dy2), ptr(w2k), ptr(add), ptr(bits),
        C_.c_int(nb), ptr(ys[0]), ptr(means[0]), ptr(invs[0]), ptr(sums[0]), ptr(ys[1]) if nb == 2 else None,
        ptr(means[1]) if nb == 2 else None, ptr(invs[1]) if nb == 2 else None, ptr(sums[1]) if nb == 2 else None, ptr(dx),
        stream_ptr()))
    torch.cuda.synchronize()
    ref = torch.where(keep, ref, torch.zeros_like(ref))
    res = _cmp(dx, ref, 2e-2)
    gf = dx.float()
    for i in range(nb):
        r1 = _cmp(sums[i][:C], gf.sum((0, 1, 2)), 2e-3)
        r2 = _cmp(sums[i][C:], (gf * (ys[i].float() - means[i]) * invs[i]).sum((0, 1, 2)), 2e-3)
        res[f'sum_g_rel_{i}'], res[f'sum_gx_rel_{i}'] = r1['rel_to_max'], r2['rel_to_max']
        res['ok'] = res['ok'] and r1['ok'] and r2['ok']
    return res


CASES['dgrad_block_bn_reduce_s2_shortcut_nb2_c128_k256'] = lambda: _dgrad_block_bn_reduce_case(4, 16, 128, 256, 3, 2, 256, nb=2)
CASES['dgrad_block_bn_reduce_s2_shortcut_nb1_c64_k128'] = lambda: _dgrad_block_bn_reduce_case(4, 32, 64, 128, 3, 2, 128, nb=1)
CASES['dgrad_block_bn_reduce_s1_addend_nb1_c128'] = lambda: _dgrad_block_bn_reduce_case(3, 16, 128, 128, 3, 1, 0, True, nb=1)
CASES['dgrad_block_bn_reduce_s2_nb2_c256_k512_n3'] = lambda: _dgrad_block_bn_reduce_case(3, 8, 256, 512, 3, 2, 512, nb=2)
CASES['bench512_dgrad_block_bn_reduce_layer2_conv1'] = lambda: _dgrad_block_bn_reduce_case(512, 32, 64, 128, 3, 2, 128, nb=1)
CASES['bench512_dgrad_block_bn_reduce_layer3_conv1'] = lambda: _dgrad_block_bn_reduce_case(512, 16, 128, 256, 3, 2, 256, nb=2)
CASES['bench512_dgrad_block_bn_reduce_layer4_conv1'] = lambda: _dgrad_block_bn_reduce_case(512, 8, 256, 512, 3, 2, 512, nb=2)


# ------------------------------------------------------------------ wgrad
def _wgrad_case(N, H, C, K, R, stride, split=0, direct=False):
    from pmdist.ops import conv as cv
    Ho = cv.out_dim(H, R, stride)
    x = _rand_bf16(N, H, H, C, seed=8)
    dy = _rand_bf16(N, Ho, Ho, K, seed=9)
    dw = cv.conv_wgrad(x, dy, R, stride, split_k=split, direct=direct)
    torch.cuda.synchronize()
    ref = torch.nn.grad.conv2d_weight(_nchw(x), (K, C, R, R), _nchw(dy), stride=stride, padding=R // 2)
    return _cmp(dw, ref.permute(0, 2, 3, 1).contiguous(), 2e-3)


for _name, _args in {
    'wgrad_3x3_s1_c64_k64_h32': (4, 32, 64, 64, 3, 1),
    'wgrad_3x3_s1_c64_k64_h32_split1': (2, 32, 64, 64, 3, 1, 1),
    'wgrad_3x3_s1_c128_k128_h16': (4, 16, 128, 128, 3, 1),
    'wgrad_3x3_s2_c64_k128_h32': (4, 32, 64, 128, 3, 2),
    'wgrad_1x1_s2_c64_k128_h32': (4, 32, 64, 128, 1, 2),
    'wgrad_1x1_s1_c256_k64_h8': (8, 8, 256, 64, 1, 1),
    'wgrad_3x3_s1_c512_k512_h4': (16, 4, 512, 512, 3, 1),
    'wgrad_3x3_s1_c256_k256_h8_n3': (3, 8, 256, 256, 3, 1),
}.items():
    CASES[_name] = (lambda a=_args: _wgrad_case(*a))
CASES['wgrad_direct_stem_c3'] = lambda: _wgrad_case(4, 32, 3, 64, 3, 1, direct=True)
CASES['wgrad_direct_generic_c16'] = lambda: _wgrad_case(2, 8, 16, 24, 3, 2, direct=True)


# ------------------------------------------------------------------ the shapes bench.py actually runs
# ResNet18-ref at 512 images per GPU (BASELINE.json config 3): every convolution layer x {fprop (+BN statistics),
# the dgrad form the engine uses for it, wgrad}.  Only at this size do the 256-wide output tiles (>= 148 CTAs), the
# 29- / 49- / 256-way split-K wgrads and the 148-CTA patch wgrad get selected.
_BENCH_LAYERS = {          # name: (H_in, C, K, R, stride)
    'stem_as_1x1': (32, 64, 64, 1, 1),
    'layer1_conv1': (32, 64, 64, 3, 1), 'layer1_conv2': (32, 64, 64, 3, 1),
    'layer2_conv1': (32, 64, 128, 3, 2), 'layer2_conv2': (16, 128, 128, 3, 1), 'layer2_shortcut': (32, 64, 128, 1, 2),
    'layer3_conv1': (16, 128, 256, 3, 2), 'layer3_conv2': (8, 256, 256, 3, 1), 'layer3_shortcut': (16, 128, 256, 1, 2),
    'layer4_conv1': (8, 256, 512, 3, 2), 'layer4_conv2': (4, 512, 512, 3, 1), 'layer4_shortcut': (8, 256, 512, 1, 2),
}
for _ln, (_h, _c, _k, _r, _s) in _BENCH_LAYERS.items():
    CASES[f'bench512_fprop_{_ln}'] = (lambda a=(512, _h, _c, _k, _r, _s): _fprop_case(*a))
    CASES[f'bench512_wgrad_{_ln}'] = (lambda a=(512, _h, _c, _k, _r, _s): _wgrad_case(*a))
# block dgrads: layer1 conv1 carries the identity shortcut as an addend, layers 2-4 fuse the projection's dgrad
CASES['bench512_dgrad_layer1_conv1_addend'] = lambda: _dgrad_case(512, 32, 64, 64, 3, 1, 0, True, krsc=True)
CASES['bench512_dgrad_layer2_conv1_shortcut'] = lambda: _dgrad_case(512, 32, 64, 128, 3, 2, 128, krsc=True)
CASES['bench512_dgrad_layer3_conv1_shortcut'] = lambda: _dgrad_case(512, 16, 128, 256, 3, 2, 256, krsc=True)
CASES['bench512_dgrad_layer4_conv1_shortcut'] = lambda: _dgrad_case(512, 8, 256, 512, 3, 2, 512, krsc=True)
# inner dgrads (conv2 of every block) also do the BatchNorm-backward reduction of bn1
CASES['bench512_dgrad_bn_reduce_addend_layer1_conv1_stem'] = lambda: _dgrad_bn_reduce_case(512, 32, 64, 64, 3, addend=True)
CASES['dgrad_bn_reduce_addend_c128_h16_n3'] = lambda: _dgrad_bn_reduce_case(3, 16, 128, 128, 3, addend=True)
CASES['bench512_dgrad_bn_reduce_layer1_conv2'] = lambda: _dgrad_bn_reduce_case(512, 32, 64, 64, 3)
CASES['bench512_dgrad_bn_reduce_layer2_conv2'] = lambda: _dgrad_bn_reduce_case(512, 16, 128, 128, 3)
CASES['bench512_dgrad_bn_reduce_layer3_conv2'] = lambda: _dgrad_bn_reduce_case(512, 8, 256, 256, 3)
CASES['bench512_dgrad_bn_reduce_layer4_conv2'] = lambda: _dgrad_bn_reduce_case(512, 4, 512, 512, 3)


def _with_env(fn, **env):
    """Runs a case with kernel switches set (the convolution builders read them per call)."""
    import os
    old = {k: os.environ.get(k) for k in env}
    os.environ.update({k: str(v) for k, v in env.items()})
    try:
        return fn()
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


# weight-resident dgrad with the fused BatchNorm-backward reduction (PMD_CONV_RESIDENT=5)
CASES['resident_dgrad_bn_reduce_layer1_conv2_n512'] = lambda: _with_env(
    lambda: _dgrad_bn_reduce_case(512, 32, 64, 64, 3), PMD_CONV_RESIDENT=5)
CASES['resident_dgrad_bn_reduce_addend_layer1_conv1_n96'] = lambda: _with_env(
    lambda: _dgrad_bn_reduce_case(96, 32, 64, 64, 3, addend=True), PMD_CONV_RESIDENT=5)
CASES['resident_dgrad_bn_reduce_h16_n7'] = lambda: _with_env(
    lambda: _dgrad_bn_reduce_case(7, 16, 64, 64, 3), PMD_CONV_RESIDENT=5)

# CTA-pair kernels (tcgen05 cta_group::2, opt-in with PMD_GEMM_2CTA=1): 128- and 256-wide tiles, K-major and MN-major B,
# with the fused BatchNorm-backward reduction, stride 1 and 2, small and bench-size grids
CASES['pair_fprop_layer2_conv2_n512'] = lambda: _with_env(lambda: _fprop_case(512, 16, 128, 128, 3, 1), PMD_GEMM_2CTA=1)
CASES['pair_fprop_layer3_conv2_n512_bn256'] = lambda: _with_env(lambda: _fprop_case(512, 8, 256, 256, 3, 1), PMD_GEMM_2CTA=1)
CASES['pair_fprop_layer4_conv1_s2_n512'] = lambda: _with_env(lambda: _fprop_case(512, 8, 256, 512, 3, 2), PMD_GEMM_2CTA=1)
CASES['pair_fprop_c128_k128_h16_n4'] = lambda: _with_env(lambda: _fprop_case(4, 16, 128, 128, 3, 1), PMD_GEMM_2CTA=1)
CASES['pair_dgrad_layer3_conv1_shortcut_n512'] = lambda: _with_env(
    lambda: _dgrad_case(512, 16, 128, 256, 3, 2, 256, krsc=True), PMD_GEMM_2CTA=1)
CASES['pair_dgrad_krsc_bn256_layer4_conv2_n512'] = lambda: _with_env(
    lambda: _dgrad_case(512, 4, 512, 512, 3, 1, krsc=True), PMD_GEMM_2CTA=1)
CASES['pair_dgrad_rsck_c128_h16_addend'] = lambda: _with_env(
    lambda: _dgrad_case(4, 16, 128, 128, 3, 1, 0, True), PMD_GEMM_2CTA=1)
CASES['pair_dgrad_bn_reduce_layer2_conv2_n512'] = lambda: _with_env(
    lambda: _dgrad_bn_reduce_case(512, 16, 128, 128, 3), PMD_GEMM_2CTA=1)
CASES['pair_dgrad_bn_reduce_deep_ring'] = lambda: _with_env(
    lambda: _dgrad_bn_reduce_case(16, 8, 256, 256, 3), PMD_GEMM_2CTA=1, PMD_GEMM2_STAGES128=4)


# ------------------------------------------------------------------ BatchNorm
@case('bn_train_forward_backward_c64')
def _bn_case(C=64, N=8, H=16, relu=True, with_res=True):
    from pmdist.ops import nn as nk
    y = _rand_bf16(N, H, H, C, seed=10) * 1.5 + 0.25
    y = y.to(torch.bfloat16)
    res = _rand_bf16(N, H, H, C, seed=11) if with_res else None
    gamma = torch.rand(C, device='cuda') + 0.5
    beta = torch.randn(C, device='cuda') * 0.1
    rm, rv = torch.zeros(C, device='cuda'), torch.ones(C, device='cuda')
    nb = torch.zeros(1, dtype=torch.int64, device='cuda')
    rows = N * H * H
    stats = nk.bn_stats(y)
    scale, shift, mean, invstd = nk.bn_finalize([stats], float(rows), gamma, beta, rm, rv, nb)
    z = nk.bn_apply(y, scale, shift, relu, res)
    dz = _rand_bf16(N, H, H, C, seed=12)
    sums = nk.bn_bwd_reduce(dz, z, y, mean, invstd, relu)
    dgamma, dbeta = torch.zeros(C, device='cuda'), torch.zeros(C, device='cuda')
    dy = nk.bn_bwd_apply(dz, z, y, mean, invstd, gamma, [sums], 0, float(rows), relu, dgamma, dbeta)
    torch.cuda.synchronize()
    # reference in fp32 autograd
    yr = y.float().requires_grad_(True)
    g_, b_ = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    rm2, rv2 = torch.zeros(C, device='cuda'), torch.ones(C, device='cuda')
    out = F.batch_norm(yr.permute(0, 3, 1, 2), rm2, rv2, g_, b_, True, 0.1, 1e-5).permute(0, 2, 3, 1)
    if with_res:
        out = out + res.float()
    if relu:
        out = F.relu(out)
    # use the kernel's own relu mask (z > 0) so bf16 rounding at the threshold cannot flip a sample
    mask = (z.float() > 0).float() if relu else torch.ones_like(out)
    pre = out if not relu else None
    gsrc = dz.float() * mask
    lin = F.batch_norm(yr.permute(0, 3, 1, 2), None, None, g_, b_, True, 0.0, 1e-5).permute(0, 2, 3, 1)
    lin.backward(gsrc)
    r = {}
    r['z'] = _cmp(z, out.detach(), 1e-2)
    r['dy'] = _cmp(dy, yr.grad, 2e-2)
    r['dgamma'] = _cmp(dgamma, g_.grad, 2e-3)
    r['dbeta'] = _cmp(dbeta, b_.grad, 2e-3)
    r['running_mean'] = _cmp(rm, rm2, 1e-3)
    r['running_var'] = _cmp(rv, rv2, 1e-3)
    r['ok'] = all(v['ok'] for v in r.values()) and int(nb.item()) == 1
    return r


CASES['bn_train_forward_backward_c512_norelu'] = lambda: _bn_case(512, 16, 4, False, False)
CASES['bn_train_forward_backward_c2048'] = lambda: _bn_case(2048, 4, 4, True, True)


@case('bn_statistics_large_mean_small_std_pivot')
def _bn_pivot_stress():
    """|mean| >> std (mean 30, std 0.05, 524,288 rows: the bench's layer1 row count): E[y^2] - E[y]^2 in fp32 cancels
    (first step, pivot 0: the variance is off by tens of percent), sums taken around the previous step's batch mean
    do not.  Oracle: torch.var_mean in fp64 over the bf16-rounded tensor.  Also: the fused finalize+apply kernel on
    pivoted sums, and the pivot rotation at the end of the backward apply."""
    import ctypes as C
    from pmdist.ops import _lib
    from pmdist.ops._lib import check, ptr, stream_ptr
    lib = _lib.lib()
    Cc, rows = 64, 524288
    g = torch.Generator(device='cuda').manual_seed(77)
    mean_c = 30.0 + torch.randn(Cc, device='cuda', generator=g)
    y = (mean_c + 0.05 * torch.randn(rows, Cc, device='cuda', generator=g)).to(torch.bfloat16).view(512, 32, 32, Cc)
    yd = y.double().view(-1, Cc)
    var_ref, mean_ref = torch.var_mean(yd, dim=0, unbiased=False)
    gamma, beta = torch.ones(Cc, device='cuda'), torch.zeros(Cc, device='cuda')
    r = {}
    pivot = torch.zeros(Cc, device='cuda')
    errs = []
    for step in range(2):
        stats = torch.zeros(2 * Cc, device='cuda')
        check(lib.pmd_bn_stats_pivot(ptr(y), C.c_longlong(rows), C.c_int(Cc), ptr(stats), ptr(pivot), stream_ptr()))
        outs = [torch.empty(Cc, device='cuda') for _ in range(4)]
        pivot_new = torch.empty(Cc, device='cuda')
        tab = torch.tensor([stats.data_ptr()], dtype=torch.int64, device='cuda')
        check(lib.pmd_bn_finalize_pivot(ptr(tab), C.c_int(1), C.c_float(rows), C.c_int(Cc), ptr(gamma), ptr(beta), None, None,
                                        None, C.c_float(0.1), C.c_float(1e-5), ptr(outs[0]), ptr(outs[1]), ptr(outs[2]),
                                        ptr(outs[3]), ptr(pivot), ptr(pivot_new), stream_ptr()))
        torch.cuda.synchronize()
        var_got = (1.0 / outs[3].double() ** 2) - 1e-5
        errs.append({'mean_abs_err': float((outs[2].double() - mean_ref).abs().max()),
                     'var_rel_err': float(((var_got - var_ref).abs() / var_ref).max())})
        pivot = pivot_new
    r['first_step_pivot0'] = dict(errs[0], ok=errs[0]['mean_abs_err'] < 1e-3)        # informational: the variance is NOT expected to hold
    r['second_step_pivot_prev_mean'] = dict(errs[1], ok=errs[1]['mean_abs_err'] < 1e-4 and errs[1]['var_rel_err'] < 2e-3)
    r['ok'] = r['first_step_pivot0']['ok'] and r['second_step_pivot_prev_mean']['ok']
    return r


@case('bn_relu_bits_mask_equals_z_mask')
def _bn_relu_bits():
    """The forward apply's ReLU bit array (1 bit per element) must be exactly (z > 0), and the backward kernels must
    produce the same sums and gradients from it (mask_mode 3) as from z (mask_mode 1), for one and two branches."""
    import ctypes as C
    from pmdist.ops import _lib, nn as nk
    from pmdist.ops._lib import check, ptr, stream_ptr

    class Br(C.Structure):
        _fields_ = [('y', C.c_void_p), ('mean', C.c_void_p), ('invstd', C.c_void_p), ('gamma', C.c_void_p),
                    ('sums', C.c_void_p), ('totals', C.c_void_p), ('dy', C.c_void_p), ('dgamma', C.c_void_p),
                    ('dbeta', C.c_void_p), ('pivot', C.c_void_p), ('pivot_new', C.c_void_p)]
    lib = _lib.lib()
    r = {}
    for Cc, N, H, nb in ((64, 16, 32, 1), (128, 8, 16, 2), (512, 8, 4, 2)):
        rows = N * H * H
        ys = [(_rand_bf16(N, H, H, Cc, seed=50 + i) * 1.3 + 0.2).to(torch.bfloat16) for i in range(2)]
        gam = [torch.rand(Cc, device='cuda') + 0.5 for _ in range(2)]
        bet = [torch.randn(Cc, device='cuda') * 0.1 for _ in range(2)]
        fin = [nk.bn_finalize([nk.bn_stats(y)], float(rows), g, b, None, None, None) for y, g, b in zip(ys, gam, bet)]
        z = torch.empty_like(ys[0])
        bits = torch.zeros(rows * Cc // 8, dtype=torch.uint8, device='cuda')
        if nb == 2:
            check(lib.pmd_bn_apply_bits(ptr(ys[0]), ptr(fin[0][0]), ptr(fin[0][1]), ptr(ys[1]), ptr(fin[1][0]), ptr(fin[1][1]),
                                        C.c_int(1), ptr(z), ptr(bits), C.c_longlong(rows), C.c_int(Cc), stream_ptr()))
        else:
            check(lib.pmd_bn_apply_bits(ptr(ys[0]), ptr(fin[0][0]), ptr(fin[0][1]), ptr(ys[1]), None, None,
                                        C.c_int(1), ptr(z), ptr(bits), C.c_longlong(rows), C.c_int(Cc), stream_ptr()))
        torch.cuda.synchronize()
        want = ((z.view(-1, 8) > 0).to(torch.int32) << torch.arange(8, device='cuda', dtype=torch.int32)).sum(1).to(torch.uint8)
        tag = f'c{Cc}_nb{nb}'
        r[f'bits_exact_{tag}'] = {'ok': bool(torch.equal(bits, want))}
        dz = _rand_bf16(N, H, H, Cc, seed=60)
        outs = {}
        for mode, mask_src in ((1, z), (3, bits)):
            sums = [torch.zeros(2 * Cc, device='cuda') for _ in range(nb)]
            dys = [torch.empty_like(ys[0]) for _ in range(nb)]
            gout = torch.empty_like(ys[0])
            brs = (Br * nb)()
            for i in range(nb):
                brs[i] = Br(ys[i].data_ptr(), fin[i][2].data_ptr(), fin[i][3].data_ptr(), gam[i].data_ptr(), sums[i].data_ptr(), 0,
                            dys[i].data_ptr(), 0, 0, 0, 0)
            check(lib.pmd_bn_bwd_reduce_multi(ptr(dz), ptr(mask_src), None, None, C.c_int(mode), brs, C.c_int(nb),
                                              C.c_longlong(rows), C.c_int(Cc), stream_ptr()))
            check(lib.pmd_bn_bwd_apply_multi(ptr(dz), ptr(mask_src), None, None, C.c_int(mode), brs, C.c_int(nb),
                                             C.c_float(rows), ptr(gout), C.c_longlong(rows), C.c_int(Cc), stream_ptr()))
            torch.cuda.synchronize()
            outs[mode] = (sums, dys, gout)
        r[f'g_out_exact_{tag}'] = {'ok': bool(torch.equal(outs[1][2], outs[3][2]))}
        for i in range(nb):
            r[f'sums{i}_{tag}'] = _cmp(outs[3][0][i], outs[1][0][i], 1e-5)
            r[f'dy{i}_{tag}'] = _cmp(outs[3][1][i], outs[1][1][i], 1e-3)
    r['ok'] = all(v['ok'] for v in r.values())
    return r


@case('bn_two_rank_tables_sum_in_rank_order')
def _bn_two_rank():
    """world=2 pointer tables on one GPU: statistics of two half-batches must combine to the
    statistics of the whole batch (the SyncBN contract, torch/nn/modules/_functions.py:102-115)."""
    from pmdist.ops import nn as nk
    C, N, H = 64, 8, 8
    y = (_rand_bf16(N, H, H, C, seed=13) * 2 + 1).to(torch.bfloat16)
    gamma, beta = torch.rand(C, device='cuda') + 0.5, torch.randn(C, device='cuda')
    halves = [y[:N // 2].contiguous(), y[N // 2:].contiguous()]
    st = [nk.bn_stats(h) for h in halves]
    rows_half = (N // 2) * H * H
    sc2, sh2, m2, i2 = nk.bn_finalize(st, float(rows_half), gamma, beta, None, None, None)
    sc1, sh1, m1, i1 = nk.bn_finalize([nk.bn_stats(y)], float(2 * rows_half), gamma, beta, None, None, None)
    torch.cuda.synchronize()
    r = {'scale': _cmp(sc2, sc1, 1e-5), 'shift': _cmp(sh2, sh1, 1e-4), 'mean': _cmp(m2, m1, 1e-5)}
    r['ok'] = all(v['ok'] for v in r.values())
    return r


@case('bn_eval_affine')
def _bn_eval():
    from pmdist.ops import nn as nk
    C = 128
    gamma, beta = torch.rand(C, device='cuda') + 0.5, torch.randn(C, device='cuda')
    rm, rv = torch.randn(C, device='cuda'), torch.rand(C, device='cuda') + 0.5
    y = _rand_bf16(4, 8, 8, C, seed=14)
    sc, sh = nk.bn_eval_affine(gamma, beta, rm, rv)
    z = nk.bn_apply(y, sc, sh, True)
    torch.cuda.synchronize()
    ref = F.relu(F.batch_norm(_nchw(y), rm, rv, gamma, beta, False, 0.1, 1e-5)).permute(0, 2, 3, 1)
    return _cmp(z, ref, 1e-2)


@case('bn_apply_residual_with_affine')
def _bn_res_affine():
    from pmdist.ops import nn as nk
    C = 128
    y, r_ = _rand_bf16(4, 8, 8, C, seed=15), _rand_bf16(4, 8, 8, C, seed=16)
    a, b, ra, rb = (torch.randn(C, device='cuda') for _ in range(4))
    z = nk.bn_apply(y, a, b, True, r_, ra, rb)
    torch.cuda.synchronize()
    ref = F.relu(y.float() * a + b + r_.float() * ra + rb)
    return _cmp(z, ref, 1e-2)


# ------------------------------------------------------------------ head / loss / optimiser / aug
@case('pool_fc_ce_forward_backward')
def _head_case(N=32, H=4, C=512, classes=10):
    from pmdist.ops import nn as nk
    z = _rand_bf16(N, H, H, C, seed=17).abs().to(torch.bfloat16)
    w = torch.randn(classes, C, device='cuda') * 0.05
    b = torch.randn(classes, device='cuda') * 0.1
    labels = torch.randint(0, classes, (N,), device='cuda')
    feat, logits = nk.pool_fc_fwd(z, w, b)
    dl, metrics = nk.ce_fwd_bwd(logits, labels, None, 1.0 / N)
    dw, db, dz = nk.pool_fc_bwd(dl, feat, w, (H, H))
    torch.cuda.synchronize()
    zr = z.float().requires_grad_(True)
    wr, br = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    lg = F.linear(zr.mean((1, 2)), wr, br)
    loss = F.cross_entropy(lg, labels)
    loss.backward()
    top = lg.topk(5, 1).indices.eq(labels.view(-1, 1))
    r = {'logits': _cmp(logits, lg.detach(), 1e-4), 'dw': _cmp(dw, wr.grad, 1e-3), 'db': _cmp(db, br.grad, 1e-3),
         'dz': _cmp(dz, zr.grad, 1e-2)}
    m = metrics.tolist()
    r['metrics'] = {'got': m, 'want': [loss.item() * N, float(top[:, :1].sum()), float(top.sum()), float(N)]}
    r['metrics']['ok'] = (abs(m[0] - loss.item() * N) < 1e-2 * N and m[1] == float(top[:, :1].sum())
                          and m[2] == float(top.sum()) and m[3] == N)
    r['ok'] = all(v['ok'] for v in r.values())
    return r


CASES['head_pool_fc_ce_large_map_28x28_c2048_split_pool'] = lambda: _head_case(8, 28, 2048, 10)
CASES['head_pool_fc_ce_n512'] = lambda: _head_case(512, 4, 512, 10)


@case('ce_valid_mask_excludes_padding')
def _ce_valid():
    from pmdist.ops import nn as nk
    N, classes = 16, 10
    logits = torch.randn(N, classes, device='cuda')
    labels = torch.randint(0, classes, (N,), device='cuda')
    valid = torch.ones(N, dtype=torch.uint8, device='cuda')
    valid[-3:] = 0
    _, m = nk.ce_fwd_bwd(logits, labels, valid, want_grad=False)
    torch.cuda.synchronize()
    want = F.cross_entropy(logits[:-3], labels[:-3], reduction='sum').item()
    ok = abs(m[0].item() - want) < 1e-3 * N and m[3].item() == N - 3
    return {'ok': ok, 'got': m.tolist(), 'want_loss_sum': want}


@case('sgd_nesterov_matches_torch_three_steps')
def _sgd_case():
    from pmdist.ops import nn as nk
    n = 100003
    p0 = torch.randn(n, device='cuda')
    p, m = p0.clone(), torch.zeros(n, device='cuda')
    pb = torch.empty(n, dtype=torch.bfloat16, device='cuda')
    pt = p0.clone().requires_grad_(True)
    opt = torch.optim.SGD([pt], lr=0.1, momentum=0.9, weight_decay=1e-4, nesterov=True)
    for step in range(3):
        g = torch.randn(n, device='cuda', generator=torch.Generator(device='cuda').manual_seed(step))
        nk.sgd_step(p, g * 4.0, m, pb, 0.1, grad_scale=0.25)      # grad_scale folds the 1/world_size
        pt.grad = g.clone()
        opt.step()
    torch.cuda.synchronize()
    r = {'param': _cmp(p, pt.detach(), 1e-5), 'bf16_copy': _cmp(pb, p, 1e-2)}
    r['ok'] = all(v['ok'] for v in r.values())
    return r


@case('augment_matches_cpu_path_bitwise')
def _aug_case():
    from pmdist.data.loader import augment_cpu, draw_aug_params
    from pmdist.data.sources import synthetic_source
    from pmdist.ops import nn as nk
    src = synthetic_source(64, 32, seed=5)
    aug = draw_aug_params(64, True, torch.Generator().manual_seed(1))
    ref = augment_cpu(src.images, aug).to(torch.bfloat16)            # NCHW
    out = nk.augment(src.images.cuda(), aug.cuda(), 8, 4)
    torch.cuda.synchronize()
    got = out[..., :3].permute(0, 3, 1, 2).cpu()
    ok = torch.equal(got, ref) and bool((out[..., 3] == 0).all())
    return {'ok': ok, 'max_abs_err': (got.float() - ref.float()).abs().max().item()}


# ------------------------------------------------------------------ step engine (single GPU)
def _mk_engine(depth=18, batch=32, hw=32, graph=True, seed=0):
    from pmdist.models.resnet import build_resnet
    from pmdist.parallel.comm import SingleComm
    from pmdist.parallel.engine import NativeEngine
    from pmdist.utils.schedule import SGDConfig
    torch.manual_seed(seed)
    model = build_resnet(depth)
    eng = NativeEngine(model, SingleComm(), SGDConfig(), torch.device('cuda', 0), batch, hw, use_graph=graph)
    return model, eng


def _mk_batch(n, hw=32, train=True, seed=0):
    from pmdist.data.loader import Batch, draw_aug_params
    from pmdist.data.sources import synthetic_source
    src = synthetic_source(n, hw, seed=seed)
    aug = draw_aug_params(n, train, torch.Generator().manual_seed(seed))
    return Batch(src.images, src.labels, torch.ones(n, dtype=torch.bool), aug, train)


def _torch_twin(model):
    """An ordinary fp32 torch copy (contiguous NCHW weights) of an engine-bound model."""
    from pmdist.models.resnet import build_resnet
    twin = build_resnet(model.spec.depth, model.spec.num_classes).cuda()
    twin.load_state_dict({k: v.detach().clone().contiguous() for k, v in model.state_dict().items()})
    return twin


def _engine_forward_case(depth):
    from pmdist.data.loader import augment_cpu
    model, eng = _mk_engine(depth, 16)
    b = _mk_batch(16, train=False)
    twin = _torch_twin(model).eval()
    eng.eval_step(b)
    eng.synchronize()
    with torch.no_grad():
        ref = twin(augment_cpu(b.u8, b.aug).cuda())
    r = _cmp(eng.logits[:16], ref, 5e-2)
    m = eng.read_metrics()
    r['metrics'] = m
    r['ok'] = r['ok'] and m[3] == 16
    eng.close()
    return r


@case('engine_resident_dataset_matches_host_staged_batches')
def _engine_resident_dataset():
    """Device-resident dataset + in-kernel Philox crop/flip + index gather == the same samples staged from the host
    with the host-side Philox parameters.  Forward-only (evaluation of the training-mode augmentation is not possible,
    so the train batches are checked through the first step's loss): evaluation logits must be BIT-identical (same
    pixels, same labels, same validity mask); training losses agree to the run-to-run noise of the fp32 atomics."""
    from pmdist.data.loader import DeviceBatchLoader
    from pmdist.data.sampler import ShardSampler
    from pmdist.data.sources import synthetic_source
    src = synthetic_source(200, 32, seed=9)
    r = {}
    outs = []
    for resident in (True, False):
        model, eng = _mk_engine(18, 32, seed=3)
        if resident:
            eng.attach_dataset(src)
        ev = DeviceBatchLoader(src, ShardSampler(len(src), 3, 2, False, 5), 32, False, seed=5)  # rank 2: wrap-around pad
        evm = []
        for ib in ev:
            eng.eval_step(ib if resident else ib.to_host_batch(src))
            eng.synchronize()
            evm.append((eng.logits[:ib.n].clone(), eng.read_metrics()))
        dl = DeviceBatchLoader(src, ShardSampler(len(src), 3, 1, True, 5), 32, True, seed=5)   # 67 samples on this rank
        dl.set_epoch(2)
        if resident:
            eng.attach_dataset(src)
        losses = []
        for ib in dl:
            eng.train_step(ib if resident else ib.to_host_batch(src), 0.05)
            m = eng.read_metrics()
            losses.append(m[0] / m[3])
        outs.append((losses, evm))
        eng.close()
    (la, ea), (lb, eb) = outs
    r['eval_logits_bit_identical'] = {'ok': all(torch.equal(x[0], y[0]) for x, y in zip(ea, eb))}
    same_counts = all(x[1][1:] == y[1][1:] for x, y in zip(ea, eb))             # top-1, top-5, counted samples
    same_loss = all(abs(x[1][0] - y[1][0]) <= 1e-5 * max(1.0, abs(y[1][0])) for x, y in zip(ea, eb))   # per-block fp32 atomics
    r['eval_metrics_identical'] = {'ok': same_counts and same_loss and sum(x[1][3] for x in ea) == 66,
                                   'counted': [x[1][3] for x in ea]}
    r['train_losses_close'] = {'ok': all(abs(x - y) <= 2e-2 * max(1.0, abs(y)) for x, y in zip(la, lb)) and len(la) == 3,
                               'resident': la, 'host': lb}
    r['ok'] = all(v['ok'] for v in r.values())
    return r


@case('engine_pipelined_steps_match_plain_steps')
def _engine_pipelined():
    """train_step_pipelined (two input slots, copy streams, metric snapshots) trains on the same batches as train_step:
    per-step losses from the cumulative snapshots must match the plain path's to the run-to-run noise of the atomics,
    counts exactly; the tail batch (smaller n) and a second epoch go through both slots."""
    from pmdist.data.loader import DeviceBatchLoader
    from pmdist.data.sampler import ShardSampler
    from pmdist.data.sources import synthetic_source
    src = synthetic_source(32 * 5 + 7, 32, seed=12)          # 6 batches per epoch, the last one with 7 samples
    outs = []
    for piped in (True, False):
        model, eng = _mk_engine(18, 32, seed=4)
        eng.attach_dataset(src)
        dl = DeviceBatchLoader(src, ShardSampler(len(src), 1, 0, True, 3), 32, True, seed=3)
        per_step, prev = [], [0.0, 0.0, 0.0, 0.0]
        fut = None

        def take(m):
            nonlocal prev
            per_step.append([a - b for a, b in zip(m, prev)])
            prev = m
        for ep in range(2):
            dl.set_epoch(ep)
            for ib in dl:
                if piped:
                    f2 = eng.train_step_pipelined(ib, 0.005)
                    if fut is not None:
                        take(fut.result())           # consumed one step late, as the API documents (4 pinned result slots)
                    fut = f2
                else:
                    eng.train_step(ib, 0.005)
                    take(eng.read_metrics(reset=False))
        if piped:
            take(fut.result())
        outs.append((per_step, eng.weights_fingerprint()))
        eng.close()
    (pa, fa), (pb, fb) = outs
    r = {'steps': {'ok': len(pa) == len(pb) == 12},
         'counts_exact': {'ok': all(x[3] == y[3] for x, y in zip(pa, pb)), 'counts': [x[3] for x in pa]},
         # same batches in the same order at a small learning rate (at lr 0.05 two runs of the SAME path drift apart
         # chaotically after three steps): a wrong or stale index vector would show up as a different batch, i.e. a
         # loss difference of 0.1-0.5
         'losses_close': {'ok': all(abs(x[0] / x[3] - y[0] / y[3]) < 5e-2 for x, y in zip(pa, pb)),
                          'pipelined': [x[0] / x[3] for x in pa], 'plain': [y[0] / y[3] for y in pb]}}
    r['ok'] = all(v['ok'] for v in r.values())
    return r


CASES['engine_eval_logits_resnet18'] = lambda: _engine_forward_case(18)
CASES['engine_eval_logits_resnet50'] = lambda: _engine_forward_case(50)
CASES['engine_eval_logits_resnet34'] = lambda: _engine_forward_case(34)


def _engine_train_case(depth, steps=1, graph=True, batch=32, fixed_floor=False):
    from pmdist.data.loader import augment_cpu
    model, eng = _mk_engine(depth, batch, graph=graph)
    twin = _torch_twin(model).train()
    twin16 = _torch_twin(model).train()          # calibration: the same step under torch bf16 autocast
    opt = torch.optim.SGD(twin.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4, nesterov=True)
    opt16 = torch.optim.SGD(twin16.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4, nesterov=True)
    before = {k: v.detach().clone() for k, v in twin.state_dict().items()}
    losses_ref, losses = [], []
    for s in range(steps):
        b = _mk_batch(batch, train=True, seed=s)
        eng.train_step(b, 0.1)
        losses.append(eng.read_metrics()[0] / batch)
        x, y = augment_cpu(b.u8, b.aug).cuda(), b.labels.cuda()
        loss = F.cross_entropy(twin(x), y)
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses_ref.append(loss.item())
        with torch.autocast('cuda', dtype=torch.bfloat16):
            loss16 = F.cross_entropy(twin16(x).float(), y)
        opt16.zero_grad()
        loss16.backward()
        opt16.step()
    eng.synchronize()
    got, want = model.state_dict(), twin.state_dict()
    cal_num = cal_a = cal_b = 0.0
    for k, v16 in twin16.state_dict().items():
        if 'num_batches' in k or 'running' in k:
            continue
        da = (v16.float() - before[k]).flatten().double()
        db = (want[k].float() - before[k]).flatten().double()
        cal_num += float(da @ db); cal_a += float(da @ da); cal_b += float(db @ db)
    cal_cos = cal_num / (math.sqrt(cal_a * cal_b) + 1e-30)
    worst_cos, worst_name = 1.0, ''
    num = den_a = den_b = 0.0
    for k in want:
        if 'num_batches' in k or 'running' in k:
            continue
        da = (got[k].float() - before[k]).flatten().double()
        db = (want[k].float() - before[k]).flatten().double()
        num += float(da @ db)
        den_a += float(da @ da)
        den_b += float(db @ db)
        cos = float(da @ db) / (float(da.norm() * db.norm()) + 1e-30)
        if cos < worst_cos:
            worst_cos, worst_name = cos, k
    total_cos = num / (math.sqrt(den_a * den_b) + 1e-30)
    rm = _cmp(got['bn1.running_mean'], want['bn1.running_mean'], 2e-2)
    rv = _cmp(got['layer4.0.bn2.running_var'], want['layer4.0.bn2.running_var'], 5e-2)
    r = {'loss': losses, 'loss_ref': losses_ref, 'update_cosine_total': total_cos,
         'update_cosine_of_torch_bf16_autocast': cal_cos, 'worst_tensor_cosine': worst_cos,
         'worst_tensor': worst_name, 'running_mean': rm, 'running_var': rv,
         'nbt': int(got['bn1.num_batches_tracked']), 'launches_per_step': eng.launches_per_step(batch),
         'batch': batch, 'steps': steps}
    if steps > 1:
        # Running statistics after several lr = 0.1 steps: the stem's (bn1) depend on the stem weights only and must
        # still agree; the deepest layer's input distribution follows two diverging weight trajectories (bf16 vs fp32
        # through 12 layers at random init), so its running_var is held to a loose 25 % of the largest entry.
        rv = _cmp(got['layer4.0.bn2.running_var'], want['layer4.0.bn2.running_var'], 0.25)
        rm = _cmp(got['bn1.running_mean'], want['bn1.running_mean'], 5e-2)
        r['running_mean'], r['running_var'] = rm, rv
    stats_ok = rm['ok'] and rv['ok']
    if fixed_floor:
        # the bench configuration (512 images): a fixed bar, no calibration
        r['ok'] = (total_cos >= 0.985 and abs(losses[0] - losses_ref[0]) < 0.02 and stats_ok and r['nbt'] == steps)
    else:
        # bf16 activations/gradients through BatchNorm's mean-removing backward are noisy at random init and batch 32;
        # the bar is "at least as faithful to fp32 as torch's own bf16 autocast step, minus a small margin"
        r['ok'] = (total_cos > min(0.98, cal_cos - 0.03) and abs(losses[0] - losses_ref[0]) < 0.05 and stats_ok
                   and r['nbt'] == steps)
    print('engine_train_case', json.dumps({k: r[k] for k in ('batch', 'steps', 'loss', 'loss_ref', 'update_cosine_total',
                                                            'update_cosine_of_torch_bf16_autocast', 'worst_tensor_cosine',
                                                            'worst_tensor')}), flush=True)
    eng.close()
    return r


CASES['engine_train_step_resnet18_matches_torch'] = lambda: _engine_train_case(18, 1)
CASES['engine_train_3steps_resnet18_graph'] = lambda: _engine_train_case(18, 3)
CASES['engine_train_step_resnet50_matches_torch'] = lambda: _engine_train_case(50, 1)
CASES['engine_train_step_resnet34_identity_shortcuts'] = lambda: _engine_train_case(34, 1)
CASES['engine_train_step_resnet18_batch512_fixed_floor'] = lambda: _engine_train_case(18, 1, batch=512, fixed_floor=True)


# every 64-channel dgrad on the weight-resident kernel: fused reductions, ReLU-bit addend (identity shortcuts)
CASES['engine_train_step_resnet18_batch512_resident_dgrads'] = lambda: _with_env(
    lambda: _engine_train_case(18, 1, batch=512, fixed_floor=True), PMD_CONV_RESIDENT=5)
CASES['engine_train_step_resnet34_resident_dgrads'] = lambda: _with_env(lambda: _engine_train_case(34, 1), PMD_CONV_RESIDENT=5)
CASES['engine_train_3steps_resnet18_graph_resident_dgrads'] = lambda: _with_env(lambda: _engine_train_case(18, 3), PMD_CONV_RESIDENT=5)


@case('engine_loss_decreases_and_tail_batch')
def _engine_learns():
    model, eng = _mk_engine(18, 64)
    from pmdist.data.loader import BatchLoader
    from pmdist.data.sampler import ShardSampler
    from pmdist.data.sources import synthetic_source
    src = synthetic_source(64 * 6 + 20, 32, seed=3).pin()       # last batch has 20 samples
    ld = BatchLoader(src, ShardSampler(len(src), 1, 0, True, 0), 64, True, pin=True)
    hist = []
    for ep in range(4):
        ld.set_epoch(ep)
        for b in ld:
            eng.train_step(b, 0.05)
        m = eng.read_metrics()
        hist.append(m[0] / m[3])
    sizes = [b.n for b in ld]
    ok = hist[-1] < hist[0] - 0.1 and all(math.isfinite(h) for h in hist) and sizes[-1] == 20
    eng.close()
    return {'ok': ok, 'loss_per_epoch': hist, 'batch_sizes': sizes}


@case('engine_checkpoint_roundtrip')
def _engine_ckpt():
    model, eng = _mk_engine(18, 16)
    eng.train_step(_mk_batch(16, seed=1), 0.1)
    sd = eng.model_state()
    opt = eng.optimizer_state()
    fp = eng.weights_fingerprint()
    eng.close()
    model2, eng2 = _mk_engine(18, 16, seed=5)
    eng2.load_model_state(sd)
    eng2.load_optimizer_state(opt)
    ok = abs(eng2.weights_fingerprint() - fp) < 1e-6 * max(1.0, abs(fp))
    keys_ok = len(sd) == 74 and all(v.is_contiguous() and v.device.type == 'cpu' for v in sd.values())
    mom_ok = all(torch.equal(a, b) for a, b in zip(opt['momentum'].values(), eng2.optimizer_state()['momentum'].values()))
    eng2.close()
    return {'ok': ok and keys_ok and mom_ok, 'fingerprint': fp, 'keys_ok': keys_ok, 'momentum_ok': mom_ok}


# ------------------------------------------------------------------ v2 kernels: im2col stem, two-branch BN backward
@case('stem_im2col_conv_matches_conv3x3')
def _stem_im2col(N=8, H=32):
    """augment_im2col + 64-wide 1x1 tensor-core conv == augment + 3x3 conv (pad 1) of the RGB image;
    its tensor-core wgrad lands in the [K][27] KRSC stem gradient."""
    import ctypes as C
    from pmdist.data.loader import augment_cpu, draw_aug_params
    from pmdist.data.sources import synthetic_source
    from pmdist.ops import _lib, conv as cv
    from pmdist.ops._lib import check, ptr, stream_ptr, PmdConvShape
    src = synthetic_source(N, H, seed=21)
    aug = draw_aug_params(N, True, torch.Generator().manual_seed(2))
    x = augment_cpu(src.images, aug).cuda().to(torch.bfloat16).float()                # NCHW, bf16-rounded
    w = _rand_bf16(64, 3, 3, 3, scale=0.2, seed=22)
    col = torch.zeros((N, H, H, 64), dtype=torch.bfloat16, device='cuda')     # columns 32..63 stay as allocated
    lib = _lib.lib()
    u8_dev, aug_dev = src.images.cuda(), aug.cuda()      # keep both alive: a freed temporary's block is reused at once
    check(lib.pmd_augment_im2col(ptr(u8_dev), ptr(aug_dev), C.c_int(N), C.c_int(H), C.c_int(H), C.c_int(8),
                                 ptr(col), stream_ptr()))
    wk = cv.krsc_from_torch(w).reshape(64, 27).contiguous()
    wp = torch.empty((64, 64), dtype=torch.bfloat16, device='cuda')
    check(lib.pmd_pad_stem_weight(ptr(wk), ptr(wp), C.c_int(64), C.c_int(27), stream_ptr()))
    y = cv.conv_fprop(col, wp.view(64, 1, 1, 64), 1)
    torch.cuda.synchronize()
    ref = F.conv2d(x, w.float(), padding=1).permute(0, 2, 3, 1)
    r = {'fprop': _cmp(y, ref, 2e-2)}
    # wgrad through the engine-internal row mask is exercised by the engine cases; here: plain 64-col wgrad
    dy = _rand_bf16(N, H, H, 64, seed=23)
    dw = cv.conv_wgrad(col, dy, 1, 1)                                               # [64,1,1,64]
    torch.cuda.synchronize()
    refw = torch.nn.grad.conv2d_weight(x, (64, 3, 3, 3), _nchw(dy), padding=1).permute(0, 2, 3, 1).reshape(64, 27)
    r['wgrad'] = _cmp(dw.view(64, 64)[:, :27], refw, 2e-3)
    r['wgrad_pad_cols_zero'] = {'ok': bool((dw.view(64, 64)[:, 27:] == 0).all())}
    r['ok'] = all(v['ok'] for v in r.values())
    return r


CASES['stem_im2col_conv_matches_conv3x3_odd_size_h20'] = lambda: _stem_im2col(5, 20)   # partial tiles of the 4 x 32 augmentation blocks


def _stem_fused(N=8, H=32, timed=0):
    """stem_fused_kernel (gather/crop/flip/normalise + stem conv + BN statistics in one persistent kernel) against the
    two-kernel form it replaces: y and x0 bit-identical, statistics equal up to summation order, labels/valid equal;
    both batch sources (staged images + parameters, resident dataset + Philox)."""
    import ctypes as C
    from pmdist.data.loader import draw_aug_params
    from pmdist.data.sources import synthetic_source
    from pmdist.ops import _lib, conv as cv
    from pmdist.ops._lib import check, ptr, stream_ptr

    class SF(C.Structure):
        _fields_ = [('ds', C.c_void_p), ('hdr', C.c_void_p), ('idx', C.c_void_p), ('labels', C.c_void_p), ('valid', C.c_void_p),
                    ('u8', C.c_void_p), ('aug', C.c_void_p), ('N', C.c_int), ('H', C.c_int), ('W', C.c_int), ('pad', C.c_int),
                    ('w_pad', C.c_void_p), ('x0', C.c_void_p), ('y', C.c_void_p), ('stats', C.c_void_p), ('pivot', C.c_void_p),
                    ('w_rsc', C.c_int)]
    lib = _lib.lib()
    M = 16384 if timed else max(2 * N, 16)      # timed: a dataset larger than a few MB, so the gather misses L2 like a real one
    src = synthetic_source(M, H, seed=31)
    w = _rand_bf16(64, 3, 3, 3, scale=0.2, seed=32)
    wk = cv.krsc_from_torch(w).reshape(64, 27).contiguous()
    wp = torch.empty((64, 64), dtype=torch.bfloat16, device='cuda')
    check(lib.pmd_pad_stem_weight(ptr(wk), ptr(wp), C.c_int(64), C.c_int(27), stream_ptr()))
    pivot = (torch.randn(64, device='cuda') * 0.1).contiguous()
    images, labels = src.images.cuda(), src.labels.cuda().long()
    r = {}
    # (a) staged images + host-drawn parameters
    aug = draw_aug_params(N, True, torch.Generator().manual_seed(4)).cuda()
    u8 = images[:N].contiguous()
    col_a = torch.zeros((N, H, H, 64), dtype=torch.bfloat16, device='cuda')
    check(lib.pmd_augment_im2col(ptr(u8), ptr(aug), C.c_int(N), C.c_int(H), C.c_int(H), C.c_int(8), ptr(col_a), stream_ptr()))
    st_a = torch.zeros(128, device='cuda')
    y_a = torch.empty((N, H, H, 64), dtype=torch.bfloat16, device='cuda')
    from pmdist.ops._lib import PmdConvShape
    shp = PmdConvShape(N, H, H, 64, 64, 1, 1)
    # statistics of the two-kernel form around the same pivot: computed here from y (fp32)
    check(lib.pmd_conv_fprop(C.byref(shp), ptr(col_a), ptr(wp), ptr(y_a), None, stream_ptr()))
    col_f = torch.zeros_like(col_a)
    y_f = torch.empty_like(y_a)
    st_f = torch.zeros(128, device='cuda')
    a = SF(0, 0, 0, 0, 0, u8.data_ptr(), aug.data_ptr(), N, H, H, 8, wp.data_ptr(), col_f.data_ptr(), y_f.data_ptr(),
           st_f.data_ptr(), pivot.data_ptr())
    check(lib.pmd_stem_fused(C.byref(a), stream_ptr()))
    torch.cuda.synchronize()
    yf = y_a.float().reshape(-1, 64) - pivot
    st_ref = torch.cat([yf.sum(0), (yf * yf).sum(0)])
    r['staged_y_bitwise'] = {'ok': bool(torch.equal(y_f, y_a))}
    r['staged_x0_bitwise'] = {'ok': bool(torch.equal(col_f, col_a))}
    r['staged_stats'] = _cmp(st_f, st_ref, 1e-4)
    # (b) resident dataset + in-kernel Philox; one wrap-around duplicate (~index)
    desc = torch.tensor([images.data_ptr(), labels.data_ptr()], dtype=torch.int64, device='cuda')
    hdr = torch.tensor([1234, 0x5EED0A06, 3, 1], dtype=torch.int64).to(torch.int32).cuda()
    idx = torch.randperm(M, generator=torch.Generator().manual_seed(6))[:N].to(torch.int32)
    idx[N - 1] = ~idx[N - 1]
    idx = idx.cuda()
    col_g = torch.zeros_like(col_a)
    lab_g = torch.zeros(N, dtype=torch.int64, device='cuda')
    val_g = torch.zeros(N, dtype=torch.uint8, device='cuda')
    check(lib.pmd_augment_im2col_gather(ptr(desc), ptr(hdr), ptr(idx), C.c_int(N), C.c_int(H), C.c_int(H), C.c_int(8),
                                        ptr(col_g), ptr(lab_g), ptr(val_g), stream_ptr()))
    y_g = torch.empty_like(y_a)
    check(lib.pmd_conv_fprop(C.byref(shp), ptr(col_g), ptr(wp), ptr(y_g), None, stream_ptr()))
    col_h = torch.zeros_like(col_a)
    y_h = torch.empty_like(y_a)
    lab_h = torch.zeros_like(lab_g)
    val_h = torch.zeros_like(val_g)
    st_h = torch.zeros(128, device='cuda')
    # this form takes the raw [64][27] weight and pads it in shared memory (what the engine does)
    b = SF(desc.data_ptr(), hdr.data_ptr(), idx.data_ptr(), lab_h.data_ptr(), val_h.data_ptr(), 0, 0, N, H, H, 8, wk.data_ptr(),
           col_h.data_ptr(), y_h.data_ptr(), st_h.data_ptr(), 0, 27)
    check(lib.pmd_stem_fused(C.byref(b), stream_ptr()))
    torch.cuda.synchronize()
    yg = y_g.float().reshape(-1, 64)
    r['gather_y_bitwise'] = {'ok': bool(torch.equal(y_h, y_g))}
    r['gather_x0_bitwise'] = {'ok': bool(torch.equal(col_h, col_g))}
    r['gather_stats_no_pivot'] = _cmp(st_h, torch.cat([yg.sum(0), (yg * yg).sum(0)]), 1e-4)
    r['gather_labels_valid'] = {'ok': bool(torch.equal(lab_h, lab_g) and torch.equal(val_h, val_g) and int(val_h[-1]) == 0)}
    # (c) evaluation form: no x0, no statistics
    hdr[3] = 0
    b.x0, b.stats = 0, 0
    y_e = torch.empty_like(y_a)
    b.y = y_e.data_ptr()
    check(lib.pmd_stem_fused(C.byref(b), stream_ptr()))
    check(lib.pmd_augment_im2col_gather(ptr(desc), ptr(hdr), ptr(idx), C.c_int(N), C.c_int(H), C.c_int(H), C.c_int(8),
                                        ptr(col_g), ptr(lab_g), ptr(val_g), stream_ptr()))
    check(lib.pmd_conv_fprop(C.byref(shp), ptr(col_g), ptr(wp), ptr(y_g), None, stream_ptr()))
    torch.cuda.synchronize()
    r['eval_y_bitwise'] = {'ok': bool(torch.equal(y_e, y_g))}
    if timed:
        b.x0, b.stats, b.y = col_h.data_ptr(), st_h.data_ptr(), y_h.data_ptr()
        hdr[3] = 1
        st_x = torch.zeros(128, device='cuda')

        def two():
            check(lib.pmd_augment_im2col_gather(ptr(desc), ptr(hdr), ptr(idx), C.c_int(N), C.c_int(H), C.c_int(H), C.c_int(8),
                                                ptr(col_g), ptr(lab_g), ptr(val_g), stream_ptr()))
            check(lib.pmd_conv_fprop(C.byref(shp), ptr(col_g), ptr(wp), ptr(y_g), ptr(st_x), stream_ptr()))

        def one():
            check(lib.pmd_stem_fused(C.byref(b), stream_ptr()))
        b_eval = SF(desc.data_ptr(), hdr.data_ptr(), idx.data_ptr(), lab_h.data_ptr(), val_h.data_ptr(), 0, 0, N, H, H, 8,
                    wp.data_ptr(), 0, y_h.data_ptr(), 0, 0)

        def one_eval():
            check(lib.pmd_stem_fused(C.byref(b_eval), stream_ptr()))
        b_x0 = SF(desc.data_ptr(), hdr.data_ptr(), idx.data_ptr(), lab_h.data_ptr(), val_h.data_ptr(), 0, 0, N, H, H, 8,
                  wp.data_ptr(), col_h.data_ptr(), y_h.data_ptr(), 0, 0)
        b_st = SF(desc.data_ptr(), hdr.data_ptr(), idx.data_ptr(), lab_h.data_ptr(), val_h.data_ptr(), 0, 0, N, H, H, 8,
                  wp.data_ptr(), 0, y_h.data_ptr(), st_h.data_ptr(), 0)

        def one_x0():
            check(lib.pmd_stem_fused(C.byref(b_x0), stream_ptr()))

        def one_st():
            check(lib.pmd_stem_fused(C.byref(b_st), stream_ptr()))
        for name, fn in (('two_kernels_us', two), ('fused_us', one), ('fused_no_x0_no_stats_us', one_eval),
                         ('fused_x0_only_us', one_x0), ('fused_stats_only_us', one_st)):
            for _ in range(5):
                fn()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(timed):
                fn()
            e1.record()
            torch.cuda.synchronize()
            r[name] = {'ok': True, 'value': e0.elapsed_time(e1) * 1000.0 / timed}
    r['ok'] = all(v['ok'] for v in r.values())
    return r


CASES['stem_fused_matches_two_kernel_form'] = lambda: _stem_fused(8, 32)
CASES['stem_fused_matches_two_kernel_form_h48_n3'] = lambda: _stem_fused(3, 48)
CASES['stem_fused_matches_two_kernel_form_w160_row_tiles'] = lambda: _stem_fused(2, 160)   # tiles inside one image row and straddling two
CASES['stem_fused_matches_two_kernel_form_224'] = lambda: _stem_fused(2, 224)
CASES['bench512_stem_fused'] = lambda: _stem_fused(512, 32, timed=50)


@case('bn_bwd_two_branch_and_mask_from_y')
def _bn_multi():
    import ctypes as C
    from pmdist.ops import _lib, nn as nk
    from pmdist.ops._lib import check, ptr, stream_ptr

    class Br(C.Structure):
        _fields_ = [('y', C.c_void_p), ('mean', C.c_void_p), ('invstd', C.c_void_p), ('gamma', C.c_void_p),
                    ('sums', C.c_void_p), ('totals', C.c_void_p), ('dy', C.c_void_p), ('dgamma', C.c_void_p),
                    ('dbeta', C.c_void_p), ('pivot', C.c_void_p), ('pivot_new', C.c_void_p)]
    lib = _lib.lib()
    Cc, N, H = 128, 8, 8
    rows = N * H * H
    ys = [(_rand_bf16(N, H, H, Cc, seed=30 + i) * 1.3 + 0.2).to(torch.bfloat16) for i in range(2)]
    gam = [torch.rand(Cc, device='cuda') + 0.5 for _ in range(2)]
    bet = [torch.randn(Cc, device='cuda') * 0.1 for _ in range(2)]
    fin = [nk.bn_finalize([nk.bn_stats(y)], float(rows), g, b, None, None, None) for y, g, b in zip(ys, gam, bet)]
    z = nk.bn_apply(ys[0], fin[0][0], fin[0][1], True, ys[1], fin[1][0], fin[1][1])      # relu(bn0(y0) + bn1(y1))
    dz = _rand_bf16(N, H, H, Cc, seed=33)
    # two-branch pass
    sums = [torch.zeros(2 * Cc, device='cuda') for _ in range(2)]
    dys = [torch.empty_like(ys[0]) for _ in range(2)]
    dgs = [torch.zeros(Cc, device='cuda') for _ in range(4)]
    brs = (Br * 2)()
    for i in range(2):
        brs[i] = Br(ys[i].data_ptr(), fin[i][2].data_ptr(), fin[i][3].data_ptr(), gam[i].data_ptr(), sums[i].data_ptr(), 0,
                    dys[i].data_ptr(), dgs[2 * i].data_ptr(), dgs[2 * i + 1].data_ptr(), 0, 0)
    check(lib.pmd_bn_bwd_reduce_multi(ptr(dz), ptr(z), None, None, C.c_int(1), brs, C.c_int(2), C.c_longlong(rows),
                                      C.c_int(Cc), stream_ptr()))
    check(lib.pmd_bn_bwd_apply_multi(ptr(dz), ptr(z), None, None, C.c_int(1), brs, C.c_int(2), C.c_float(rows), None,
                                     C.c_longlong(rows), C.c_int(Cc), stream_ptr()))
    torch.cuda.synchronize()
    r = {}
    for i in range(2):   # oracle: the single-branch kernels
        s1 = nk.bn_bwd_reduce(dz, z, ys[i], fin[i][2], fin[i][3], True)
        dg, db = torch.zeros(Cc, device='cuda'), torch.zeros(Cc, device='cuda')
        d1 = nk.bn_bwd_apply(dz, z, ys[i], fin[i][2], fin[i][3], gam[i], [s1], 0, float(rows), True, dg, db)
        torch.cuda.synchronize()
        r[f'dy{i}'] = _cmp(dys[i], d1, 1e-3)     # one bf16 ulp where the atomically-summed totals differ in the last bit
        r[f'sums{i}'] = _cmp(sums[i], s1, 1e-4)
        r[f'dgamma{i}'] = _cmp(dgs[2 * i], dg, 1e-4)
    # mask recomputed from y (no residual): must equal the z-based mask
    z1 = nk.bn_apply(ys[0], fin[0][0], fin[0][1], True)
    sA, sB = torch.zeros(2 * Cc, device='cuda'), torch.zeros(2 * Cc, device='cuda')
    dyA, dyB = torch.empty_like(z1), torch.empty_like(z1)
    for mode, zz, ss, dd in ((1, z1, sA, dyA), (2, None, sB, dyB)):
        b1 = (Br * 1)()
        b1[0] = Br(ys[0].data_ptr(), fin[0][2].data_ptr(), fin[0][3].data_ptr(), gam[0].data_ptr(), ss.data_ptr(), 0,
                   dd.data_ptr(), 0, 0, 0, 0)
        check(lib.pmd_bn_bwd_reduce_multi(ptr(dz), ptr(zz), ptr(fin[0][0]), ptr(fin[0][1]), C.c_int(mode), b1, C.c_int(1),
                                          C.c_longlong(rows), C.c_int(Cc), stream_ptr()))
        check(lib.pmd_bn_bwd_apply_multi(ptr(dz), ptr(zz), ptr(fin[0][0]), ptr(fin[0][1]), C.c_int(mode), b1, C.c_int(1),
                                         C.c_float(rows), None, C.c_longlong(rows), C.c_int(Cc), stream_ptr()))
    torch.cuda.synchronize()
    r['mask_from_y_dy'] = _cmp(dyB, dyA, 1e-3)
    r['mask_from_y_sums'] = _cmp(sB, sA, 1e-5)
    r['ok'] = all(v['ok'] for v in r.values())
    return r


@case('bn_finalize_apply_fused_matches_separate_kernels')
def _bn_fused_fwd():
    """One launch (finalize + apply, incl. the projection shortcut's BatchNorm) == stats -> finalize -> apply."""
    import ctypes as C
    from pmdist.ops import _lib, nn as nk
    from pmdist.ops._lib import check, ptr, stream_ptr

    class Fin(C.Structure):
        _fields_ = [('stats', C.c_void_p), ('gamma', C.c_void_p), ('beta', C.c_void_p), ('running_mean', C.c_void_p),
                    ('running_var', C.c_void_p), ('num_batches', C.c_void_p), ('scale', C.c_void_p), ('shift', C.c_void_p),
                    ('mean', C.c_void_p), ('invstd', C.c_void_p), ('pivot', C.c_void_p), ('pivot_new', C.c_void_p)]
    lib = _lib.lib()
    Cc, N, H = 128, 8, 8
    rows = N * H * H
    ys = [(_rand_bf16(N, H, H, Cc, seed=40 + i) * 1.3 + 0.2).to(torch.bfloat16) for i in range(2)]
    gam = [torch.rand(Cc, device='cuda') + 0.5 for _ in range(2)]
    bet = [torch.randn(Cc, device='cuda') * 0.1 for _ in range(2)]
    stats = [nk.bn_stats(y) for y in ys]
    # separate kernels (oracle), with their own running-stat buffers
    rmA = [torch.zeros(Cc, device='cuda') for _ in range(2)]
    rvA = [torch.ones(Cc, device='cuda') for _ in range(2)]
    nbA = [torch.zeros(1, dtype=torch.int64, device='cuda') for _ in range(2)]
    fin = [nk.bn_finalize([st], float(rows), g, b, rm, rv, nb) for st, g, b, rm, rv, nb in zip(stats, gam, bet, rmA, rvA, nbA)]
    z_ref = nk.bn_apply(ys[0], fin[0][0], fin[0][1], True, ys[1], fin[1][0], fin[1][1])
    # fused kernel
    rmB = [torch.zeros(Cc, device='cuda') for _ in range(2)]
    rvB = [torch.ones(Cc, device='cuda') for _ in range(2)]
    nbB = [torch.zeros(1, dtype=torch.int64, device='cuda') for _ in range(2)]
    outs = [[torch.empty(Cc, device='cuda') for _ in range(4)] for _ in range(2)]
    fs = []
    for i in range(2):
        fs.append(Fin(stats[i].data_ptr(), gam[i].data_ptr(), bet[i].data_ptr(), rmB[i].data_ptr(), rvB[i].data_ptr(),
                      nbB[i].data_ptr(), outs[i][0].data_ptr(), outs[i][1].data_ptr(), outs[i][2].data_ptr(),
                      outs[i][3].data_ptr(), 0, 0))
    z = torch.empty_like(ys[0])
    check(lib.pmd_bn_finalize_apply(C.byref(fs[0]), C.byref(fs[1]), C.c_float(rows), C.c_float(0.1), C.c_float(1e-5),
                                    ptr(ys[0]), ptr(ys[1]), C.c_int(1), ptr(z), C.c_longlong(rows), C.c_int(Cc), stream_ptr()))
    z1 = torch.empty_like(ys[0])          # single BatchNorm, plain residual
    f_single = Fin(stats[0].data_ptr(), gam[0].data_ptr(), bet[0].data_ptr(), 0, 0, 0, outs[0][0].data_ptr(),
                   outs[0][1].data_ptr(), outs[0][2].data_ptr(), outs[0][3].data_ptr(), 0, 0)
    check(lib.pmd_bn_finalize_apply(C.byref(f_single), None, C.c_float(rows), C.c_float(0.1), C.c_float(1e-5),
                                    ptr(ys[0]), ptr(ys[1]), C.c_int(1), ptr(z1), C.c_longlong(rows), C.c_int(Cc), stream_ptr()))
    torch.cuda.synchronize()
    z1_ref = nk.bn_apply(ys[0], fin[0][0], fin[0][1], True, ys[1])
    torch.cuda.synchronize()
    r = {'z': _cmp(z, z_ref, 1e-3), 'z_plain_residual': _cmp(z1, z1_ref, 1e-3)}
    for i in range(2):
        r[f'scale{i}'] = _cmp(outs[i][0], fin[i][0], 1e-6)
        r[f'mean{i}'] = _cmp(outs[i][2], fin[i][2], 1e-6)
        r[f'invstd{i}'] = _cmp(outs[i][3], fin[i][3], 1e-6)
        r[f'running_var{i}'] = _cmp(rvB[i], rvA[i], 1e-6)
        r[f'nbt{i}'] = {'ok': int(nbB[i]) == 1}
    r['ok'] = all(v['ok'] for v in r.values())
    return r
