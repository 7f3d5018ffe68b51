"""pmd_bn_bwd_apply_static through the C ABI.

Some hosts of this library may not allocate device memory at all -- they have given the GPU's memory to
their own pool, or they cannot tolerate the device-wide synchronisation a cudaMalloc causes in the middle of
a step.  pmd_bn_bwd_apply needs a small scratch vector and gets it from cudaMalloc; the _static variant
keeps that scratch in a __device__ array inside the library.  This test drives both entry points with ctypes
on the same inputs, twice in a row with different shapes (the built-in scratch is shared between calls and
still holds the previous call's values), and checks that an oversized request is refused cleanly.

It lives in a file that sorts last so that an `-x` run has already executed every other test if this fails,
and it makes no assumption about what ran before it in the process.
"""
import ctypes as C

import pytest
import torch

F32 = C.c_float


def _inputs(n, h, c, seed):
    g = torch.Generator(device='cuda').manual_seed(seed)
    bf = lambda *s: torch.randn(*s, device='cuda', generator=g).to(torch.bfloat16)
    y = (bf(n, h, h, c) * 1.5 + 0.25).to(torch.bfloat16)
    z = torch.relu(bf(n, h, h, c))
    dz = bf(n, h, h, c)
    mean = torch.randn(c, device='cuda', generator=g) * 0.1
    invstd = torch.rand(c, device='cuda', generator=g) + 0.5
    gamma = torch.rand(c, device='cuda', generator=g) + 0.5
    # two "ranks" worth of partial sums, combined in rank order by the entry point
    sums = [torch.randn(2 * c, device='cuda', generator=g) for _ in range(2)]
    return y, z, dz, mean, invstd, gamma, sums


def _call(fn, y, z, dz, mean, invstd, gamma, sums, rows_total):
    c = y.shape[-1]
    rows = y.numel() // c
    table = torch.tensor([s.data_ptr() for s in sums], dtype=torch.int64, device='cuda')
    dy = torch.empty_like(y)
    g_out = torch.empty_like(y)
    dgamma = torch.zeros(c, device='cuda')
    dbeta = torch.zeros(c, device='cuda')
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr())
    rc = fn(p(dz), p(z), p(y), p(mean), p(invstd), p(gamma), p(table), C.c_int(len(sums)), p(sums[0]),
            F32(rows_total), C.c_int(1), p(dy), p(dgamma), p(dbeta), p(g_out), C.c_longlong(rows), C.c_int(c), stream)
    torch.cuda.synchronize()
    return rc, dy, dgamma, dbeta, g_out


def _close(a, b, rel=1e-3):
    a, b = a.float(), b.float()
    scale = b.abs().max().item() + 1e-12
    return (a - b).abs().max().item() <= rel * scale and not torch.isnan(a).any()


@pytest.mark.gpu
def test_static_scratch_entry_point_matches_allocating_one_twice_in_a_row():
    from pmdist.ops import _lib
    lib = _lib.lib()
    lib.pmd_last_error.restype = C.c_char_p
    # call 1: wide layer; call 2: a narrower one, so stale entries of call 1 remain in the shared scratch
    for seed, (n, h, c) in enumerate([(8, 8, 256), (4, 16, 64)]):
        args = _inputs(n, h, c, seed)
        rows_total = float(2 * n * h * h)
        rc_ref, dy_ref, dg_ref, db_ref, g_ref = _call(lib.pmd_bn_bwd_apply, *args, rows_total)
        rc_new, dy_new, dg_new, db_new, g_new = _call(lib.pmd_bn_bwd_apply_static, *args, rows_total)
        assert rc_ref == 0, lib.pmd_last_error()
        assert rc_new == 0, lib.pmd_last_error()
        assert _close(dy_new, dy_ref) and _close(g_new, g_ref), (seed, c)
        assert _close(dg_new, dg_ref) and _close(db_new, db_ref), (seed, c)
        assert dy_ref.float().abs().max() > 0


@pytest.mark.gpu
def test_static_scratch_entry_point_refuses_more_channels_than_it_holds():
    from pmdist.ops import _lib
    lib = _lib.lib()
    lib.pmd_last_error.restype = C.c_char_p
    args = _inputs(2, 2, 2048, 7)
    before = _lib.launch_count()
    rc, *_ = _call(lib.pmd_bn_bwd_apply_static, *args, 16.0)
    assert rc == -1
    msg = lib.pmd_last_error().decode()
    assert '1024' in msg and '2048' in msg, msg
    assert _lib.launch_count() == before            # nothing was launched
    rc_ok, *_ = _call(lib.pmd_bn_bwd_apply, *args, 16.0)       # the allocating entry point still takes it
    assert rc_ok == 0
