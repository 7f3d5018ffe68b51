"""Launches one convolution shape a few times (no validation) so that ncu can capture a steady-state
instance:  ncu --set full --import-source on -k regex:conv_gemm -s 2 -c 1 python tools/prof_conv.py"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pmdist.ops import conv as cv  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--n', type=int, default=256)
ap.add_argument('--h', type=int, default=32)
ap.add_argument('--c', type=int, default=64)
ap.add_argument('--k', type=int, default=64)
ap.add_argument('--r', type=int, default=3)
ap.add_argument('--iters', type=int, default=4)
ap.add_argument('--timed', type=int, default=1, help='launches inside the timed region (mean is printed)')
ap.add_argument('--what', default='fprop', choices=['fprop', 'wgrad', 'dgrad'])
a = ap.parse_args()
x = torch.randn(a.n, a.h, a.h, a.c, device='cuda').to(torch.bfloat16)
w = (torch.randn(a.k, a.r, a.r, a.c, device='cuda') * 0.05).to(torch.bfloat16)
stats = torch.zeros(2 * a.k, device='cuda')
dy = torch.randn(a.n, a.h, a.h, a.k, device='cuda').to(torch.bfloat16)
wt = w.permute(1, 2, 3, 0).contiguous()
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(a.iters - 1 + a.timed):
    if i == a.iters - 1:
        ev0.record()
    if a.what == 'fprop':
        cv.conv_fprop(x, w, 1, stats)
    elif a.what == 'dgrad':
        cv.conv_dgrad(dy, w, tuple(x.shape), a.r, 1, krsc=True)   # the engine's path: KRSC weights, MN-major B
    else:
        cv.conv_wgrad(x, dy, a.r, 1)
ev1.record()
torch.cuda.synchronize()
print(f'{a.what} n={a.n} h={a.h} c={a.c} k={a.k}: {ev0.elapsed_time(ev1) * 1e3 / a.timed:.1f} us')
