"""Times the fprop / dgrad kernels of the ResNet18-ref layers at the bench batch (512) under different kernel
switches (CTA pairs on/off, pipeline depth), cold L2 (a 512 MB write between launches), CUDA events per launch.

    python tools/conv_sweep.py            # prints one JSON line per (layer, pass, setting)
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pmdist.ops import conv as cv  # noqa: E402

LAYERS = {'layer2_conv2': (16, 128, 128, 3, 1), 'layer3_conv2': (8, 256, 256, 3, 1), 'layer4_conv2': (4, 512, 512, 3, 1),
          'layer2_conv1': (32, 64, 128, 3, 2), 'layer3_conv1': (16, 128, 256, 3, 2), 'layer4_conv1': (8, 256, 512, 3, 2)}
SETTINGS = [{'PMD_GEMM_2CTA': '0'}, {'PMD_GEMM_2CTA': '1'}, {'PMD_GEMM_2CTA': '1', 'PMD_GEMM2_STAGES128': '4', 'PMD_GEMM2_STAGES256': '3'}]
if os.environ.get('PMD_SWEEP') == 'bn':      # tile width instead of CTA pairs
    SETTINGS = [{}, {'PMD_GEMM_BN': '64'}, {'PMD_GEMM_BN': '128'}, {'PMD_GEMM_BN': '256'}]
N = 512
flush = torch.empty(512 << 20, dtype=torch.uint8, device='cuda')


def timed(fn, reps=6):
    out = []
    for _ in range(reps):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        out.append(a.elapsed_time(b) * 1e3)
    out.sort()
    return out[len(out) // 2]


for name, (h, c, k, r, st) in LAYERS.items():
    ho = cv.out_dim(h, r, st)
    x = torch.randn(N, h, h, c, device='cuda').to(torch.bfloat16)
    w = (torch.randn(k, r, r, c, device='cuda') * 0.05).to(torch.bfloat16)
    dy = torch.randn(N, ho, ho, k, device='cuda').to(torch.bfloat16)
    stats = torch.zeros(2 * k, device='cuda')
    for setting in SETTINGS:
        for kk in ('PMD_GEMM_2CTA', 'PMD_GEMM2_STAGES128', 'PMD_GEMM2_STAGES256', 'PMD_GEMM_BN'):
            os.environ.pop(kk, None)
        os.environ.update(setting)
        f = timed(lambda: cv.conv_fprop(x, w, st, stats))
        d = timed(lambda: cv.conv_dgrad(dy, w, tuple(x.shape), r, st, krsc=True))
        print(json.dumps({'layer': name, 'setting': setting, 'fprop_us': round(f, 1), 'dgrad_us': round(d, 1)}), flush=True)
