"""Per-CTA pipeline timeline of the tensor-core convolution kernels (csrc/conv_tc.cu, pmd_conv_set_trace).

The kernels stamp clock64() at fixed milestones (setup done, first TMA issued, first stage landed, last MMA
issued, accumulator ready, tile staged, epilogue done, exit).  This tool runs one shape, reads the records back
and prints the median / p90 duration of every phase in SM cycles, the CTA lifetime, how many CTAs shared an SM
at a time and the kernel's span measured with %globaltimer.  That is the profile ncu cannot give: where inside
a CTA the time goes.

    python tools/conv_trace.py --n 512 --h 32 --c 64 --k 64 [--what dgrad] [--json out.json]
"""
import argparse
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pmdist.ops import _lib  # noqa: E402
from pmdist.ops import conv as cv  # noqa: E402

SLOTS = 12
NAMES = ['entry', 'setup', 'first_load', 'first_full', 'mma_issued', 'acc_ready', 'staged', 'epi_done', 'exit']
PHASES = [('setup (bar init, TMEM alloc, sync)', 0, 1), ('first TMA issue', 1, 2), ('first stage latency', 2, 3),
          ('mainloop (first full -> last MMA issued)', 3, 4), ('MMA drain (issued -> acc ready)', 4, 5),
          ('epilogue convert+stage', 5, 6), ('stats + store read', 6, 7), ('teardown sync', 7, 8),
          ('CTA lifetime', 0, 8)]


def pct(v, q):
    v = sorted(v)
    return v[min(len(v) - 1, int(q * len(v)))]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--n', type=int, default=512)
    ap.add_argument('--h', type=int, default=32)
    ap.add_argument('--c', type=int, default=64)
    ap.add_argument('--k', type=int, default=64)
    ap.add_argument('--r', type=int, default=3)
    ap.add_argument('--what', default='fprop', choices=['fprop', 'dgrad'])
    ap.add_argument('--warm', type=int, default=3)
    ap.add_argument('--json', default=None)
    a = ap.parse_args()

    x = torch.randn(a.n, a.h, a.h, a.c, device='cuda').to(torch.bfloat16)
    w = (torch.randn(a.k, a.r, a.r, a.c, device='cuda') * 0.05).to(torch.bfloat16)
    dy = torch.randn(a.n, a.h, a.h, a.k, device='cuda').to(torch.bfloat16)
    stats = torch.zeros(2 * a.k, device='cuda')

    def run():
        if a.what == 'fprop':
            cv.conv_fprop(x, w, 1, stats)
        else:
            cv.conv_dgrad(dy, w, tuple(x.shape), a.r, 1, krsc=True)

    for _ in range(a.warm):
        run()
    torch.cuda.synchronize()
    cap = 1 << 16
    buf = torch.zeros(cap * SLOTS, dtype=torch.int64, device='cuda')
    lib = _lib.lib()
    lib.pmd_conv_set_trace.argtypes = [C.c_void_p, C.c_int]
    lib.pmd_conv_set_trace(C.c_void_p(buf.data_ptr()), C.c_int(cap))
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    run()
    ev1.record()
    torch.cuda.synchronize()
    lib.pmd_conv_set_trace(C.c_void_p(0), C.c_int(0))
    rec = buf.view(cap, SLOTS).cpu()
    used = rec[:, 9] != 0
    rec = rec[used].tolist()
    out = {'shape': vars(a), 'ctas': len(rec), 'event_us': ev0.elapsed_time(ev1) * 1e3, 'phases': {}}
    print(f'{a.what} n={a.n} h={a.h} c={a.c} k={a.k} r={a.r}: {len(rec)} CTAs traced, '
          f'launch {out["event_us"]:.1f} us (traced run)')
    resident = (a.r == 3 and a.c == 64 and a.k == 64 and a.h % 16 == 0 and os.environ.get('PMD_CONV_RESIDENT', '1') != '0'
                and os.environ.get('PMD_GEMM_PERSISTENT', '0') == '0')
    if resident:
        # weight-resident persistent kernel: slots 2..7 are per-role totals of WAIT cycles over the CTA's tiles
        names = [('setup', None), ('producer: wait free slot', 2), ('MMA: wait patch landed', 3),
                 ('MMA: wait accumulator free', 4), ('epilogue g0: wait accumulator ready', 5),
                 ('epilogue g0: busy (ld, convert, stage, stats)', 6), ('epilogue g0: wait store read', 7)]
        life = [r[8] - r[0] for r in rec]
        print(f'  CTA lifetime p50 {pct(life, .5)} cycles, tiles per CTA p50 {pct([r[11] for r in rec], .5)}')
        out['phases']['lifetime'] = {'p50': pct(life, .5), 'max': max(life)}
        for name, slot in names:
            d = [(r[1] - r[0]) if slot is None else r[slot] for r in rec]
            out['phases'][name] = {'p50': pct(d, .5), 'p90': pct(d, .9)}
            per_tile = pct([x / max(1, r[11]) for x, r in zip(d, rec)], .5)
            print(f'  {name:48s} p50 {pct(d, .5):8d}  p90 {pct(d, .9):8d} cycles   ({per_tile:7.0f} per tile)')
    for name, i0, i1 in ([] if resident else PHASES):
        d = [r[i1] - r[i0] for r in rec if r[i1] and r[i0]]
        if not d:
            continue
        out['phases'][name] = {'p50': pct(d, .5), 'p90': pct(d, .9), 'min': min(d), 'max': max(d)}
        print(f'  {name:44s} p50 {pct(d, .5):7d}  p90 {pct(d, .9):7d}  min {min(d):7d}  max {max(d):7d}  cycles')
    tiles = [r[11] for r in rec]
    g0 = min(r[9] for r in rec)
    g1 = max(r[9] for r in rec)
    sms = {}
    for r in rec:
        sms.setdefault(r[10], []).append(r)
    # residency: for each SM, the peak number of CTAs whose [entry, exit] clock intervals overlap
    peak = []
    busy = []
    for sm, rs in sms.items():
        evs = []
        for r in rs:
            evs.append((r[0], 1))
            evs.append((r[8], -1))
        evs.sort()
        cur = mx = 0
        for _, dlt in evs:
            cur += dlt
            mx = max(mx, cur)
        peak.append(mx)
        busy.append(max(r[8] for r in rs) - min(r[0] for r in rs))
    out.update({'sms_used': len(sms), 'ctas_per_sm_peak_p50': pct(peak, .5), 'tiles_per_cta_p50': pct(tiles, .5),
                'sm_busy_cycles_p50': pct(busy, .5), 'sm_busy_cycles_max': max(busy),
                'last_cta_start_minus_first_ns': g1 - g0})
    print(f'  SMs used {len(sms)}, CTAs per SM {len(rec) / len(sms):.1f}, resident at once (p50 of per-SM peak) '
          f'{pct(peak, .5)}, tiles per CTA p50 {pct(tiles, .5)}')
    print(f'  SM busy span p50 {pct(busy, .5)} max {max(busy)} cycles; last CTA started {g1 - g0} ns after the first')
    if a.json:
        with open(a.json, 'w') as f:
            json.dump(out, f, indent=1)


if __name__ == '__main__':
    main()
