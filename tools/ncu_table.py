"""Turns `ncu -i step.ncu-rep --page raw --csv` of one training step into the per-kernel table kept in profiles/.

    ncu --set full --clock-control none --import-source on -s <skip> -c <launches of one step> -o step \
        python bench.py --no-graph --steps 1 --warmup 3 --no-exposed-comm        # right after the same command ran plain
    ncu -i step.ncu-rep --page raw --csv > step_raw.csv
    python tools/ncu_table.py step_raw.csv "title" > profiles/<name>.md

Columns: duration, registers, warps active, tensor-pipe activity (% of active cycles, relative to the NOMINAL rate),
DRAM throughput (%), DRAM bytes read / written, L2 throughput (%); then the same aggregated per kernel name.
"""
from __future__ import annotations

import csv
import re
import sys
from collections import OrderedDict

COLS = OrderedDict([
    ('us', 'gpu__time_duration.sum'),
    ('regs', 'launch__registers_per_thread'),
    ('warps active %', 'sm__warps_active.avg.pct_of_peak_sustained_active'),
    ('tensor pipe %', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active'),
    ('DRAM %', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'),
    ('DRAM rd MB', 'dram__bytes_read.sum'),
    ('DRAM wr MB', 'dram__bytes_write.sum'),
    ('L2 (lts) %', 'lts__throughput.avg.pct_of_peak_sustained_elapsed'),
])
UNIT_SCALE = {'Kbyte': 1e-3, 'Mbyte': 1.0, 'Gbyte': 1e3, 'byte': 1e-6, 'ns': 1e-3, 'us': 1.0, 'ms': 1e3}


def short(name: str) -> str:
    name = re.sub(r'^void\s+', '', name)
    name = re.sub(r'\(.*$', '', name)
    name = re.sub(r'^.*?unnamed>::', '', name)
    return name.replace('pmd::', '').replace('(anonymous namespace)::', '')


def num(x: str) -> float:
    try:
        return float(x.replace(',', ''))
    except ValueError:
        return float('nan')


def main() -> int:
    if len(sys.argv) < 2:
        print(__doc__)
        return 2
    rows = list(csv.reader(open(sys.argv[1])))
    title = sys.argv[2] if len(sys.argv) > 2 else 'ncu --set full, one training step'
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    missing = [c for c in COLS.values() if c not in ix]
    if missing:
        print('missing columns: ' + ', '.join(missing), file=sys.stderr)
        return 1
    kn, gs = ix['Kernel Name'], ix['Grid Size']
    print(f'# {title}\n')
    print('| # | kernel | grid | ' + ' | '.join(COLS) + ' |')
    print('|---|---|---|' + '---|' * len(COLS))
    agg = OrderedDict()
    for n, r in enumerate(data):
        vals = []
        for label, col in COLS.items():
            v = num(r[ix[col]]) * UNIT_SCALE.get(units[ix[col]], 1.0)
            vals.append(v)
        name = short(r[kn])
        fmt = [f'{vals[0]:.1f}', f'{vals[1]:.0f}'] + [f'{v:.1f}' for v in vals[2:]]
        print(f'| {n} | {name} | {r[gs]} | ' + ' | '.join(fmt) + ' |')
        a = agg.setdefault(name, [0, 0.0, 0.0, 0.0])
        a[0] += 1
        a[1] += vals[0]
        a[2] += vals[0] * vals[3]
        a[3] += vals[0] * vals[4]
    total = sum(a[1] for a in agg.values())
    tens = sum(a[2] for a in agg.values())
    conv_t = sum(a[1] for k, a in agg.items() if k.startswith('conv') or k.startswith('stem'))
    conv_tens = sum(a[2] for k, a in agg.items() if k.startswith('conv') or k.startswith('stem'))
    print('\n## Per kernel (time-weighted)\n')
    print('| kernel | launches | us | tensor pipe % | DRAM % |')
    print('|---|---|---|---|---|')
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f'| {name} | {a[0]} | {a[1]:.1f} | {a[2] / a[1]:.1f} | {a[3] / a[1]:.1f} |')
    print(f'| **all {len(data)} launches** | | {total:.1f} | {tens / total:.1f} | |')
    if conv_t > 0:
        print(f'| **convolution kernels only** | | {conv_t:.1f} | {conv_tens / conv_t:.1f} | |')
    return 0


if __name__ == '__main__':
    sys.exit(main())
