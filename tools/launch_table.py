"""Turns an ncu per-launch CSV (--metrics gpu__time_duration.sum --csv) of `bench.py --no-graph` into the per-kernel
table of ONE training step kept under profiles/.  A step is the launch range from the engine's first kernel (the fused
stem kernel, or the augmentation kernel of the two-kernel form) to just before the next step's first kernel.

    ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/l.csv \
        python bench.py --no-graph --steps 1 --warmup 3
    python tools/launch_table.py gpurun_out/l.csv > profiles/step_launches.md
"""
import csv
import re
import sys
from collections import OrderedDict


def short(name: str) -> str:
    name = re.sub(r'^void\s+', '', name)
    name = re.sub(r'\(.*$', '', name)
    name = name.replace('pmd::', '').replace('<unnamed>::', '').replace('unnamed>::', '')
    return name.strip()


def main():
    rows = []
    with open(sys.argv[1], newline='') as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get('Metric Name') != 'gpu__time_duration.sum':
            continue
        rows.append((short(r['Kernel Name']), float(r['Metric Value'].replace(',', '')) / 1e3, r['Grid Size']))
    starts = [i for i, r in enumerate(rows) if r[0].startswith('augment') or r[0].startswith('stem_fused')]
    step = None
    if len(starts) >= 2:            # the last complete step: from one first kernel to just before the next
        step = rows[starts[-2]:starts[-1]]
        while step and not (step[-1][0].startswith('allreduce') or step[-1][0].startswith('pad_stem_weight') or
                            step[-1][0].startswith('conv_wgrad')):
            step.pop()              # host-side torch kernels between steps (metric reads, copies) are not part of it
    if not step:
        step = rows
    agg = OrderedDict()
    for name, us, _ in step:
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += us
    total = sum(v[1] for v in agg.values())
    title = sys.argv[2] if len(sys.argv) > 2 else 'One training step'
    print(f'# {title}\n')
    print('ncu gpu__time_duration.sum per kernel (serialised, one kernel at a time, cold caches)\n')
    print('| kernel | launches | us | share |\n|---|---|---|---|')
    for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f'| {name} | {n} | {us:.1f} | {100 * us / total:.1f}% |')
    print(f'| **total (serialised)** | {len(step)} | {total:.1f} | |')
    print('\n## Launch order\n')
    print('| # | kernel | grid | us |\n|---|---|---|---|')
    for i, (name, us, grid) in enumerate(step):
        print(f'| {i} | {name} | {grid} | {us:.1f} |')


if __name__ == '__main__':
    main()
