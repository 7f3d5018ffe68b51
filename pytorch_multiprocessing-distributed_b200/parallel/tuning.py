"""Per-bucket choice of the gradient all-reduce algorithm from MEASURED timings.

BASELINE.json north star (1)/(2): "one-shot or two-shot P2P reduction, plus an NVLS-multicast variant
where the host exposes it, chosen per bucket size from measured bus bandwidth".  `tools/allreduce_sweep.py`
times every algorithm x CTA count over 1 KB .. 256 MB on the box at hand and writes a JSON table; tables
measured on the build's 8xB200 NVSwitch box are committed under `pmdist/parallel/tables/` and used when no
fresh one is supplied (env PMDIST_ALLREDUCE_TABLE).  NCCL picks ring/tree/NVLS per size with its internal
tuner; here the table is explicit, inspectable and re-measurable.
"""
from __future__ import annotations

import json
import math
import os
from typing import Dict, List, Optional, Tuple

_HERE = os.path.dirname(os.path.abspath(__file__))
_ALGOS = ('oneshot', 'twoshot', 'nvls')


class AllreduceTable:
    def __init__(self, world: int, rows: List[dict], source: str = '') -> None:
        self.world, self.rows, self.source = world, sorted(rows, key=lambda r: r['bytes']), source

    @staticmethod
    def load(world: int, path: Optional[str] = None) -> Optional['AllreduceTable']:
        cands = [path, os.environ.get('PMDIST_ALLREDUCE_TABLE'),
                 os.path.join(_HERE, 'tables', f'allreduce_w{world}.json')]
        for c in cands:
            if c and os.path.exists(c):
                with open(c) as f:
                    d = json.load(f)
                if int(d.get('world', world)) == world:
                    return AllreduceTable(world, d['rows'], c)
        return None

    def choose(self, nbytes: int, multicast: bool) -> Tuple[str, int, float]:
        """(algo, CTAs, expected microseconds) of the fastest measured variant at the nearest size."""
        row = min(self.rows, key=lambda r: abs(math.log(max(r['bytes'], 1)) - math.log(max(nbytes, 1))))
        best = None
        for k, v in row.items():
            if not k.endswith('_us') or k.startswith('nccl'):
                continue
            algo, blocks = k[:-3].rsplit('_b', 1)
            if algo == 'nvls' and not multicast:
                continue
            if best is None or v < best[2]:
                best = (algo, int(blocks), float(v))
        # scale the expectation linearly with size when the nearest row is not exact (bandwidth regime)
        return best


def choose_bucket_algos(world: int, bucket_bytes: List[int], multicast: bool, forced: str = 'auto',
                        default_blocks: int = 32) -> Tuple[List[str], List[int], str]:
    """Per-bucket (algo, CTAs).  `forced` in {'oneshot','twoshot','nvls'} overrides the table."""
    if world == 1:
        return ['oneshot'] * len(bucket_bytes), [default_blocks] * len(bucket_bytes), 'single rank'
    if forced != 'auto':
        return [forced] * len(bucket_bytes), [default_blocks] * len(bucket_bytes), f'forced {forced}'
    tab = AllreduceTable.load(world)
    if tab is None:
        # no measurement for this world size: latency-optimal one-shot below 256 KB, owner-computes above
        algos = ['oneshot' if b < (256 << 10) else ('nvls' if multicast else 'twoshot') for b in bucket_bytes]
        return algos, [default_blocks if b < (4 << 20) else 96 for b in bucket_bytes], 'heuristic (no table)'
    algos, blocks = [], []
    for b in bucket_bytes:
        a, k, _ = tab.choose(b, multicast)
        algos.append(a)
        blocks.append(k)
    return algos, blocks, f'measured table {os.path.basename(tab.source)}'
