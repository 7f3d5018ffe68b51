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


def measure_bucket_algos(comm, bucket_bytes: List[int], multicast: bool, iters: int = 20):
    """Bring-up microbenchmark (SURVEY.md N2'): times every all-reduce algorithm x CTA count on THIS box for the
    model's real bucket sizes (CUDA events, device barrier before, max over ranks) and returns the fastest per
    bucket.  Used when no committed table matches the world size (or PMDIST_TUNE=1).  ~50 ms for ResNet18."""
    import ctypes as C
    import struct

    import torch

    from ..ops import _lib
    from ..ops._lib import check
    from .peer import ALGOS
    big = max(bucket_bytes)
    try:
        in_off, out_off = comm.alloc(big + 64), comm.alloc(big + 64)
    except Exception:
        return None                                  # heap too small for scratch: fall back to the table / heuristic
    lib = _lib.lib()
    algos = ['oneshot', 'twoshot'] + (['nvls'] if multicast else [])
    pick_a, pick_b, notes = [], [], []
    cache = {}
    for nbytes in bucket_bytes:
        n = nbytes // 16 * 4
        if n in cache:
            a_, b_, us_ = cache[n]
        else:
            best = None
            for algo in algos:
                for blocks in ((8, 32) if nbytes <= (1 << 20) else (32, 96)):
                    def fn():
                        check(lib.pmd_allreduce_f32(C.byref(comm.sym), C.c_ulonglong(in_off), C.c_ulonglong(out_off),
                                                    C.c_longlong(n), C.c_int(ALGOS[algo]), C.c_int(blocks), C.c_int(2000),
                                                    _lib.stream_ptr()))
                    for _ in range(3):
                        fn()
                    torch.cuda.synchronize()
                    comm.barrier()
                    comm.device_barrier()
                    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    t0.record()
                    for _ in range(iters):
                        fn()
                    t1.record()
                    torch.cuda.synchronize()
                    us = t0.elapsed_time(t1) * 1e3 / iters
                    us = max(struct.unpack('d', v)[0] for v in comm.all_gather_bytes(struct.pack('d', us)))
                    if best is None or us < best[2]:
                        best = (algo, blocks, us)        # every rank sees the same max, hence the same choice
            a_, b_, us_ = cache[n] = best
        pick_a.append(a_)
        pick_b.append(b_)
        notes.append('%d B: %s x%d %.1f us' % (nbytes, a_, b_, us_))
    return pick_a, pick_b, '; '.join(notes)


def choose_bucket_algos(world: int, bucket_bytes: List[int], multicast: bool, forced: str = 'auto',
                        default_blocks: int = 32, comm=None) -> Tuple[List[str], List[int], str]:
    """Per-bucket (algo, CTAs).  `forced` in {'oneshot','twoshot','nvls'} overrides the table."""
    if world == 1:
        return ['oneshot'] * len(bucket_bytes), [default_blocks] * len(bucket_bytes), 'single rank'
    if forced != 'auto':
        return [forced] * len(bucket_bytes), [default_blocks] * len(bucket_bytes), f'forced {forced}'
    tab = AllreduceTable.load(world)
    if (tab is None or os.environ.get('PMDIST_TUNE', '0') == '1') and comm is not None:
        measured = measure_bucket_algos(comm, bucket_bytes, multicast)
        if measured is not None:
            return measured[0], measured[1], 'measured at bring-up (%s)' % measured[2]
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
