"""Topology discovery for the peer transport (SURVEY.md N2', W3).

NCCL builds a topology graph from NVML/sysfs before it picks rings and trees.  On a single NVSwitch box every
GPU reaches every peer at full NVLink bandwidth, so what the transport must know is much smaller: is every pair
P2P-reachable (otherwise fall back to the NCCL/host path), do atomics work across the link, how many NVLink
links are up, and is multicast (NVLS) offered.  `probe()` collects that from the CUDA runtime (through
libpmdist) and NVML; `python -m pmdist.parallel.topology` prints it.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List

from ..ops import _lib


def probe(max_devices: int = 16) -> Dict:
    lib = _lib.lib()
    n = C.c_int(0)
    acc = (C.c_int * (max_devices * max_devices))()
    rank = (C.c_int * (max_devices * max_devices))()
    atom = (C.c_int * (max_devices * max_devices))()
    mc = (C.c_int * max_devices)()
    if lib.pmd_topology_probe(C.c_int(max_devices), C.byref(n), acc, rank, atom, mc) != 0:
        return {'devices': 0, 'error': lib.pmd_last_error().decode()}
    nd = n.value
    mat = lambda a: [[int(a[i * max_devices + j]) for j in range(nd)] for i in range(nd)]
    out = {'devices': nd, 'can_access_peer': mat(acc), 'p2p_performance_rank': mat(rank), 'p2p_native_atomics': mat(atom),
           'multicast_supported': [int(mc[i]) for i in range(nd)], 'nvlink_links_up': nvlink_links(nd)}
    out['all_pairs_p2p'] = all(all(r) for r in out['can_access_peer']) if nd else False
    return out


def nvlink_links(n_devices: int) -> List[int]:
    """Active NVLink links per GPU (18 on a B200 in an NVSwitch box); -1 where NVML cannot say."""
    try:
        import pynvml
        pynvml.nvmlInit()
    except Exception:
        return [-1] * n_devices
    res = []
    for i in range(n_devices):
        try:
            h = pynvml.nvmlDeviceGetHandleByIndex(i)
            up = 0
            for link in range(32):
                try:
                    if pynvml.nvmlDeviceGetNvLinkState(h, link):
                        up += 1
                except Exception:
                    break
            res.append(up)
        except Exception:
            res.append(-1)
    return res


def describe(p: Dict) -> str:
    if not p.get('devices'):
        return 'no CUDA devices (%s)' % p.get('error', '')
    n = p['devices']
    lines = [f"{n} GPU(s); all pairs P2P: {p['all_pairs_p2p']}; multicast/NVLS: {p['multicast_supported']}; "
             f"NVLink links up: {p['nvlink_links_up']}"]
    lines.append('P2P performance rank (row = from, col = to):')
    for r in p['p2p_performance_rank']:
        lines.append('  ' + ' '.join('%2d' % v for v in r))
    return '\n'.join(lines)


if __name__ == '__main__':
    print(describe(probe()))
