"""Kernel tier: every hand-written sm_100a kernel against a plain PyTorch fp32
reference of the same op.  Cases live in tools/gpu_cases.py (shared with tools/gpu_probe.py)."""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tools'))
import gpu_cases  # noqa: E402


@pytest.mark.gpu
@pytest.mark.parametrize('name', sorted(gpu_cases.CASES))
def test_kernel_matches_torch_reference(name):
    res = gpu_cases.CASES[name]()
    log = os.environ.get('PMD_PROBE_LOG')          # PMD_PROBE_LOG=file: keep every case's error statistics (profiles/)
    if log:
        import json
        with open(log, 'a') as f:
            f.write(json.dumps({'case': name, **res}, default=str) + '\n')
    assert res['ok'], res


@pytest.mark.gpu
def test_native_library_is_loaded_and_counts_launches():
    import torch
    from pmdist.ops import _lib, nn as nk
    before = _lib.launch_count()
    nk.bn_stats(torch.zeros(8, 8, 8, 64, dtype=torch.bfloat16, device='cuda'))
    assert _lib.launch_count() == before + 1
    assert any('libpmdist.so' in ln for ln in open('/proc/self/maps'))


@pytest.mark.gpu
def test_topology_probe_reports_devices():
    from pmdist.parallel.topology import describe, probe
    p = probe()
    assert p['devices'] >= 1 and len(p['can_access_peer']) == p['devices']
    assert all(p['can_access_peer'][i][i] == 1 for i in range(p['devices']))
    assert 'GPU(s)' in describe(p)
