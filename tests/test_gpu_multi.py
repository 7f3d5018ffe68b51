"""Collective and multi-rank end-to-end tiers (need >= 2 B200s on one box): P2P all-reduce (one-shot /
two-shot / NVLS) vs the exact sum, bitwise-identical results on every rank, flag reuse across thousands
of iterations, and W-rank training == single-rank training on the same global batches."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.multigpu
def test_two_rank_collectives_and_engine(tmp_path):
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'tools', 'multi_gpu_checks.py'), '--gpus', '2', '--quick',
                        '--out', str(tmp_path)], capture_output=True, text=True, timeout=1200)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    s = json.load(open(tmp_path / 'summary.json'))
    assert s['ok'], s
