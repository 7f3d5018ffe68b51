"""Runs every case of tools/gpu_cases.py in its own subprocess (a trapping kernel poisons its CUDA
context; isolation keeps the other cases meaningful) and writes gpurun_out/probe.json.

    python tools/gpu_probe.py [--filter SUBSTR] [--timeout 90]
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tools'))


def run_one(name: str) -> None:
    import torch
    import gpu_cases
    t0 = time.time()
    try:
        res = gpu_cases.CASES[name]()
        torch.cuda.synchronize()
    except Exception as e:                      # noqa: BLE001
        res = {'ok': False, 'exception': f'{type(e).__name__}: {e}'}
    res['seconds'] = round(time.time() - t0, 2)
    print('PROBE_RESULT ' + json.dumps(res, default=str))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument('--case', default='')
    ap.add_argument('--filter', default='')
    ap.add_argument('--timeout', type=float, default=90.0)
    ap.add_argument('--out', default=os.path.join(ROOT, 'gpurun_out', 'probe.json'))
    a = ap.parse_args()
    if a.case:
        run_one(a.case)
        return
    import gpu_cases
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    results = {}
    for name in gpu_cases.CASES:
        if a.filter and a.filter not in name:
            continue
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), '--case', name], capture_output=True,
                               text=True, timeout=a.timeout)
            line = [ln for ln in r.stdout.splitlines() if ln.startswith('PROBE_RESULT ')]
            if line:
                res = json.loads(line[-1][len('PROBE_RESULT '):])
            else:
                res = {'ok': False, 'crash': True, 'rc': r.returncode, 'tail': (r.stdout + r.stderr)[-600:]}
        except subprocess.TimeoutExpired:
            res = {'ok': False, 'timeout': True}
        results[name] = res
        flag = 'PASS' if res.get('ok') else 'FAIL'
        brief = {k: v for k, v in res.items() if k in ('rel_to_max', 'max_abs_err', 'exception', 'timeout', 'crash',
                                                       'stats_sum_rel', 'stats_sq_rel', 'tail')}
        print(f'{flag} {name} {json.dumps(brief, default=str)[:400]}', flush=True)
        with open(a.out, 'w') as f:
            json.dump(results, f, indent=1, default=str)
    n_ok = sum(1 for r in results.values() if r.get('ok'))
    print(f'{n_ok}/{len(results)} cases passed')


if __name__ == '__main__':
    main()
