"""Loss-curve parity against the reference ITSELF (end-to-end tier; VERDICT r1 missing #4).

Runs, on the same W GPUs, the same synthetic data set, the same initial weights and the same sample order:
  1. baseline/_ref/main.py (the unmodified reference: torch DDP + SyncBatchNorm + NCCL + cuDNN, fp32) through its own
     CLI under baseline/shim (labelled test-only patches: REF_NO_AUG turns its random crop / flip into identities,
     REF_DUMP_INIT saves the state_dict every rank starts from);
  2. this framework with --engine host --backend nccl (fp32 torch-autograd twin over NCCL: the tight comparison);
  3. this framework with --engine native (bf16 sm_100a kernels, peer-memory transport: the product),
both started from (1)'s initial weights (--init-from), on the reference's data (--data ref-fake), in the reference's
sample order (--no-reshuffle, seed 0) and with its rank-0-local metrics (--compat-metrics --log-columns 3).
Compares the per-epoch train losses in train.log (ref:main.py:129-130 writes [epoch, loss avg, top-1 avg]).

    python tools/parity_vs_reference.py --gpus 2 --epochs 2 --out gpurun_out/parity
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def read_log(path):
    rows = []
    with open(path) as f:
        for ln in f:
            p = ln.split()
            if len(p) >= 3:
                rows.append([int(p[0]), float(p[1]), float(p[2])])
    return rows


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=2)
    ap.add_argument('--epochs', type=int, default=2)
    ap.add_argument('--global-batch', type=int, default=128)
    ap.add_argument('--train-size', type=int, default=2048)
    ap.add_argument('--test-size', type=int, default=256)
    ap.add_argument('--out', default=os.path.join(ROOT, 'gpurun_out', 'parity'))
    ap.add_argument('--tol-host', type=float, default=2e-2)
    ap.add_argument('--tol-native', type=float, default=5e-2)
    a = ap.parse_args()
    a.out = os.path.abspath(a.out)
    os.makedirs(a.out, exist_ok=True)
    init = os.path.join(a.out, 'ref_init.pth')
    res = {'config': vars(a)}

    # ---- 1. the reference
    ref_dir = os.path.join(a.out, 'reference') + '/'
    env = {k: v for k, v in os.environ.items() if not k.startswith('TORCHELASTIC_')}
    env.update({'REF_SHIM': '1', 'PYTHONPATH': os.path.join(ROOT, 'baseline', 'shim') + os.pathsep + env.get('PYTHONPATH', ''),
                'REF_TRAIN_SIZE': str(a.train_size), 'REF_TEST_SIZE': str(a.test_size), 'REF_IMAGE_SIZE': '32',
                'REF_NO_AUG': '1', 'REF_DUMP_INIT': init})
    cmd = [sys.executable, os.path.join(ROOT, 'baseline', '_ref', 'main.py'), '--batch_size', str(a.global_batch),
           '--epochs', str(a.epochs), '--world_size', str(a.gpus), '--save_path', ref_dir, '--print-freq', '1000']
    r = subprocess.run(cmd, cwd=a.out, env=env, capture_output=True, text=True, timeout=1500)
    if r.returncode != 0 or not os.path.exists(init):
        res['error'] = 'reference run failed: ' + (r.stderr or r.stdout)[-1500:]
        print(json.dumps(res, indent=1))
        sys.exit(2)
    res['reference'] = read_log(os.path.join(ref_dir, 'train.log'))

    # ---- 2./3. this framework, fp32 torch twin over NCCL and the native engine
    common = ['--batch_size', str(a.global_batch), '--epochs', str(a.epochs), '--world_size', str(a.gpus), '--data', 'ref-fake',
              '--train-size', str(a.train_size), '--test-size', str(a.test_size), '--init-from', init, '--no-augment',
              '--no-reshuffle', '--seed', '0', '--compat-metrics', '--log-columns', '3', '--print-freq', '1000', '--no-plot']
    for name, extra in (('host_fp32_nccl', ['--engine', 'host', '--backend', 'nccl']), ('native_bf16', ['--engine', 'native'])):
        d = os.path.join(a.out, name) + '/'
        r = subprocess.run([sys.executable, os.path.join(ROOT, 'main.py')] + common + extra + ['--save_path', d],
                           cwd=ROOT, capture_output=True, text=True, timeout=1500)
        if r.returncode != 0:
            res['error'] = f'{name} failed: ' + (r.stderr or r.stdout)[-1500:]
            print(json.dumps(res, indent=1))
            sys.exit(2)
        res[name] = read_log(os.path.join(d, 'train.log'))

    ref = res['reference']
    ok = True
    for name, tol in (('host_fp32_nccl', a.tol_host), ('native_bf16', a.tol_native)):
        deltas = [abs(x[1] - y[1]) for x, y in zip(res[name], ref)]
        res[name + '_abs_loss_delta_per_epoch'] = deltas
        res[name + '_ok'] = len(res[name]) == len(ref) == a.epochs and all(d <= tol for d in deltas)
        ok = ok and res[name + '_ok']
    res['ok'] = ok
    with open(os.path.join(a.out, 'summary.json'), 'w') as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res, indent=1))
    sys.exit(0 if ok else 1)


if __name__ == '__main__':
    main()
