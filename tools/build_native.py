"""Builds libpmdist.so (all of csrc/) for sm_100a with nvcc, in-tree.

    python tools/build_native.py [--force] [--verbose]

Output: pytorch_multiprocessing-distributed_b200/_lib/libpmdist.so (git-ignored; travels to the
GPU box with the gpurun snapshot).  Objects are cached under build/ and recompiled when a source
or any header is newer.  No torch headers are involved: the library exposes a plain C ABI
(include/pmdist.h) that the Python package binds with ctypes, so a full rebuild takes seconds.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import glob
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, 'csrc')
OUT_DIR = os.path.join(ROOT, 'pytorch_multiprocessing-distributed_b200', '_lib')
OUT = os.path.join(OUT_DIR, 'libpmdist.so')
OBJ_DIR = os.path.join(ROOT, 'build', 'obj')

NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '--extended-lambda', '-Xcompiler', '-fPIC', '-Xcompiler', '-fvisibility=hidden',
              '-Xcompiler', '-Wno-unused-function', '-DPMD_BUILD']


def find_nvcc() -> str:
    for cand in (os.environ.get('NVCC'), shutil.which('nvcc'), '/usr/local/cuda/bin/nvcc'):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError('nvcc not found (set NVCC=/path/to/nvcc)')


def sources():
    return sorted(glob.glob(os.path.join(CSRC, '*.cu')) + glob.glob(os.path.join(CSRC, '*.cpp')))


def headers_mtime() -> float:
    hs = glob.glob(os.path.join(CSRC, '*.cuh')) + glob.glob(os.path.join(CSRC, '*.h')) + \
        glob.glob(os.path.join(ROOT, 'include', '*.h')) + [os.path.abspath(__file__)]
    return max(os.path.getmtime(h) for h in hs)


def build(force: bool = False, verbose: bool = False) -> str:
    nvcc = find_nvcc()
    os.makedirs(OBJ_DIR, exist_ok=True)
    os.makedirs(OUT_DIR, exist_ok=True)
    hm = headers_mtime()
    jobs = []
    objs = []
    for src in sources():
        obj = os.path.join(OBJ_DIR, os.path.basename(src) + '.o')
        objs.append(obj)
        stale = force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hm)
        if stale:
            cmd = [nvcc] + NVCC_FLAGS + (['-x', 'cu'] if src.endswith('.cpp') else []) + ['-c', src, '-o', obj]
            jobs.append((src, cmd))

    def run(job):
        src, cmd = job
        if verbose:
            print(' '.join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        return src, r.returncode, r.stdout + r.stderr

    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for src, rc, out in ex.map(run, jobs):
                if rc != 0:
                    raise RuntimeError(f'nvcc failed for {src}:\n{out}')
                if verbose and out.strip():
                    print(out)
    need_link = bool(jobs) or not os.path.exists(OUT) or any(os.path.getmtime(o) > os.path.getmtime(OUT) for o in objs)
    if need_link:
        cmd = [nvcc, '-shared', '-gencode', 'arch=compute_100a,code=sm_100a', '-o', OUT] + objs + \
              ['-Xcompiler', '-fPIC', '-cudart', 'static', '-lpthread', '-ldl', '-lrt']
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError('link failed:\n' + r.stdout + r.stderr)
    return OUT


if __name__ == '__main__':
    ap = argparse.ArgumentParser()
    ap.add_argument('--force', action='store_true')
    ap.add_argument('--verbose', action='store_true')
    a = ap.parse_args()
    print(build(a.force, a.verbose))
