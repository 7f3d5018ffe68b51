"""Gradient all-reduce bucket-size sweep (BASELINE.json config 5): hand-written P2P kernels
(one-shot / two-shot / NVLS multimem) vs NCCL on the same box, 1 KB .. 256 MB plus the real sizes of
the reference model (SyncBN 512 B - 4 KB, DDP buckets 9.6 / 10.0 / 19.6 MB).

Timing: CUDA events on the launching stream around `iters` back-to-back calls after warm-up, device
barrier before, max over ranks.  Bus bandwidth = 2*(W-1)/W * bytes / time (NCCL's convention).

    python tools/allreduce_sweep.py --gpus 8 --out gpurun_out/sweep8.json
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import struct
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SIZES = [1 << 10, 4 << 10, 16 << 10, 64 << 10, 256 << 10, 1 << 20, 4 << 20, 9_622_784, 9_990_184, 16 << 20,
         19_612_968, 64 << 20, 256 << 20]


def _worker(rank: int, world: int, out: str, max_bytes: int, nvlink_proof: bool = False) -> None:
    import torch
    import torch.distributed as dist
    from pmdist.ops import _lib
    from pmdist.ops._lib import check
    from pmdist.parallel.peer import ALGOS, PeerComm
    torch.cuda.set_device(rank)
    dev = torch.device('cuda', rank)
    sizes = [s for s in SIZES if s <= max_bytes]
    big = max(sizes)
    comm = PeerComm(rank, world, dev, heap_bytes=2 * big + (64 << 20))
    lib = _lib.lib()
    in_off = comm.alloc(big + 64)
    out_off = comm.alloc(big + 64)
    comm.view(in_off, (big // 4,), torch.float32).normal_()
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    nccl_buf = torch.randn(big // 4, device=dev)
    algos = ['oneshot', 'twoshot'] + (['nvls'] if comm.info.multicast_mapped else [])

    def time_it_n(fn, iters):
        for _ in range(10):
            fn()
        torch.cuda.synchronize()
        comm.barrier()
        comm.device_barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        torch.cuda.synchronize()
        us = a.elapsed_time(b) * 1e3 / iters
        return max(struct.unpack('d', v)[0] for v in comm.all_gather_bytes(struct.pack('d', us)))

    rows = []
    for nbytes in sizes:
        n = nbytes // 4 // 4 * 4
        iters = 200 if nbytes <= (1 << 20) else (50 if nbytes <= (32 << 20) else 10)
        row = {'bytes': n * 4}

        def time_it(fn):
            for _ in range(5):
                fn()
            torch.cuda.synchronize()
            comm.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(iters):
                fn()
            b.record()
            torch.cuda.synchronize()
            us = a.elapsed_time(b) * 1e3 / iters
            allv = comm.all_gather_bytes(struct.pack('d', us))
            return max(struct.unpack('d', v)[0] for v in allv)

        for algo in algos:
            for blocks in ((8, 32) if nbytes <= (1 << 20) else (32, 96)):
                def fn(algo=algo, blocks=blocks):
                    check(lib.pmd_allreduce_f32(C.byref(comm.sym), C.c_ulonglong(in_off), C.c_ulonglong(out_off),
                                                C.c_longlong(n), C.c_int(ALGOS[algo]), C.c_int(blocks), C.c_int(2000),
                                                _lib.stream_ptr()))
                us = time_it(fn)
                key = f'{algo}_b{blocks}'
                row[key + '_us'] = round(us, 2)
                row[key + '_busbw_GBs'] = round(2 * (world - 1) / world * n * 4 / us / 1e3, 1)
        view = nccl_buf[:n]
        us = time_it(lambda: dist.all_reduce(view))
        row['nccl_us'] = round(us, 2)
        row['nccl_busbw_GBs'] = round(2 * (world - 1) / world * n * 4 / us / 1e3, 1)
        best = min((v, k) for k, v in row.items() if k.endswith('_us') and not k.startswith('nccl'))
        row['best_p2p'] = best[1][:-3]
        row['speedup_vs_nccl'] = round(row['nccl_us'] / best[0], 2)
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    # ---- SyncBatchNorm push exchange (csrc/comm.cu small_allreduce_push_kernel): the [2C] vectors of the model's
    # BatchNorm layers, one and two vectors per launch, back-to-back launches (PDL), device time per exchange
    class SmallX(C.Structure):
        _fields_ = [('local', C.c_void_p), ('total', C.c_void_p), ('recv_off', C.c_ulonglong), ('n', C.c_int),
                    ('slot', C.c_int)]
    push_rows = []
    for c in (64, 128, 256, 512, 2048):
        n = 2 * c
        loc = [torch.randn(n, device=dev) for _ in range(2)]
        tot = [torch.empty(n, device=dev) for _ in range(2)]
        recv = [comm.alloc(2 * world * n * 8) for _ in range(2)]
        for off in recv:
            comm.view(off, (2 * world * n * 2,), torch.int32).zero_()
        torch.cuda.synchronize()
        comm.barrier()
        xs = (SmallX * 2)(SmallX(loc[0].data_ptr(), tot[0].data_ptr(), recv[0], n, 3000 + 2 * c % 97),
                          SmallX(loc[1].data_ptr(), tot[1].data_ptr(), recv[1], n, 3100 + 2 * c % 97))
        row = {'channels': c, 'bytes_per_rank': n * 4}
        iters = 500
        for nvec in (1, 2):
            us = time_it_n(lambda nvec=nvec: check(lib.pmd_small_allreduce_push(C.byref(comm.sym), xs, C.c_int(nvec),
                                                                               _lib.stream_ptr())), iters)
            row[f'push_{nvec}vec_us'] = round(us, 2)
        view = nccl_buf[:n]
        row['nccl_allreduce_us'] = round(time_it_n(lambda: dist.all_reduce(view), iters), 2)
        push_rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    # ---- NVLink-side proof: the driver's per-link byte counters of THIS rank's GPU around 1000 launches of the
    # fused-path collective on the whole ResNet18 gradient slab (19.6 MB); expected traffic per GPU = 2 S (W-1)/W
    # for the two-shot kernel (NVLS: S (W-1)/W up for the in-switch reduction + the multicast stores)
    proof = None
    if nvlink_proof:
        import subprocess

        def counters():
            r = subprocess.run(['nvidia-smi', 'nvlink', '-gt', 'd', '-i', str(rank)], capture_output=True, text=True)
            tx = rx = 0
            for ln in r.stdout.splitlines():
                p = ln.replace(':', ' ').split()
                if 'Tx' in p and 'KiB' in p:
                    tx += int(p[p.index('KiB') - 1])
                if 'Rx' in p and 'KiB' in p:
                    rx += int(p[p.index('KiB') - 1])
            return tx, rx, r.stdout[-400:]
        n = 19_612_960 // 16 * 4
        proof = {'bytes': n * 4, 'launches': 1000}
        for algo in algos[1:]:                      # twoshot, nvls
            torch.cuda.synchronize()
            comm.barrier()
            t0, r0, _ = counters()
            for _ in range(1000):
                check(lib.pmd_allreduce_f32(C.byref(comm.sym), C.c_ulonglong(in_off), C.c_ulonglong(out_off), C.c_longlong(n),
                                            C.c_int(ALGOS[algo]), C.c_int(32), C.c_int(2000), _lib.stream_ptr()))
            torch.cuda.synchronize()
            comm.barrier()
            t1, r1, tail = counters()
            proof[algo] = {'tx_KiB': t1 - t0, 'rx_KiB': r1 - r0,
                           'tx_bytes_per_launch': (t1 - t0) * 1024 / 1000, 'rx_bytes_per_launch': (r1 - r0) * 1024 / 1000,
                           'expected_two_shot_bytes_per_launch_each_way': 2 * n * 4 * (world - 1) / world}
            if t1 == t0:
                proof[algo]['raw_tail'] = tail
    if rank == 0:
        with open(out, 'w') as f:
            json.dump({'world': world, 'transport': comm.describe(), 'rows': rows, 'push_exchange': push_rows,
                       'nvlink_counters_rank0': proof}, f, indent=1)
    dist.destroy_process_group()
    comm.close()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=2)
    ap.add_argument('--max-mb', type=int, default=256)
    ap.add_argument('--out', default=os.path.join(ROOT, 'gpurun_out', 'allreduce_sweep.json'))
    ap.add_argument('--nvlink-proof', action='store_true', help='nvidia-smi nvlink byte counters around 1000 launches')
    a = ap.parse_args()
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    from pmdist.parallel import launch
    launch.spawn(_worker, a.gpus, args=(a.out, a.max_mb << 20, a.nvlink_proof), timeout=1200)


if __name__ == '__main__':
    main()
