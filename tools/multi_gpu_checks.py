"""Collective + multi-rank engine checks on N GPUs of one box (collective and
end-to-end tiers).  Spawns one process per GPU with pmdist.parallel.launch and writes one JSON per
rank under --out.

    python tools/multi_gpu_checks.py --gpus 2 [--out gpurun_out/multi2]
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _worker(rank: int, world: int, out_dir: str, quick: bool) -> None:
    import ctypes as C

    import torch
    from pmdist.ops import _lib
    from pmdist.parallel.peer import ALGOS, PeerComm
    torch.cuda.set_device(rank)
    dev = torch.device('cuda', rank)
    res = {'rank': rank, 'world': world}
    comm = PeerComm(rank, world, dev, heap_bytes=(1 << 30) if quick else (3 << 29))
    res['transport'] = comm.describe()
    lib = _lib.lib()

    # ---- plain all-reduce, every algorithm, several sizes; result must be identical on every rank
    algos = ['oneshot', 'twoshot'] + (['nvls'] if comm.info.multicast_mapped else [])
    ar = {}
    for algo in algos:
        # quick: up to 20 MB; full: 1 KB .. 256 MB (collective tier)
        for n in ((256, 4096, 1 << 20, 5 << 20) if quick else (256, 4096, 65536, 1 << 20, 5 << 20, 1 << 24, 1 << 26)):
            g = torch.Generator(device='cuda').manual_seed(1000 * rank + 7)
            x = torch.randn(n, device=dev, generator=g)
            want = torch.zeros(n, device=dev)
            for r in range(world):
                gr = torch.Generator(device='cuda').manual_seed(1000 * r + 7)
                want += torch.randn(n, device=dev, generator=gr)
            got = x.clone()
            comm.all_reduce_sum_(got, algo=algo)
            torch.cuda.synchronize()
            err = (got - want).abs().max().item()
            ar[f'{algo}_{n}'] = {'max_abs_err': err, 'fingerprint': float(got.double().sum())}
    res['allreduce'] = ar
    res['allreduce_ok'] = all(v['max_abs_err'] < 1e-4 for v in ar.values())

    # ---- broadcast
    t = torch.full((100003,), float(rank + 1), device=dev)
    comm.broadcast_(t, src=0)
    torch.cuda.synchronize()
    res['broadcast_ok'] = bool((t == 1.0).all())

    # ---- flag / epoch reuse: many back-to-back device barriers and small all-reduces on one slot
    iters = 300 if quick else 10000
    t0 = time.time()
    v = torch.ones(256, device=dev)
    for _ in range(iters):
        comm.device_barrier()
    for _ in range(iters // 10):
        comm.all_reduce_sum_(v, algo='oneshot')
        v.div_(world)
    torch.cuda.synchronize()
    res['flag_reuse_ok'] = bool(torch.allclose(v, torch.ones_like(v)))
    res['flag_reuse_seconds'] = round(time.time() - t0, 2)

    # ---- metric reduction
    m = comm.reduce_metrics([1.5 * (rank + 1), rank, 2.0, 10.0])
    res['metric_ok'] = m == [1.5 * sum(range(1, world + 1)), float(sum(range(world))), 2.0 * world, 10.0 * world]

    # ---- multi-rank engine: weights bitwise equal on all ranks; W ranks x B/W == 1 rank x B
    from pmdist.data.loader import Batch, draw_aug_params
    from pmdist.data.sources import synthetic_source
    from pmdist.models.resnet import build_resnet
    from pmdist.parallel.comm import SingleComm
    from pmdist.parallel.engine import NativeEngine
    from pmdist.utils.schedule import SGDConfig
    GB = (16 if quick else 512) * world          # full: the bench configuration, 512 images per rank
    src = synthetic_source(GB * 3, 32, seed=11)
    eng_res = {}
    for algo in algos + ['auto']:               # auto = measured table per bucket + push all-reduce for the tail bucket
        torch.manual_seed(100 + rank)            # deliberately different initial weights per rank (D12)
        model = build_resnet(18)
        eng = NativeEngine(model, comm, SGDConfig(), dev, GB // world, 32, algo=algo)
        fps, losses = [eng.weights_fingerprint()], []
        for s in range(3):
            idx = torch.arange(s * GB + rank, (s + 1) * GB, world)       # DistributedSampler-style shard
            b = Batch(src.images[idx], src.labels[idx], torch.ones(len(idx), dtype=torch.bool),
                      draw_aug_params(len(idx), False, None), True)
            eng.train_step(b, 0.05)
            fps.append(eng.weights_fingerprint())
            mm = eng.read_metrics(reduce=True)
            losses.append(mm[0] / mm[3])
        sd = {k: v.clone() for k, v in eng.model_state().items()}
        opt = eng.optimizer_state()
        eng_res[algo] = {'fingerprints': fps, 'losses': losses,
                         'bn_rm': float(sd['bn1.running_mean'].double().sum()),
                         'mom_fp': float(sum(v.double().sum() for v in opt['momentum'].values()))}
        eng.close()
    res['engine'] = eng_res

    if rank == 0:
        # oracle: the same three global batches on ONE rank (same initial weights as rank 0)
        torch.manual_seed(100)
        model = build_resnet(18)
        eng = NativeEngine(model, SingleComm(), SGDConfig(), dev, GB, 32)
        fps, losses = [eng.weights_fingerprint()], []
        for s in range(3):
            idx = torch.arange(s * GB, (s + 1) * GB)
            b = Batch(src.images[idx], src.labels[idx], torch.ones(GB, dtype=torch.bool),
                      draw_aug_params(GB, False, None), True)
            eng.train_step(b, 0.05)
            fps.append(eng.weights_fingerprint())
            mm = eng.read_metrics()
            losses.append(mm[0] / mm[3])
        sd1 = eng.model_state()
        res['single_rank_oracle'] = {'fingerprints': fps, 'losses': losses,
                                     'bn_rm': float(sd1['bn1.running_mean'].double().sum())}
        eng.close()
    comm.barrier()
    comm.close()
    with open(os.path.join(out_dir, f'rank{rank}.json'), 'w') as f:
        json.dump(res, f, indent=1)


def summarize(out_dir: str, world: int) -> dict:
    rs = [json.load(open(os.path.join(out_dir, f'rank{r}.json'))) for r in range(world)]
    ok = all(r['allreduce_ok'] and r['broadcast_ok'] and r['flag_reuse_ok'] and r['metric_ok'] for r in rs)
    summary = {'world': world, 'transport': rs[0]['transport'], 'collectives_ok': ok, 'algos': {}}
    # all-reduce results bitwise identical on every rank
    for k in rs[0]['allreduce']:
        same = len({r['allreduce'][k]['fingerprint'] for r in rs}) == 1
        ok = ok and same
    summary['allreduce_identical_on_all_ranks'] = ok
    oracle = rs[0]['single_rank_oracle']
    for algo in rs[0]['engine']:
        fps = [r['engine'][algo]['fingerprints'] for r in rs]
        same = all(f == fps[0] for f in fps)
        moved = len(set(fps[0])) == len(fps[0])
        l_multi, l_one = rs[0]['engine'][algo]['losses'], oracle['losses']
        close = all(abs(a - b) < 0.03 for a, b in zip(l_multi, l_one))
        fp_close = all(abs(a - b) < 1e-2 * max(1.0, abs(b)) for a, b in zip(fps[0], oracle['fingerprints']))
        bn_same = len({r['engine'][algo]['bn_rm'] for r in rs}) == 1
        summary['algos'][algo] = {'weights_bitwise_equal_across_ranks': same, 'weights_moved': moved,
                                  'loss_matches_single_rank': close, 'fingerprint_close_to_single_rank': fp_close,
                                  'running_stats_equal_across_ranks': bn_same, 'losses': l_multi, 'losses_single': l_one,
                                  'fingerprints': fps[0], 'fingerprints_single': oracle['fingerprints']}
        ok = ok and same and moved and close and fp_close and bn_same
    summary['ok'] = ok
    return summary


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=2)
    ap.add_argument('--out', default=os.path.join(ROOT, 'gpurun_out', 'multi'))
    ap.add_argument('--quick', action='store_true')
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    from pmdist.parallel import launch
    launch.spawn(_worker, a.gpus, args=(a.out, a.quick), timeout=900)
    s = summarize(a.out, a.gpus)
    with open(os.path.join(a.out, 'summary.json'), 'w') as f:
        json.dump(s, f, indent=1)
    print(json.dumps(s, indent=1))
    sys.exit(0 if s['ok'] else 1)


if __name__ == '__main__':
    main()
