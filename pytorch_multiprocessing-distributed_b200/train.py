"""Per-rank trainer and the job entry point.

Parity targets:
  * ``run_model``  (ref:main.py:180-188): create save_path, snapshot the launching script into it,
    spawn ``world_size`` ranks, join.
  * ``main(rank, world_size)`` (ref:main.py:32-84): bring-up, loaders, model, engine, loggers,
    epoch loop (LR schedule -> train -> validate -> checkpoint), rank-0 plot, teardown.
  * ``train`` / ``validate`` (ref:main.py:87-171) incl. the console line formats
    (ref:main.py:119-127, 162-170) and the per-epoch log rows (ref:main.py:129-130, 169-171).

Behavioural fixes relative to the reference (``--compat-metrics`` restores the
reference numbers for A/B): metrics are reduced over ranks (D1, D2), the LR table is evaluated on
every rank (D3), the sampler reshuffles per epoch (D4), evaluation ignores wrap-around duplicates
(D5), top-5 is computed and logged (D6, D7), loss/accuracy are accumulated on the device and read
once per print interval instead of two ``.item()`` syncs per step (D13).
"""
from __future__ import annotations

import os
import shutil
import sys
import time
from typing import Optional

import torch

from .cli import parse_args
from .data.loader import DeviceBatchLoader, make_device_loaders, make_loaders
from .data.sources import make_sources
from .models.resnet import model_from_name
from .parallel import launch
from .parallel.comm import make_comm
from .utils import checkpoint as ckpt
from .utils.meters import AverageMeter, Logger
from .utils.schedule import SGDConfig


def gpu_list(spec: str, world_size: int):
    """``--gpu 4,5,6,7`` with exactly world_size entries names the device of every rank; anything else (the
    reference's default '7' included) is ignored, as the reference ignores the flag entirely (ref:main.py:26)."""
    parts = [p.strip() for p in str(spec).split(',') if p.strip() != '']
    if len(parts) != world_size or not all(p.isdigit() for p in parts) or len(set(parts)) != len(parts):
        return None
    return [int(p) for p in parts]


def resolve_device(args, rank: int) -> torch.device:
    want = args.device
    if want == 'auto':
        want = 'cuda' if torch.cuda.is_available() else 'cpu'
    if want == 'cuda':
        n = max(torch.cuda.device_count(), 1)
        ids = gpu_list(getattr(args, 'gpu', ''), int(getattr(args, 'world_size', 0) or 0))
        index = ids[rank] if ids is not None and all(i < n for i in ids) else rank % n
        torch.cuda.set_device(index)
        return torch.device('cuda', index)
    return torch.device('cpu')


def apply_deterministic(args) -> None:
    """--deterministic: the same command line twice gives bit-identical checkpoints (the reference leaves
    cudnn.benchmark on and is not reproducible, ref:main.py:45).  Only the torch-autograd engine qualifies: the native
    kernels reduce split-K and statistics partials with fp32 atomics, whose order varies from run to run."""
    if not getattr(args, 'deterministic', False):
        return
    if args.engine == 'native':
        raise SystemExit('--deterministic needs the torch-autograd engine (--engine host or auto)')
    args.engine = 'host'
    if args.backend in ('auto', 'p2p'):
        args.backend = 'auto_host'
    os.environ.setdefault('CUBLAS_WORKSPACE_CONFIG', ':4096:8')      # before the first cuBLAS call of this process
    torch.use_deterministic_algorithms(True, warn_only=True)
    torch.backends.cudnn.deterministic = True
    torch.backends.cudnn.benchmark = False


def build_engine(args, model, comm, device, sgd):
    engine = args.engine
    if engine == 'auto':
        engine = 'native' if device.type == 'cuda' else 'host'
    if engine == 'native':
        from .parallel.engine import NativeEngine
        return NativeEngine(model, comm, sgd, device, per_rank_batch=max(1, int(args.batch_size / comm.world_size)),
                            image_size=args.image_size, sync_bn=not args.no_sync_bn,
                            use_graph=not args.no_graph, bucket_mb=args.bucket_mb, algo=args.algo)
    from .parallel.host_engine import HostEngine
    kw = {}
    if args.bucket_mb > 0:
        kw = dict(first_bucket_bytes=int(args.bucket_mb * (1 << 20)), bucket_bytes=int(args.bucket_mb * (1 << 20)))
    return HostEngine(model, comm, sgd, device, sync_bn=not args.no_sync_bn, **kw)


def _make_comm(args, rank, world_size, device):
    backend = args.backend
    if backend == 'auto_host':        # --deterministic: library all-reduce (fixed ring order), no peer-memory kernels
        backend = ('nccl' if device.type == 'cuda' else 'gloo') if world_size > 1 else 'single'
    if backend == 'auto':
        backend = ('p2p' if device.type == 'cuda' else 'gloo') if world_size > 1 else 'single'
    if backend == 'p2p':
        from .models.resnet import MODEL_ALIASES, resnet_spec
        from .parallel.engine import heap_bytes_for
        from .parallel.peer import PeerComm
        depth = MODEL_ALIASES.get(args.model.lower(), 18)
        return PeerComm(rank, world_size, device, heap_bytes_for(resnet_spec(depth, args.num_classes).num_parameters()))
    if world_size == 1 and backend in ('gloo', 'single'):
        backend = 'single'
    return make_comm(rank, world_size, backend, device)


def train_epoch(engine, loader, epoch: int, lr: float, args, rank: int, logger: Optional[Logger], comm) -> list:
    """ref:main.py:87-130."""
    batch_time, data_time, losses, top1 = AverageMeter(), AverageMeter(), AverageMeter(), AverageMeter()
    tot = [0.0, 0.0, 0.0, 0.0]
    end = time.time()
    pending_n = 0
    nb = len(loader)
    if isinstance(loader, DeviceBatchLoader):
        engine.attach_dataset(loader.source)
    for i, batch in enumerate(loader):
        if args.max_steps and i >= args.max_steps:
            break
        data_time.update(time.time() - end)
        engine.train_step(batch, lr)
        pending_n += batch.n
        if i % args.print_freq == 0 or i == nb - 1:
            m = engine.read_metrics(reset=True)          # the only host sync in the loop
            if m[3] > 0:
                losses.update(m[0] / m[3], int(m[3]))
                top1.update(100.0 * m[1] / m[3], int(m[3]))
            tot = [a + b for a, b in zip(tot, m)]
            batch_time.update(time.time() - end)
            if rank == 0 and i % args.print_freq == 0:
                print('Epoch: [{0}][{1}/{2}]\t'
                      'Time {batch_time.val:.3f} ({batch_time.avg:.3f})\t'
                      'Data {data_time.val:.3f} ({data_time.avg:.3f})\t'
                      'Loss {loss.val:.4f} ({loss.avg:.4f})\t'
                      'Prec {top1.val:.3f}% ({top1.avg:.3f}%)'.format(
                          epoch, i, nb, batch_time=batch_time, data_time=data_time, loss=losses, top1=top1),
                      flush=True)
        else:
            batch_time.update(time.time() - end)
        end = time.time()
    m = engine.read_metrics(reset=True)
    tot = [a + b for a, b in zip(tot, m)]
    g = tot if args.compat_metrics else comm.reduce_metrics(tot)
    n = max(g[3], 1.0)
    row = [epoch, g[0] / n, 100.0 * g[1] / n, 100.0 * g[2] / n]
    if rank == 0 and logger is not None:
        logger.write(row[:args.log_columns])
    return row


def validate(engine, loader, epoch: int, args, rank: int, logger: Optional[Logger], comm, mode: str = 'test') -> list:
    """ref:main.py:134-171."""
    batch_time, losses = AverageMeter(), AverageMeter()
    tot = [0.0, 0.0, 0.0, 0.0]
    end = time.time()
    nb = len(loader)
    if isinstance(loader, DeviceBatchLoader):
        engine.attach_dataset(loader.source)
    for i, batch in enumerate(loader):
        if args.max_steps and i >= args.max_steps:
            break
        engine.eval_step(batch)
        if i % args.print_freq == 0 or i == nb - 1:
            m = engine.read_metrics(reset=True)
            if m[3] > 0:
                losses.update(m[0] / m[3], int(m[3]))
            tot = [a + b for a, b in zip(tot, m)]
            batch_time.update(time.time() - end)
            if rank == 0 and i % args.print_freq == 0:
                print(mode, ': [{0}/{1}]\t'
                            'Time {batch_time.val:.3f} ({batch_time.avg:.3f})\t'
                            'Loss {loss.val:.4f} ({loss.avg:.4f})'.format(
                                i, nb, batch_time=batch_time, loss=losses), flush=True)
        else:
            batch_time.update(time.time() - end)
        end = time.time()
    m = engine.read_metrics(reset=True)
    tot = [a + b for a, b in zip(tot, m)]
    if args.compat_metrics:
        # the reference divides rank 0's local hit count by the FULL dataset (ref:main.py:168)
        n = max(tot[3], 1.0)
        row = [epoch, tot[0] / n, 100.0 * tot[1] / loader.dataset_len, 100.0 * tot[2] / loader.dataset_len]
    else:
        g = comm.reduce_metrics(tot)
        n = max(g[3], 1.0)
        row = [epoch, g[0] / n, 100.0 * g[1] / n, 100.0 * g[2] / n]
    if rank == 0:
        print('Accuracy {:.2f}'.format(row[2]), flush=True)
        if logger is not None:
            logger.write(row[:args.log_columns])
    return row


def worker(rank: int, world_size: int, args) -> None:
    """ref:main.py:32-84, one call per rank."""
    apply_deterministic(args)
    torch.manual_seed(args.seed)
    device = resolve_device(args, rank)
    if device.type == 'cpu':
        try:
            torch.set_num_threads(max(1, len(os.sched_getaffinity(0))))
        except (AttributeError, RuntimeError):
            pass
    comm = _make_comm(args, rank, world_size, device)
    engine = None
    try:
        train_src, test_src = make_sources(args.data, args.data_root, args.image_size, args.train_size,
                                           args.test_size, args.num_classes, args.seed)
        if int(args.batch_size / world_size) * world_size != args.batch_size and rank == 0:
            print(f'warning: batch_size {args.batch_size} is not divisible by world_size {world_size}; '
                  f'per-rank batch is int({args.batch_size}/{world_size}) as in the reference', flush=True)
        if rank == 0:
            print("-------------------Make loader-------------------")
            print('Train Dataset :', len(train_src), '   Test Dataset :', len(test_src), flush=True)

        model = model_from_name(args.model, train_src.num_classes)
        sgd = SGDConfig(args.lr, args.momentum, args.weight_decay, not args.no_nesterov,
                        tuple(args.milestones), args.gamma)
        start_epoch = 1
        resume_state = None
        if args.init_from:
            model.load_state_dict(ckpt.load_model(args.init_from))
        if args.resume:
            mpth, spth, e = ckpt.find_resume(args.resume)
            model.load_state_dict(ckpt.load_model(mpth))
            start_epoch = e + 1
            if spth:
                resume_state = ckpt.load_train_state(spth)
        engine = build_engine(args, model, comm, device, sgd)
        if engine.kind == 'native' and not args.host_dataset:
            # both datasets live in GPU memory (CIFAR-10: 180 MB of 180 GB); a step's input is its index vector
            train_loader, test_loader = make_device_loaders(train_src, test_src, args.batch_size, world_size, rank,
                                                            args.seed, compat_eval_shuffle=args.compat_metrics,
                                                            augment=not args.no_augment)
        else:
            pin = device.type == 'cuda'
            if pin:
                train_src.pin(); test_src.pin()
            train_loader, test_loader = make_loaders(train_src, test_src, args.batch_size, world_size, rank,
                                                     args.seed, pin, compat_eval_shuffle=args.compat_metrics,
                                                     augment=not args.no_augment)
        if resume_state is not None:
            engine.load_optimizer_state(resume_state['optimizer'])

        train_logger = Logger(os.path.join(args.save_path, 'train.log'))
        test_logger = Logger(os.path.join(args.save_path, 'test.log'))
        if rank == 0 and not args.resume:
            train_logger.truncate(); test_logger.truncate()

        for epoch in range(start_epoch, args.epochs + 1):
            lr = sgd.lr_at(epoch)
            train_loader.set_epoch(0 if args.no_reshuffle else epoch)
            test_loader.set_epoch(0)
            train_epoch(engine, train_loader, epoch, lr, args, rank, train_logger, comm)
            validate(engine, test_loader, epoch, args, rank, test_logger, comm, 'test')
            last = epoch == int(args.epochs)
            if last or (args.save_every and epoch % args.save_every == 0):
                opt_state = engine.optimizer_state()       # collective: momentum slices live on their owner rank
                if rank == 0:
                    ckpt.save_model(args.save_path, epoch, engine.model_state())
                    ckpt.save_train_state(args.save_path, epoch, opt_state,
                                          {'args': {k: v for k, v in vars(args).items()}})
        if rank == 0 and not args.no_plot and args.epochs >= start_epoch:
            from .utils import plot_curves
            plot_curves.draw_plot(args.save_path)
        comm.barrier()
    finally:
        if engine is not None:
            engine.close()
        comm.close()


def run_model(args, script: Optional[str] = None) -> None:
    """ref:main.py:180-188."""
    os.makedirs(args.save_path, exist_ok=True)
    script = script or sys.argv[0]
    if script and os.path.isfile(script):
        dst = os.path.join(args.save_path, 'main.py')
        if os.path.exists(dst):
            os.chmod(dst, 0o644)          # the reference's copy inherits r--r--r-- and blocks reruns
        shutil.copy(script, dst)
    if os.environ.get('WORLD_SIZE') and os.environ.get('RANK') is not None and 'TORCHELASTIC_RUN_ID' in os.environ:
        # already launched one-process-per-GPU by torchrun: be a rank, not a launcher
        worker(int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), args)
        return
    launch.spawn(worker, args.world_size, args=(args,), timeout=args.timeout or None)


def main(argv=None) -> None:
    os.environ['PYTHONWARNINGS'] = 'ignore:semaphore_tracker:UserWarning'
    run_model(parse_args(argv))


if __name__ == '__main__':
    main()
