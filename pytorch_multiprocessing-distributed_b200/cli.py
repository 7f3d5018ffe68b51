"""Flag system: a superset of the reference's seven argparse flags.

Parity target: ref:main.py:21-30.  Every reference command line parses unchanged:
``--batch_size`` (global batch, default 64), ``--epochs`` (20), ``--model`` ('res'),
``--save_path`` ('./test/'), ``--gpu`` ('7'; parsed and ignored by the reference; here a comma list with
exactly world_size distinct device ordinals binds rank i to the i-th of them, train.gpu_list, anything else is
ignored), ``--print-freq/-p`` (10),
``--world_size`` (2).

Unlike the reference, flags are parsed inside ``parse_args`` (not at import), travel to the ranks
as a plain dict (not through a re-imported module global, ref:main.py:30), and everything the
reference hard-codes (lr, momentum, wd, milestones, crop padding, dataset root, backend, port) has
a flag with the reference's value as its default.
"""
from __future__ import annotations

import argparse
from typing import List, Optional


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(
        prog='pmdist', description='B200-native multiprocess data-parallel ResNet training')
    # --- the reference's flags, same names/defaults (ref:main.py:22-28)
    p.add_argument('--batch_size', default=64, type=int, help='GLOBAL batch size (split over ranks)')
    p.add_argument('--epochs', default=20, type=int, help='Total number of epochs to run')
    p.add_argument('--model', default='res', type=str,
                   help='res|res18|res34|res50|res101|res152 (res == the reference ResNet18 [1,1,1,1])')
    p.add_argument('--save_path', default='./test/', type=str, help='Savefiles directory')
    p.add_argument('--gpu', default='7', type=str, help='"4,5,6,7" (world_size entries): device of each rank; otherwise ignored like the reference does')
    p.add_argument('--print-freq', '-p', default=10, type=int, metavar='N', help='print frequency')
    p.add_argument('--world_size', default=2, type=int, help='number of GPUs / processes')
    # --- optimiser (reference values hard-coded at ref:main.py:51-59)
    p.add_argument('--lr', default=0.1, type=float)
    p.add_argument('--momentum', default=0.9, type=float)
    p.add_argument('--weight-decay', default=1e-4, type=float)
    p.add_argument('--no-nesterov', action='store_true')
    p.add_argument('--milestones', default='60,80', type=str)
    p.add_argument('--gamma', default=0.1, type=float)
    # --- data (ref:data.py)
    p.add_argument('--data', default='auto', choices=['auto', 'cifar10', 'synthetic', 'ref-fake'],
                   help="auto = CIFAR-10 from --data-root if present, else synthetic")
    p.add_argument('--data-root', default='./cifar10_data', type=str)
    p.add_argument('--image-size', default=32, type=int, help='synthetic image side (32 or 224)')
    p.add_argument('--train-size', default=50000, type=int, help='synthetic train set size')
    p.add_argument('--test-size', default=10000, type=int, help='synthetic test set size')
    p.add_argument('--num-classes', default=10, type=int)
    # --- engine / transport
    p.add_argument('--engine', default='auto', choices=['auto', 'native', 'host'],
                   help='native = sm_100a kernels + NVLink peer memory; host = torch autograd path')
    p.add_argument('--backend', default='auto', choices=['auto', 'p2p', 'nccl', 'gloo'],
                   help='transport; auto = p2p on B200, gloo on CPU')
    p.add_argument('--device', default='auto', choices=['auto', 'cuda', 'cpu'])
    p.add_argument('--no-sync-bn', action='store_true', help='per-rank BatchNorm statistics')
    p.add_argument('--no-graph', action='store_true', help='native engine: do not capture CUDA graphs')
    p.add_argument('--deterministic', action='store_true',
                   help='bitwise-reproducible runs: the torch-autograd engine under torch.use_deterministic_algorithms '
                        '(cuDNN deterministic, fixed cuBLAS workspace), NCCL / gloo all-reduce; the native engine '
                        'accumulates split-K and BatchNorm partial sums with fp32 atomics and is not offered here')
    p.add_argument('--host-dataset', action='store_true',
                   help='native engine: keep the dataset in pinned host memory and copy every batch (default: the '
                        'dataset is uploaded once and lives in GPU memory; a step copies only its indices)')
    p.add_argument('--bucket-mb', default=0.0, type=float,
                   help='gradient bucket cap in MiB (0 = engine default; 25 reproduces torch DDP)')
    p.add_argument('--algo', default='auto', choices=['auto', 'oneshot', 'twoshot', 'nvls', 'nccl'],
                   help='gradient all-reduce algorithm of the native engine')
    # --- run control
    p.add_argument('--seed', default=0, type=int)
    p.add_argument('--init-from', default='', type=str,
                   help='initial weights from a state_dict file (reference key names); epoch numbering starts at 1')
    p.add_argument('--no-augment', action='store_true', help='training batches without random crop / flip')
    p.add_argument('--no-reshuffle', action='store_true',
                   help='same sample order every epoch, as the reference (it never calls set_epoch; SURVEY.md D4)')
    p.add_argument('--resume', default='', type=str, help='model_E.pth or a run directory')
    p.add_argument('--save-every', default=0, type=int, help='checkpoint every N epochs (0 = last only)')
    p.add_argument('--compat-metrics', action='store_true',
                   help="log rank 0's LOCAL numbers and divide test hits by the full set (reference defects D1/D2)")
    p.add_argument('--log-columns', default=4, type=int, choices=[3, 4],
                   help='4 = epoch loss top1 top5; 3 = the exact format ref plot_curves.py reads')
    p.add_argument('--max-steps', default=0, type=int, help='stop each epoch after N steps (smoke runs)')
    p.add_argument('--timeout', default=0.0, type=float, help='watchdog for the whole job, seconds')
    p.add_argument('--no-plot', action='store_true')
    return p


def parse_args(argv: Optional[List[str]] = None) -> argparse.Namespace:
    args = build_parser().parse_args(argv)
    args.milestones = tuple(int(x) for x in str(args.milestones).split(',') if x)
    if args.world_size < 1:
        raise SystemExit('--world_size must be >= 1')
    return args
