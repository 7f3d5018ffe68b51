"""Environment shim that lets the UNMODIFIED reference (baseline/_ref/main.py) run on this image.

Activated only when REF_SHIM=1, via PYTHONPATH, so it also applies inside the reference's own
mp.spawn children and DataLoader workers (they re-import everything).  It never edits the
reference; each patch exists because of one reference line:

  collections.Iterable alias        <- ref:utils.py:1 (removed from `collections` in Python 3.10)
  matplotlib stub (if not importable) <- ref:plot_curves.py:5, 19-37
  torchvision.datasets.CIFAR10 -> synthetic CIFAR-shaped dataset <- ref:data.py:21-28
      (download=False, no dataset on disk, no network); the reference's own PIL transforms,
      DistributedSampler and DataLoader workers still run on it
  CPU-only mode (REF_CPU=1): .cuda() -> identity, backend -> gloo, no SyncBN, DDP without device_ids

With REF_TIMING_OUT set, it also records per-iteration timing from OUTSIDE the training loop:
a CUDA event pair around DDP.forward ... SGD.step (device time of the step) and the wall-clock
interval between consecutive training forwards (the reference's true end-to-end iteration: its
DataLoader, H2D copy, step and the two .item() syncs).  Results go to $REF_TIMING_OUT.rank<r>.json
when the reference calls dist.destroy_process_group() (ref:main.py:84).
"""
import os
import sys

if os.environ.get('REF_SHIM') == '1':
    import collections
    import collections.abc
    if not hasattr(collections, 'Iterable'):
        collections.Iterable = collections.abc.Iterable

    try:
        import matplotlib  # noqa: F401
    except Exception:
        import types
        mpl = types.ModuleType('matplotlib')
        plt = types.ModuleType('matplotlib.pyplot')

        def _noop(*a, **k):
            return None

        for _n in ('plot', 'xlabel', 'ylabel', 'legend', 'title', 'close', 'figure'):
            setattr(plt, _n, _noop)

        def _savefig(path, *a, **k):
            with open(path, 'wb') as f:
                f.write(b'placeholder: matplotlib is not installed in this image\n')

        plt.savefig = _savefig
        mpl.pyplot = plt
        mpl.use = _noop
        sys.modules['matplotlib'] = mpl
        sys.modules['matplotlib.pyplot'] = plt

    def _install_fake_cifar():
        import torchvision.datasets as tvd
        from ref_fake_cifar import FakeCIFAR10
        tvd.CIFAR10 = FakeCIFAR10

    try:
        _install_fake_cifar()
    except Exception as _e:      # torchvision missing: the reference cannot run anyway
        sys.stderr.write('ref shim: could not install the synthetic CIFAR10: %r\n' % (_e,))

    if os.environ.get('REF_GLOBAL_POOL') == '1':
        # BASELINE.json config 4 (224x224): the reference head is F.avg_pool2d(out, 4) + Linear(512*expansion, 10)
        # (ref:model/resnet.py:102-104) and raises on anything but a 4x4 final map.  The labelled variant used
        # for that config only: pool over the whole remaining map (identical at 32x32).  SURVEY.md 7.5.
        import torch.nn.functional as _F
        _avg = _F.avg_pool2d

        def _avg_pool2d(x, kernel_size, *a, **k):
            if kernel_size == 4 and x.dim() == 4 and x.shape[-1] != 4:
                return _F.adaptive_avg_pool2d(x, 1)
            return _avg(x, kernel_size, *a, **k)

        _F.avg_pool2d = _avg_pool2d

    if os.environ.get('REF_CPU') == '1':
        import torch
        import torch.distributed as dist
        import torch.nn as nn
        from torch.nn.parallel import DistributedDataParallel as DDP
        torch.Tensor.cuda = lambda self, *a, **k: self
        nn.Module.cuda = lambda self, *a, **k: self
        _orig_init = dist.init_process_group

        def _init(backend=None, *a, **k):
            return _orig_init('gloo', *a, **k)

        dist.init_process_group = _init
        nn.SyncBatchNorm.convert_sync_batchnorm = classmethod(lambda cls, m, process_group=None: m)
        _ddp_init = DDP.__init__

        def _ddp(self, module, device_ids=None, **k):
            return _ddp_init(self, module, **k)

        DDP.__init__ = _ddp

    if os.environ.get('REF_TIMING_OUT'):
        import json
        import time

        import torch
        import torch.distributed as dist
        from torch.nn.parallel import DistributedDataParallel as DDP

        _rec = {'start_ev': [], 'end_ev': [], 'wall': []}
        _fwd = DDP.forward

        def _forward(self, *a, **k):
            if self.training and torch.cuda.is_available():
                _rec['wall'].append(time.perf_counter())
                ev = torch.cuda.Event(enable_timing=True)
                ev.record()
                _rec['start_ev'].append(ev)
            return _fwd(self, *a, **k)

        DDP.forward = _forward
        _step = torch.optim.SGD.step

        def _sgd_step(self, *a, **k):
            r = _step(self, *a, **k)
            if torch.cuda.is_available():
                ev = torch.cuda.Event(enable_timing=True)
                ev.record()
                _rec['end_ev'].append(ev)
            return r

        torch.optim.SGD.step = _sgd_step
        _destroy = dist.destroy_process_group

        def _dump_and_destroy(*a, **k):
            try:
                rank = dist.get_rank()
                torch.cuda.synchronize()
                n = min(len(_rec['start_ev']), len(_rec['end_ev']))
                dev_ms = [_rec['start_ev'][i].elapsed_time(_rec['end_ev'][i]) for i in range(n)]
                wall = _rec['wall']
                iter_ms = [(wall[i + 1] - wall[i]) * 1e3 for i in range(len(wall) - 1)]
                flags = {'cudnn_allow_tf32': torch.backends.cudnn.allow_tf32,
                         'matmul_allow_tf32': torch.backends.cuda.matmul.allow_tf32,
                         'cudnn_benchmark': torch.backends.cudnn.benchmark}
                with open('%s.rank%d.json' % (os.environ['REF_TIMING_OUT'], rank), 'w') as f:
                    json.dump({'rank': rank, 'device_ms': dev_ms, 'iter_wall_ms': iter_ms, 'flags': flags}, f)
            finally:
                return _destroy(*a, **k)

        dist.destroy_process_group = _dump_and_destroy
