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

Labelled VARIANTS (never the stock arm; each is named in the JSON line that reports it):
  REF_GLOBAL_POOL=1       head pools the whole final map (224x224 inputs; ref:model/resnet.py:102 hard-codes 4x4)
  REF_MODEL=ResNet50      main.py's `resnet.ResNet18()` (ref:main.py:39-40, the only model it can build) resolves to
                          another factory of the SAME file (ref:model/resnet.py:112-125)
  REF_NO_AUG=1            RandomCrop / RandomHorizontalFlip become identities and REF_DUMP_INIT=<file> saves rank 0's
                          initial state_dict right after DDP(model): only tools/parity_vs_reference.py uses them, to feed
                          both frameworks the same tensors so that their loss curves can be compared number by number
  REF_AUTOCAST_BF16=1     the reference's model runs under torch.autocast(bf16) with channels_last weights and inputs:
                          a like-for-like precision baseline for this framework's bf16 path

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

    if os.environ.get('REF_MODEL'):
        import importlib.abc

        class _RebindResNet18(importlib.abc.MetaPathFinder):
            """After the reference's model/resnet.py has been executed, ResNet18 names the requested factory."""

            def find_spec(self, fullname, path, target=None):
                if fullname != 'model.resnet':
                    return None
                for finder in sys.meta_path:
                    if finder is self or not hasattr(finder, 'find_spec'):
                        continue
                    spec = finder.find_spec(fullname, path, target)
                    if spec is not None and spec.loader is not None:
                        run = spec.loader.exec_module

                        def exec_module(module, _run=run):
                            _run(module)
                            module.ResNet18 = getattr(module, os.environ['REF_MODEL'])

                        spec.loader.exec_module = exec_module
                        return spec
                return None

        sys.meta_path.insert(0, _RebindResNet18())

    if os.environ.get('REF_AUTOCAST_BF16') == '1':
        import torch
        from torch.nn.parallel import DistributedDataParallel as _DDP
        _ddp_init0 = _DDP.__init__

        def _ddp_init_cl(self, module, *a, **k):
            return _ddp_init0(self, module.to(memory_format=torch.channels_last), *a, **k)

        _DDP.__init__ = _ddp_init_cl
        _ddp_fwd0 = _DDP.forward

        def _ddp_fwd_autocast(self, *a, **k):
            a = tuple(t.contiguous(memory_format=torch.channels_last) if torch.is_tensor(t) and t.dim() == 4 else t
                      for t in a)
            with torch.autocast('cuda', dtype=torch.bfloat16):
                out = _ddp_fwd0(self, *a, **k)
            return out.float()

        _DDP.forward = _ddp_fwd_autocast

    if os.environ.get('REF_NO_AUG') == '1':
        import torchvision.transforms as _T

        class _Identity:
            def __init__(self, *a, **k):
                pass

            def __call__(self, img):
                return img

        import torchvision.transforms.transforms as _TT      # the reference names them transforms.transforms.X
        for _m in (_T, _TT):                                 # (ref:data.py:12-13)
            _m.RandomCrop = _Identity
            _m.RandomHorizontalFlip = _Identity

    if os.environ.get('REF_DUMP_INIT'):
        import torch
        from torch.nn.parallel import DistributedDataParallel as _DDP1
        _ddp_init1 = _DDP1.__init__

        def _ddp_init_dump(self, module, *a, **k):
            r = _ddp_init1(self, module, *a, **k)        # after the rank-0 broadcast: what every rank starts from
            import torch.distributed as _dist
            if _dist.get_rank() == 0:
                torch.save({kk: v.detach().cpu().clone() for kk, v in self.state_dict().items()},
                           os.environ['REF_DUMP_INIT'])
            return r

        _DDP1.__init__ = _ddp_init_dump

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
