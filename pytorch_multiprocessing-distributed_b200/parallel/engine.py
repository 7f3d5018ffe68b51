"""NativeEngine: Python owner of the C++ step engine (csrc/engine.cu, include/pmdist_engine.h).

It lays the model's parameters out in the flat bucket slab (parallel/buckets.py), hands the layout
to the C++ engine, and rebinds every ``nn.Parameter`` / BatchNorm buffer of the PyTorch module to a
*view* of engine memory -- so ``model.state_dict()``, checkpointing and ``load_state_dict`` keep
working with the reference's key names (ref:main.py:75-77) while the training step itself never
touches Python or autograd.  Same interface as HostEngine, so train.py drives either.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional

import torch
import torch.nn as nn

from ..data.loader import Batch, IndexBatch
from ..models.resnet import ResNet
from ..ops import _lib
from ..ops._lib import check
from ..utils.schedule import SGDConfig
from .buckets import params_of, plan_buckets
from .peer import ALGOS, PeerComm

MiB = 1 << 20
MAX_BUCKETS = 24
LL_MAX_ELEMS = 16384           # PMD_LL_MAX_ELEMS (include/pmdist_comm.h)
ALIGN = 64


class PmdEngineConfig(C.Structure):
    _fields_ = [
        ('bottleneck', C.c_int), ('blocks', C.c_int * 4), ('num_classes', C.c_int), ('image_hw', C.c_int),
        ('max_batch', C.c_int), ('n_convs', C.c_int),
        ('conv_w_off', C.POINTER(C.c_longlong)), ('bn_g_off', C.POINTER(C.c_longlong)),
        ('bn_b_off', C.POINTER(C.c_longlong)),
        ('conv_w_bucket', C.POINTER(C.c_int)), ('bn_g_bucket', C.POINTER(C.c_int)), ('bn_b_bucket', C.POINTER(C.c_int)),
        ('fc_w_off', C.c_longlong), ('fc_b_off', C.c_longlong), ('fc_w_bucket', C.c_int), ('fc_b_bucket', C.c_int),
        ('slab_elems', C.c_longlong), ('n_buckets', C.c_int),
        ('bucket_start', C.POINTER(C.c_longlong)), ('bucket_numel', C.POINTER(C.c_longlong)),
        ('d_bn_float', C.c_void_p), ('bn_rm_off', C.POINTER(C.c_longlong)), ('bn_rv_off', C.POINTER(C.c_longlong)),
        ('d_bn_nbt', C.c_void_p),
        ('momentum', C.c_float), ('weight_decay', C.c_float), ('nesterov', C.c_int),
        ('bn_momentum', C.c_float), ('bn_eps', C.c_float),
        ('sync_bn', C.c_int), ('use_graph', C.c_int), ('algo', C.c_int), ('comm_blocks', C.c_int),
        ('bucket_algo', C.POINTER(C.c_int)), ('bucket_blocks', C.POINTER(C.c_int)),
    ]


class PmdEngineBuffers(C.Structure):
    _fields_ = [('param', C.c_ulonglong), ('grad', C.c_ulonglong), ('mom', C.c_ulonglong), ('bf16', C.c_ulonglong),
                ('in_u8', C.c_ulonglong), ('in_aug', C.c_ulonglong), ('in_labels', C.c_ulonglong),
                ('in_valid', C.c_ulonglong), ('in_index', C.c_ulonglong), ('metrics_snap', C.c_ulonglong),
                ('logits', C.c_ulonglong),
                ('metrics', C.c_ulonglong),
                ('slab_elems', C.c_longlong), ('activation_bytes', C.c_longlong), ('heap_bytes_used', C.c_longlong)]


def heap_bytes_for(num_params: int) -> int:
    """fp32 param + grad + momentum, bf16 copy, BN partials, SyncBN push-exchange receive areas, scratch."""
    return int(num_params * 14 * 1.02) + 160 * MiB


def _arr(ctype, values):
    return (ctype * len(values))(*values)


def alias_tensor_f64(addr: int, device) -> torch.Tensor:
    """double[4] view of the engine's metric accumulators."""
    from .peer import alias_tensor
    return alias_tensor(addr, 32, device).view(torch.float64)


class MetricsFuture:
    """Result of :meth:`NativeEngine.read_metrics_async`: ``[loss_sum, top1, top5, count]`` once the copy has landed."""

    def __init__(self, slot: torch.Tensor, event: 'torch.cuda.Event') -> None:
        self._slot, self._event = slot, event

    def done(self) -> bool:
        return self._event.query()

    def result(self) -> List[float]:
        self._event.synchronize()
        return [float(x) for x in self._slot.tolist()]


class NativeEngine:
    kind = 'native'

    def __init__(self, model: ResNet, comm, sgd: SGDConfig, device: torch.device, per_rank_batch: int,
                 image_size: int = 32, sync_bn: bool = True, use_graph: bool = True, bucket_mb: float = 0.0,
                 algo: str = 'auto', comm_blocks: int = 32) -> None:
        if not isinstance(comm, PeerComm):
            # single rank without a transport object, or a torch transport: the engine still needs a heap
            self._own_comm = PeerComm(getattr(comm, 'rank', 0), getattr(comm, 'world_size', 1), device,
                                      heap_bytes_for(sum(p.numel() for p in model.parameters())))
            comm = self._own_comm
        else:
            self._own_comm = None
        self.comm, self.sgd, self.device, self.model = comm, sgd, device, model
        self.lib = _lib.lib()
        self.max_batch = per_rank_batch
        spec = model.spec
        self.spec = spec
        infos = params_of(model)
        comm.verify_same('model structure', repr([(i.name, i.numel) for i in infos]).encode())

        first_cap, cap = (1 * MiB, 4 * MiB) if bucket_mb <= 0 else (int(bucket_mb * MiB),) * 2
        # multi-rank: the last gradients to arrive (the stem) get a bucket of their own, small enough for the barrier-free
        # push all-reduce -- nothing is left to overlap the step's final all-reduce, so its latency is what is exposed
        import os
        tail_ll = comm.world_size > 1 and os.environ.get('PMD_TAIL_LL', '1') != '0' and algo == 'auto'
        while True:
            buckets, total = plan_buckets(infos, first_cap, cap, align_elems=ALIGN,
                                          tail_cap_bytes=LL_MAX_ELEMS * 4 if tail_ll else 0)
            if len(buckets) <= MAX_BUCKETS:
                break
            first_cap, cap = first_cap * 2, cap * 2
        self.buckets, self.slab_elems = buckets, total
        offset_of: Dict[str, int] = {}
        bucket_of: Dict[str, int] = {}
        for bi, b in enumerate(buckets):
            for e in b.entries:
                offset_of[e.name], bucket_of[e.name] = e.offset, bi
        self._offset_of = offset_of
        starts = [b.start for b in buckets]
        numels = [(b.numel + ALIGN - 1) // ALIGN * ALIGN for b in buckets]

        convs = spec.all_convs()
        n = len(convs)
        # BatchNorm buffers: one flat tensor that the module's buffers become views of
        widths = [c.cout for c in convs]
        self.bn_float = torch.zeros(2 * sum(widths), dtype=torch.float32, device=device)
        self.bn_nbt = torch.zeros(n, dtype=torch.int64, device=device)
        rm_off, rv_off, pos = [], [], 0
        mods = dict(model.named_modules())
        for i, c in enumerate(convs):
            bn: nn.BatchNorm2d = mods[c.bn]
            rm_off.append(pos)
            rv_off.append(pos + c.cout)
            self.bn_float[pos:pos + c.cout].copy_(bn.running_mean)
            self.bn_float[pos + c.cout:pos + 2 * c.cout].copy_(bn.running_var)
            self.bn_nbt[i] = int(bn.num_batches_tracked)
            bn._buffers['running_mean'] = self.bn_float[pos:pos + c.cout]
            bn._buffers['running_var'] = self.bn_float[pos + c.cout:pos + 2 * c.cout]
            bn._buffers['num_batches_tracked'] = self.bn_nbt[i]
            pos += 2 * c.cout

        cfg = PmdEngineConfig()
        cfg.bottleneck = 1 if spec.kind == 'bottleneck' else 0
        cfg.blocks = (C.c_int * 4)(*spec.blocks_per_stage)
        cfg.num_classes, cfg.image_hw, cfg.max_batch, cfg.n_convs = spec.num_classes, image_size, per_rank_batch, n
        keep = []      # ctypes arrays must outlive the create call

        def arr(ctype, vals):
            a = _arr(ctype, vals)
            keep.append(a)
            return a

        cfg.conv_w_off = arr(C.c_longlong, [offset_of[f'{c.name}.weight'] for c in convs])
        cfg.bn_g_off = arr(C.c_longlong, [offset_of[f'{c.bn}.weight'] for c in convs])
        cfg.bn_b_off = arr(C.c_longlong, [offset_of[f'{c.bn}.bias'] for c in convs])
        cfg.conv_w_bucket = arr(C.c_int, [bucket_of[f'{c.name}.weight'] for c in convs])
        cfg.bn_g_bucket = arr(C.c_int, [bucket_of[f'{c.bn}.weight'] for c in convs])
        cfg.bn_b_bucket = arr(C.c_int, [bucket_of[f'{c.bn}.bias'] for c in convs])
        cfg.fc_w_off, cfg.fc_b_off = offset_of['linear.weight'], offset_of['linear.bias']
        cfg.fc_w_bucket, cfg.fc_b_bucket = bucket_of['linear.weight'], bucket_of['linear.bias']
        cfg.slab_elems, cfg.n_buckets = total, len(buckets)
        cfg.bucket_start, cfg.bucket_numel = arr(C.c_longlong, starts), arr(C.c_longlong, numels)
        cfg.d_bn_float = self.bn_float.data_ptr()
        cfg.bn_rm_off, cfg.bn_rv_off = arr(C.c_longlong, rm_off), arr(C.c_longlong, rv_off)
        cfg.d_bn_nbt = self.bn_nbt.data_ptr()
        cfg.momentum, cfg.weight_decay, cfg.nesterov = sgd.momentum, sgd.weight_decay, int(sgd.nesterov)
        cfg.bn_momentum, cfg.bn_eps = 0.1, 1e-5
        cfg.sync_bn, cfg.use_graph = int(sync_bn), int(use_graph)
        if algo == 'nccl':
            raise ValueError("--algo nccl belongs to --engine host (the NCCL baseline path)")
        if algo == 'nvls' and not comm.info.multicast_mapped:
            algo = 'twoshot'
        from .tuning import choose_bucket_algos
        self.bucket_algos, self.bucket_blocks, self.algo_source = choose_bucket_algos(
            comm.world_size, [nn_ * 4 for nn_ in numels], bool(comm.info.multicast_mapped), algo, comm_blocks, comm=comm)
        if tail_ll and numels[-1] <= LL_MAX_ELEMS:
            self.bucket_algos[-1] = 'll'
            self.bucket_blocks[-1] = 1
            self.algo_source += '; last bucket (%d B): push all-reduce' % (numels[-1] * 4)
        self.algo = '+'.join(sorted(set(self.bucket_algos)))
        cfg.algo, cfg.comm_blocks = ALGOS[self.bucket_algos[0]], comm_blocks
        cfg.bucket_algo = arr(C.c_int, [ALGOS[a_] for a_ in self.bucket_algos])
        cfg.bucket_blocks = arr(C.c_int, [min(int(b_), 128) for b_ in self.bucket_blocks])
        self._h = C.c_void_p()
        check(self.lib.pmd_engine_create(C.byref(cfg), comm.handle, C.byref(self._h)))
        bufs = PmdEngineBuffers()
        check(self.lib.pmd_engine_buffers(self._h, C.byref(bufs)))
        self.bufs = bufs
        from .peer import alias_tensor
        self.flat_p = alias_tensor(bufs.param, total * 4, device).view(torch.float32)
        self.flat_g = alias_tensor(bufs.grad, total * 4, device).view(torch.float32)
        self.flat_m = alias_tensor(bufs.mom, total * 4, device).view(torch.float32)
        hw = image_size
        self.in_u8 = alias_tensor(bufs.in_u8, per_rank_batch * hw * hw * 3, device).view(per_rank_batch, hw, hw, 3)
        self.in_aug = alias_tensor(bufs.in_aug, per_rank_batch * 12, device).view(torch.int32).view(per_rank_batch, 3)
        self.in_labels = alias_tensor(bufs.in_labels, per_rank_batch * 8, device).view(torch.int64)
        self.in_valid = alias_tensor(bufs.in_valid, per_rank_batch, device)
        both = alias_tensor(bufs.in_index, 2 * (4 + per_rank_batch) * 4, device).view(torch.int32).view(2, 4 + per_rank_batch)
        self.in_index_slots = both                   # two input slots (train_step_pipelined alternates between them)
        self.in_index = both[0]
        self.metrics_snap = alias_tensor(bufs.metrics_snap, 64, device).view(torch.float64).view(2, 4)
        self._pipe = None                            # lazily created state of train_step_pipelined
        self._datasets: Dict[int, tuple] = {}      # id(source) -> (images, labels) resident in device memory
        self._attached: Optional[int] = None
        self.logits = alias_tensor(bufs.logits, per_rank_batch * spec.num_classes * 4, device).view(torch.float32).view(
            per_rank_batch, spec.num_classes)
        # parameters become views of the slab; 4-D conv weights are stored KRSC (channels-last)
        for name, p in model.named_parameters():
            off = offset_of[name]
            flat = self.flat_p[off:off + p.numel()]
            if p.dim() == 4:
                k, c, r, s = p.shape
                view = flat.view(k, r, s, c).permute(0, 3, 1, 2)
            else:
                view = flat.view(p.shape)
            view.copy_(p.data.to(device))
            p.data = view
            p.requires_grad_(False)
        self.step_count = 0
        self._pending_events: List[torch.cuda.Event] = []
        self._metric_slots = None                  # read_metrics_async(): pinned host slots + alias of the device accumulators
        self._metric_dev = None
        self._metric_next = 0
        self.sync_params(broadcast=True)

    # ------------------------------------------------------------------ plumbing
    def sync_params(self, broadcast: bool = False) -> None:
        check(self.lib.pmd_engine_sync_params(self._h, C.c_int(int(broadcast)), _lib.stream_ptr()))

    def _stage(self, batch: Batch, with_valid: bool) -> int:
        if self._attached is not None:
            self.detach_dataset()                  # host-staged batch: back to the staging-buffer input kernel
        n = batch.n
        if n > self.max_batch:
            raise ValueError(f'batch of {n} exceeds the engine capacity {self.max_batch}')
        self.in_u8[:n].copy_(batch.u8, non_blocking=True)
        self.in_aug[:n].copy_(batch.aug, non_blocking=True)
        self.in_labels[:n].copy_(batch.labels, non_blocking=True)
        if with_valid:
            self.in_valid[:n].copy_(batch.valid.to(torch.uint8), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._pending_events.append(ev)
        if len(self._pending_events) > 1:          # the loader has two pinned staging slots
            self._pending_events.pop(0).synchronize()
        return n

    def h2d_bytes(self, n: int, resident: bool = False) -> int:
        """Bytes copied host -> device per step of n samples (device-resident dataset: header + indices only)."""
        if resident:
            return (4 + n) * 4
        hw = self.in_u8.shape[1]
        return n * (hw * hw * 3 + 12 + 8)

    # ------------------------------------------------------------------ device-resident dataset
    def attach_dataset(self, source) -> None:
        """Uploads ``source`` (uint8 [M,H,W,3] + int64 labels) to device memory once (CIFAR-10: 150 MB of 180 GB) and
        makes it the engine's batch source: steps then take an :class:`IndexBatch`.  Several sources (train / test)
        may be uploaded; the one named here is the active one."""
        key = id(source)
        if key not in self._datasets:
            self._datasets[key] = (source.images.to(self.device, non_blocking=True).contiguous(),
                                   source.labels.to(self.device, non_blocking=True).contiguous())
        if self._attached != key:
            img, lab = self._datasets[key]
            check(self.lib.pmd_engine_set_dataset(self._h, C.c_void_p(img.data_ptr()), C.c_void_p(lab.data_ptr()),
                                                  C.c_longlong(img.shape[0])))
            self._attached = key

    def detach_dataset(self) -> None:
        if self._attached is not None:
            check(self.lib.pmd_engine_set_dataset(self._h, C.c_void_p(0), C.c_void_p(0), C.c_longlong(0)))
            self._attached = None

    def _stage_index(self, batch: IndexBatch) -> int:
        if self._attached is None:
            raise RuntimeError('IndexBatch needs a device-resident dataset: call attach_dataset(source) first')
        n = batch.n
        if n > self.max_batch:
            raise ValueError(f'batch of {n} exceeds the engine capacity {self.max_batch}')
        self.in_index[:4 + n].copy_(batch.packed[:4 + n], non_blocking=True)     # the step's only H2D copy
        ev = torch.cuda.Event()
        ev.record()
        self._pending_events.append(ev)
        if len(self._pending_events) > 2:          # the loader rotates 4 pinned slots: never run more than 2 copies ahead
            self._pending_events.pop(0).synchronize()
        return n

    def train_step(self, batch, lr: float) -> None:
        n = self._stage_index(batch) if isinstance(batch, IndexBatch) else self._stage(batch, False)
        check(self.lib.pmd_engine_train_step(self._h, C.c_int(n), C.c_float(lr), _lib.stream_ptr()))
        self.step_count += 1

    def train_step_pipelined(self, batch: IndexBatch, lr: float) -> 'MetricsFuture':
        """One training step whose host<->device traffic is taken off the GPU's critical path (device-resident dataset):

        * the step's index vector is copied host->device on a COPY stream into one of two device slots, as soon as the
          step that last used that slot has finished -- i.e. while the previous step is still running;
        * the step itself (one CUDA-graph launch) goes to the current stream right behind the previous step;
        * its result -- the cumulative ``[loss_sum, top1, top5, count]`` as of the end of THIS step, snapshotted by the
          step's last graph node -- is copied device->host into a pinned slot on a second copy stream while the next
          step runs.  The returned future's ``result()`` waits for just that copy.

        Every step still has its H2D input copy and its D2H result read; consecutive steps nevertheless run back to back
        on the GPU.  Metrics are cumulative here (difference two results for per-step values; ``read_metrics()`` clears)."""
        if self._attached is None:
            raise RuntimeError('train_step_pipelined needs a device-resident dataset: call attach_dataset(source) first')
        n = batch.n
        if n > self.max_batch:
            raise ValueError(f'batch of {n} exceeds the engine capacity {self.max_batch}')
        if self._pipe is None:
            self._pipe = {'h2d': torch.cuda.Stream(self.device), 'd2h': torch.cuda.Stream(self.device), 'step': 0,
                          'slot_free': [None, None],
                          'host': [torch.empty(4, dtype=torch.float64).pin_memory() for _ in range(4)]}
        p = self._pipe
        slot = p['step'] & 1
        cur = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(p['h2d']):
            if p['slot_free'][slot] is not None:
                p['h2d'].wait_event(p['slot_free'][slot])            # the step two back has consumed this slot
            else:
                p['h2d'].wait_stream(cur)                            # first use: behind whatever ran before (set-up, warm-up)
            self.in_index_slots[slot, :4 + n].copy_(batch.packed[:4 + n], non_blocking=True)
            copied = torch.cuda.Event()
            copied.record(p['h2d'])
        cur.wait_event(copied)
        check(self.lib.pmd_engine_train_step_slot(self._h, C.c_int(n), C.c_float(lr), C.c_int(slot), _lib.stream_ptr()))
        done = torch.cuda.Event()
        done.record(cur)
        p['slot_free'][slot] = done
        out = p['host'][p['step'] % len(p['host'])]
        with torch.cuda.stream(p['d2h']):
            p['d2h'].wait_event(done)
            out.copy_(self.metrics_snap[slot], non_blocking=True)
            landed = torch.cuda.Event()
            landed.record(p['d2h'])
        p['step'] += 1
        self.step_count += 1
        return MetricsFuture(out, landed)

    def train_step_staged(self, n: int, lr: float) -> None:
        """Step on whatever is already in the device staging buffers (kernel-only benchmarking)."""
        check(self.lib.pmd_engine_train_step(self._h, C.c_int(n), C.c_float(lr), _lib.stream_ptr()))
        self.step_count += 1

    def eval_step(self, batch) -> None:
        n = self._stage_index(batch) if isinstance(batch, IndexBatch) else self._stage(batch, True)
        check(self.lib.pmd_engine_eval_step(self._h, C.c_int(n), _lib.stream_ptr()))

    def read_metrics(self, reset: bool = True, reduce: bool = False) -> List[float]:
        out = (C.c_double * 4)()
        check(self.lib.pmd_engine_read_metrics(self._h, C.c_int(int(reduce)), C.c_int(int(reset)), out,
                                               _lib.stream_ptr()))
        return [float(x) for x in out]

    def read_metrics_async(self, reset: bool = True) -> 'MetricsFuture':
        """Like :meth:`read_metrics` (this rank only) but without stalling the host: the device->host copy of the
        accumulators into a pinned slot (and the reset) are enqueued behind the steps issued so far, and the
        returned future's ``result()`` waits for just that copy.  A loop that calls ``result()`` one step late keeps
        the next step's launch off the critical path::

            eng.train_step(b0, lr); f = eng.read_metrics_async()
            eng.train_step(b1, lr); g = eng.read_metrics_async(); loss0 = f.result()[0]
        """
        if self._metric_slots is None:
            self._metric_dev = alias_tensor_f64(self.bufs.metrics, self.device)
            self._metric_slots = [torch.empty(4, dtype=torch.float64).pin_memory() for _ in range(4)]
        slot = self._metric_slots[self._metric_next % len(self._metric_slots)]
        self._metric_next += 1
        slot.copy_(self._metric_dev, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        if reset:
            self._metric_dev.zero_()
        return MetricsFuture(slot, ev)

    def set_comm_stub(self, mask: int) -> None:
        """Measurement switch (collective: every rank must make the same call): bit 0 keeps the SyncBatchNorm
        statistics rank-local, bit 1 reduces gradient buckets over this rank alone -- the step then runs the kernels
        of a 1-GPU run.  ``step_ms(mask 0) - step_ms(mask 3)`` is the communication the step could not hide
        ("exposed comm ms/step", BASELINE.json).  Ranks diverge while a stub is active."""
        check(self.lib.pmd_engine_set_comm_stub(self._h, C.c_int(int(mask))))

    def launches_per_step(self, n: Optional[int] = None) -> int:
        return int(self.lib.pmd_engine_launches_per_step(self._h, C.c_int(n or self.max_batch)))

    def synchronize(self) -> None:
        torch.cuda.synchronize(self.device)

    # ------------------------------------------------------------------ state
    def model_state(self) -> Dict[str, torch.Tensor]:
        self.synchronize()
        return {k: v.detach().to('cpu').contiguous().clone() for k, v in self.model.state_dict().items()}

    def load_model_state(self, sd: Dict[str, torch.Tensor]) -> None:
        own = self.model.state_dict()
        for k, v in sd.items():
            own[k].copy_(v.to(self.device))
        self.sync_params(broadcast=False)

    def _momentum_full(self) -> torch.Tensor:
        """Under the owner-computes algorithms rank r only keeps the momentum of slice r of every
        bucket current; collect the live slices from the peers' heaps (collective: all ranks call)."""
        w = self.comm.world_size
        m = self.flat_m.clone()
        if w == 1 or all(a in ('oneshot', 'll') for a in self.bucket_algos):
            return m
        self.comm.barrier()
        mom_off = self.bufs.mom - self.comm.info.local_base
        for b, b_algo in zip(self.buckets, self.bucket_algos):
            if b_algo in ('oneshot', 'll'):
                continue            # every rank updated the whole bucket: momentum is replicated
            n4 = ((b.numel + ALIGN - 1) // ALIGN * ALIGN) // 4
            per = (n4 + w - 1) // w
            for r in range(w):
                lo, hi = b.start + 4 * per * r, b.start + 4 * min(per * (r + 1), n4)
                if hi > lo and r != self.comm.rank:
                    peer = self.comm.peer_view(r, mom_off + lo * 4, (hi - lo,), torch.float32)
                    m[lo:hi].copy_(peer)
        self.comm.barrier()
        return m

    def optimizer_state(self) -> dict:
        m = self._momentum_full()
        out = {}
        for name, p in self.model.named_parameters():
            off = self._offset_of[name]
            flat = m[off:off + p.numel()]
            if p.dim() == 4:
                k, c, r, s = p.shape
                flat = flat.view(k, r, s, c).permute(0, 3, 1, 2)
            out[name] = flat.reshape(p.shape).contiguous().cpu()
        return {'momentum': out, 'step': self.step_count, 'order': [e.name for b in self.buckets for e in b.entries]}

    def load_optimizer_state(self, st: dict) -> None:
        for name, p in self.model.named_parameters():
            off = self._offset_of[name]
            v = st['momentum'][name].to(self.device).view(p.shape)
            flat = self.flat_m[off:off + p.numel()]
            if p.dim() == 4:
                k, c, r, s = p.shape
                flat.view(k, r, s, c).copy_(v.permute(0, 2, 3, 1))
            else:
                flat.copy_(v.reshape(-1))
        self.step_count = int(st.get('step', 0))

    def weights_fingerprint(self) -> float:
        self.synchronize()
        return float(self.flat_p.double().sum())

    def close(self) -> None:
        if self._h:
            torch.cuda.synchronize(self.device)
            self.lib.pmd_engine_destroy(self._h)
            self._h = C.c_void_p()
        if self._own_comm is not None:
            self._own_comm.close()
            self._own_comm = None
