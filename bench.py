"""Headline benchmark: ResNet training images/sec on N B200s of one node.

    python bench.py --gpus 1 --steps 50 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus 8 --steps 50 --warmup 5
    python bench.py --impl reference --gpus 1 --steps 20 --warmup 5      # the unmodified reference

Metric / config (BASELINE.json): "ResNet images/sec at 1/2/4/8 B200 (device-timed, max over ranks)",
config 3 = weak scaling, the reference's CIFAR ResNet18 ([1,1,1,1], ref:model/resnet.py:108-109),
synthetic 32x32 data, 512 images per GPU, SGD-Nesterov + SyncBN + data-parallel gradient averaging,
i.e. one full training step (forward, loss, backward, gradient all-reduce, optimizer) per "step".

Two numbers are measured on the own arm:
  value  -- device-timed (CUDA events around exactly K steps, barrier + synchronize on both sides,
            max over ranks) with the step's inputs already in device memory;
  e2e    -- the same K steps through the public API (BatchLoader -> engine.train_step -> read_metrics):
            every step copies its uint8 batch + labels + augmentation parameters from pinned host memory
            and reads the loss/accuracy accumulators back to the host.
One training step touches ~1.4 GB of activations per 512 images, far more than the 126 MB L2, so no
explicit L2 flush is needed between iterations (stated in "config").
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BASELINE_PUBLISHED = None      # BASELINE.json: "published": {} -> vs_baseline is null
DEPTH_OF = {'res': 18, 'res18': 18, 'resnet18': 18, 'res34': 34, 'resnet34': 34, 'res50': 50, 'resnet50': 50,
            'res101': 101, 'resnet101': 101, 'res152': 152, 'resnet152': 152}
REF_FACTORY = {18: 'ResNet18', 34: 'ResNet34', 50: 'ResNet50', 101: 'ResNet101', 152: 'ResNet152'}
BLOCKS = {18: '[1,1,1,1] BasicBlock', 34: '[3,4,6,3] BasicBlock', 50: '[3,4,6,3] Bottleneck',
          101: '[3,4,23,3] Bottleneck', 152: '[3,8,36,3] Bottleneck'}


def shared_config(depth: int, gb: int, per_gpu: int, image: int, world: int) -> dict:
    """The workload description both arms print under "config" with identical keys AND values (the driver compares
    them); what differs between the arms (precision, kernels, transport) goes under "impl_detail"."""
    return {'model': f'CIFAR ResNet{depth} {BLOCKS[depth]} (model/resnet.py:{"108-109" if depth == 18 else "112-125"})',
            'global_batch': gb, 'per_gpu_batch': per_gpu, 'seq_len': None, 'image': f'{image}x{image}',
            'parallelism': f'dp{world}', 'sync_bn': True, 'optimizer': 'SGD-Nesterov lr0.1 m0.9 wd1e-4',
            'weights': 'random init', 'head': 'avg_pool2d(4)' if image == 32 else 'global average pool (variant: the '
            'reference head rejects non-32x32 inputs, ref:model/resnet.py:102-104)'}


# ----------------------------------------------------------------------------- clocks sampler
class ClockSampler:
    """Samples SM clock and throttle reasons with NVML while the timed region runs."""
    REASONS = {0x0000000000000004: 'sw_power_cap', 0x0000000000000008: 'hw_slowdown',
               0x0000000000000020: 'sw_thermal_slowdown', 0x0000000000000040: 'hw_thermal_slowdown',
               0x0000000000000080: 'hw_power_brake_slowdown', 0x0000000000000002: 'applications_clocks_setting',
               0x0000000000000010: 'sync_boost', 0x0000000000000100: 'display_clock_setting'}

    def __init__(self, index: int, period: float = 0.005) -> None:
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self) -> None:
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        if self.nv is not None:
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._t is not None:
            self._t.join()

    def summary(self) -> dict:
        s = sorted(self.samples)
        return {'sm_mhz': (s[len(s) // 2] if s else None), 'sm_max_mhz': self.max_mhz, 'reasons': sorted(self.reasons),
                'samples': len(s)}


# ----------------------------------------------------------------------------- own arm
def run_native(a) -> None:
    import torch
    from pmdist.data.loader import BatchLoader, DeviceBatchLoader
    from pmdist.data.sampler import ShardSampler
    from pmdist.data.sources import synthetic_source
    from pmdist.models.resnet import MODEL_ALIASES, build_resnet, resnet_spec
    from pmdist.ops import _lib
    from pmdist.parallel.engine import NativeEngine, heap_bytes_for
    from pmdist.parallel.peer import PeerComm
    from pmdist.utils.schedule import SGDConfig

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', str(rank)))
    if world != a.gpus:
        raise SystemExit(f'--gpus {a.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run --nproc-per-node {a.gpus}')
    torch.cuda.set_device(local)
    device = torch.device('cuda', local)
    depth = MODEL_ALIASES[a.model]
    spec = resnet_spec(depth)
    comm = PeerComm(rank, world, device, heap_bytes_for(spec.num_parameters()))
    torch.manual_seed(0)
    model = build_resnet(depth)
    B = a.per_gpu_batch
    eng = NativeEngine(model, comm, SGDConfig(), device, B, a.image_size, sync_bn=True, use_graph=not a.no_graph,
                       bucket_mb=a.bucket_mb, algo=a.algo)
    src = synthetic_source(B * 8, a.image_size, seed=rank).pin()
    if a.host_dataset:
        loader = BatchLoader(src, ShardSampler(len(src), 1, 0, True, rank), B, True, seed=rank, pin=True)
    else:
        # the framework's default data path: the dataset is uploaded to GPU memory once; per step the host copies
        # the batch's header + index vector from pinned memory, gather + crop/flip (Philox) + normalise run in-kernel
        loader = DeviceBatchLoader(src, ShardSampler(len(src), 1, 0, True, rank), B, True, seed=rank)
        eng.attach_dataset(src)
    lr = 0.1

    def batches():
        ep = 0
        while True:
            loader.set_epoch(ep)
            ep += 1
            for b in loader:
                if b.n == B:
                    yield b

    it = batches()
    # ---- warm-up (first call per batch size is eager, the second captures the CUDA graph)
    for _ in range(max(a.warmup, 3)):
        eng.train_step(next(it), lr)
    eng.read_metrics()
    torch.cuda.synchronize()

    # ---- device-timed region: inputs resident in device memory.  The ranks are aligned ON THE DEVICE (a barrier
    # kernel in the stream right before the start event), so the window holds K steps and no host-side entry skew.
    def timed_steps(k: int) -> float:
        comm.barrier()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if world > 1:
            comm.device_barrier()
        marks = []
        t0.record()
        for _ in range(k):
            eng.train_step_staged(B, lr)
            if trace_steps:                      # diagnostic only (PMD_BENCH_STEP_EVENTS=1): one event per step
                marks.append(torch.cuda.Event(enable_timing=True))
                marks[-1].record()
        t1.record()
        torch.cuda.synchronize()
        comm.barrier()
        if marks:
            prev, per = t0, []
            for m_ in marks:
                per.append(round(prev.elapsed_time(m_), 4))
                prev = m_
            step_traces.append(per)
        return t0.elapsed_time(t1)

    trace_steps = os.environ.get('PMD_BENCH_STEP_EVENTS', '0') != '0'
    step_traces = []
    launches0 = _lib.launch_count()
    with ClockSampler(local) as clk:
        dev_ms = timed_steps(a.steps)
    launches = _lib.launch_count() - launches0
    loss_dev = eng.read_metrics()

    # ---- end-to-end region: public API, H2D of every batch from pinned memory, D2H of the metrics
    comm.barrier()
    # Every step's batch is copied host -> device from pinned memory inside this region and every step's metrics are
    # read back (which synchronises).  The loader's gather of the NEXT batch (into its other pinned slot) is issued
    # before that read, so it overlaps the running step.  (Also tried: H2D one step ahead on a copy stream into a
    # second set of device buffers + device-to-device copy -- the extra stream operations cost more host time than
    # the 63 us PCIe transfer they hide: 1.81 vs 1.69 ms/step.)
    # With PMD_BENCH_ASYNC_METRICS=1 the read of step i's metrics is enqueued right behind step i but consumed on
    # the host after step i+1 has been launched (NativeEngine.read_metrics_async), so the launch leaves the critical path.
    async_metrics = os.environ.get('PMD_BENCH_ASYNC_METRICS', '1') != '0'
    pipelined = async_metrics and not a.host_dataset and os.environ.get('PMD_BENCH_PIPELINED', '1') != '0'
    nxt = next(it)
    fut = None
    if pipelined:
        eng.train_step_pipelined(nxt, lr).result()       # (untimed) creates the copy streams, captures the slot-1 graph
        eng.train_step_pipelined(next(it), lr).result()
        nxt = next(it)
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        if pipelined:
            # H2D of this step's index vector on a copy stream (overlapping the previous step), one graph launch, D2H of
            # this step's metric snapshot on a second copy stream (overlapping the next step)
            f2 = eng.train_step_pipelined(nxt, lr)
            nxt = next(it)
            if fut is not None:
                m = fut.result()                 # the previous step's cumulative loss / accuracy, on the host by now
            fut = f2
            continue
        eng.train_step(nxt, lr)                  # H2D of this step's batch from pinned memory + one graph launch
        if async_metrics:
            f2 = eng.read_metrics_async(reset=False)   # D2H of the (cumulative) metrics, enqueued behind this step
            nxt = next(it)
            if fut is not None:
                m = fut.result()                 # the previous step's loss / accuracy, on the host by now
            fut = f2
        else:
            nxt = next(it)
            m = eng.read_metrics()               # device -> host read of the step's loss / accuracy
    if fut is not None:
        m = fut.result()
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    if async_metrics:
        m = eng.read_metrics()                   # (outside the timed region) totals of the K steps; clears the accumulators
    comm.barrier()

    # ---- exposed communication: the same K steps with the collectives stubbed out (BatchNorm statistics and
    # gradient buckets stay rank-local: the kernels of a 1-GPU run on every GPU at once).  Last, because the
    # ranks' weights diverge under the stub.
    lps = eng.launches_per_step(B)
    stub_ms = {}
    if world > 1 and not a.no_exposed_comm:
        for name, mask in (('no_comm', 3), ('no_bn_exchange', 1), ('no_grad_allreduce', 2)):
            eng.set_comm_stub(mask)
            for _ in range(3):
                eng.train_step_staged(B, lr)       # eager, capture, first replay
            torch.cuda.synchronize()
            stub_ms[name] = timed_steps(a.steps)
        eng.set_comm_stub(0)
    import struct
    names = sorted(stub_ms)
    gathered = comm.all_gather_bytes(struct.pack('%dd' % (2 + len(names)), dev_ms, e2e_ms, *[stub_ms[k] for k in names]))
    cols = list(zip(*[struct.unpack('%dd' % (2 + len(names)), g) for g in gathered]))
    dev_max, e2e_max = max(cols[0]), max(cols[1])
    stub_max = {k: max(cols[2 + i]) for i, k in enumerate(names)}
    if step_traces:
        os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
        with open(os.path.join(ROOT, 'gpurun_out', f'step_ms_n{world}_rank{rank}.json'), 'w') as f:
            json.dump(step_traces, f)
    if rank == 0:
        gb = B * world
        value = gb * a.steps / (dev_max / 1e3)
        out = {
            'metric': 'ResNet training images/sec (device-timed, max over ranks)', 'value': round(value, 1),
            'unit': 'images/s', 'n_gpus': world, 'steps': a.steps, 'warmup': max(a.warmup, 3),
            'ms_per_step': round(dev_max / a.steps, 4), 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': BASELINE_PUBLISHED, 'dtype': 'bf16', 'data': 'synthetic', 'impl': 'pmdist-native',
            'config': shared_config(depth, gb, B, a.image_size, world),
            'impl_detail': {'grad_allreduce': eng.algo, 'grad_allreduce_choice': eng.algo_source, 'transport': comm.describe(),
                            'cuda_graph': not a.no_graph, 'precision': 'bf16 activations / gradients, fp32 master weights',
                            'l2': 'no flush: per-step working set (~2.7 MB/image of activations+gradients) exceeds '
                                  'the 126 MB L2'},
            'clocks': clk.summary(),
            'e2e': {'value': round(gb * a.steps / (e2e_max / 1e3), 1), 'unit': 'images/s',
                    'ms_per_step': round(e2e_max / a.steps, 4),
                    'h2d_bytes_per_step': eng.h2d_bytes(B, resident=not a.host_dataset), 'd2h_bytes_per_step': 32,
                    'path': ('BatchLoader(pinned uint8 batch) -> NativeEngine.train_step -> read_metrics' if a.host_dataset else
                             'DeviceBatchLoader(pinned header + index vector; dataset resident in HBM, in-kernel gather + '
                             'Philox crop/flip) -> NativeEngine.train_step -> read_metrics') +
                            (' (train_step_pipelined: index H2D on a copy stream during the previous step, metric snapshot D2H on '
                             'a second copy stream during the next step, consumed on the host one step later)' if pipelined else
                             (' (every step\'s metrics copied D2H into a pinned slot behind the step, consumed on the host one step '
                              'later: read_metrics_async)' if async_metrics else ''))},
            'exposed_comm_ms': (round((dev_max - stub_max['no_comm']) / a.steps, 4) if stub_max else
                                (0.0 if world == 1 else None)),
            'exposed_comm': ({'ms_per_step_full': round(dev_max / a.steps, 4),
                              **{f'ms_per_step_{k}': round(v / a.steps, 4) for k, v in stub_max.items()},
                              'how': 'same build, same K steps, collectives stubbed by NativeEngine.set_comm_stub '
                                     '(exchanges read local data only); exposed = full - no_comm'} if stub_max else None),
            'gpu_launches': int(launches), 'launches_per_step': lps,
            'loss_after_bench': (m[0] / m[3]) if m[3] else None,
            'achieved_tflops_per_gpu': round(spec.train_flops_per_image(a.image_size) * B * a.steps / (dev_max / 1e3) / 1e12, 2),
        }
        print(json.dumps(out), flush=True)
    eng.close()
    comm.close()


# ----------------------------------------------------------------------------- reference arm
def run_reference(a) -> None:
    """Runs baseline/_ref/main.py (the unmodified reference) through its own CLI; under torchrun only
    rank 0 launches it -- the reference spawns its own world_size processes (ref:main.py:185-188)."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    ref_main = os.path.join(ROOT, 'baseline', '_ref', 'main.py')
    if not os.path.exists(ref_main):
        print(json.dumps({'impl': 'reference', 'unavailable': 'baseline/_ref missing: run baseline/install_reference.sh'}))
        return
    try:
        import torch
        ngpu = torch.cuda.device_count()
    except Exception:
        ngpu = 0
    if ngpu < a.gpus:
        print(json.dumps({'impl': 'reference', 'unavailable': f'needs {a.gpus} CUDA devices, found {ngpu}'}))
        return
    import tempfile
    W, K, N, B = max(a.warmup, 3), a.steps, a.gpus, a.per_gpu_batch
    work = tempfile.mkdtemp(prefix='ref_bench_')
    gb = B * N
    # Scrub everything torchrun exported: in particular TORCHELASTIC_USE_AGENT_STORE=True would make the
    # reference's own init_process_group (ref:main.py:193) look for the agent's TCPStore on ITS port 20080
    # instead of hosting one, and every rank would wait for a store nobody serves.
    torchrun_vars = {'RANK', 'LOCAL_RANK', 'WORLD_SIZE', 'MASTER_ADDR', 'MASTER_PORT', 'GROUP_RANK', 'ROLE_RANK',
                     'LOCAL_WORLD_SIZE', 'ROLE_WORLD_SIZE', 'GROUP_WORLD_SIZE', 'ROLE_NAME', 'OMP_NUM_THREADS',
                     'NCCL_ASYNC_ERROR_HANDLING', 'TORCH_NCCL_ASYNC_ERROR_HANDLING'}
    env = {k: v for k, v in os.environ.items() if k not in torchrun_vars and not k.startswith('TORCHELASTIC_')}
    env.update({'REF_SHIM': '1', 'PYTHONPATH': os.path.join(ROOT, 'baseline', 'shim') + os.pathsep + env.get('PYTHONPATH', ''),
                'REF_TRAIN_SIZE': str(gb * (W + K)), 'REF_TEST_SIZE': str(gb), 'REF_IMAGE_SIZE': str(a.image_size),
                'REF_TIMING_OUT': os.path.join(work, 'timing')})
    if a.image_size != 32:
        env['REF_GLOBAL_POOL'] = '1'      # labelled variant: the reference head cannot take non-32x32 inputs
    depth = DEPTH_OF[a.model]          # (nothing of this framework is imported on the reference arm)
    variant = []
    if depth != 18:
        # the reference CLI can only build ResNet18 (ref:main.py:39-40): labelled variant, same file's other factory
        env['REF_MODEL'] = REF_FACTORY[depth]
        variant.append(f'main.py builds resnet.{REF_FACTORY[depth]} instead of ResNet18 (REF_MODEL shim)')
    if a.image_size != 32:
        variant.append('global average pool head (REF_GLOBAL_POOL shim)')
    bf16 = a.impl == 'reference-bf16'
    if bf16:
        env['REF_AUTOCAST_BF16'] = '1'
        variant.append('model under torch.autocast(bf16) + channels_last (REF_AUTOCAST_BF16 shim)')
    impl_name = 'reference-bf16-autocast' if bf16 else 'reference'
    cmd = [sys.executable, ref_main, '--batch_size', str(gb), '--epochs', '1', '--world_size', str(N),
           '--save_path', os.path.join(work, 'out') + '/', '--print-freq', '1000']
    local = int(os.environ.get('LOCAL_RANK', '0'))
    with ClockSampler(local) as clk:
        r = subprocess.run(cmd, cwd=work, env=env, capture_output=True, text=True, timeout=1500)
    if r.returncode != 0:
        print(json.dumps({'impl': impl_name, 'unavailable': 'reference run failed: ' + (r.stderr or r.stdout)[-300:].replace('\n', ' ')}))
        return
    dev, wall, flags = [], [], {}
    for rk in range(N):
        with open(os.path.join(work, f'timing.rank{rk}.json')) as f:
            t = json.load(f)
        dev.append(sum(t['device_ms'][W:W + K]))
        wall.append(sum(t['iter_wall_ms'][W:W + K - 1]) * K / max(K - 1, 1))
        flags = t['flags']
    dev_max, e2e_max = max(dev), max(wall)
    out = {'metric': 'ResNet training images/sec (device-timed, max over ranks)', 'value': round(gb * K / (dev_max / 1e3), 1),
           'unit': 'images/s', 'n_gpus': N, 'steps': K, 'warmup': W, 'ms_per_step': round(dev_max / K, 4),
           'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
           'dtype': 'bf16 autocast (fp32 weights, optimizer, BatchNorm)' if bf16 else 'fp32 (cuDNN TF32 convs allowed)',
           'data': 'synthetic', 'impl': impl_name,
           'config': shared_config(depth, gb, B, a.image_size, N),
           'impl_detail': {'source': 'baseline/_ref/main.py (byte-identical copy of /root/reference) through its own CLI',
                           'stack': 'torch DDP + SyncBatchNorm over NCCL, cuDNN (benchmark=True), cuBLAS',
                           'variant': variant or 'none: stock reference', 'torch_flags': flags,
                           'timing': 'CUDA events DDP.forward..SGD.step from baseline/shim/sitecustomize.py'},
           'clocks': clk.summary(),
           'e2e': {'value': round(gb * K / (e2e_max / 1e3), 1), 'unit': 'images/s', 'ms_per_step': round(e2e_max / K, 4),
                   'h2d_bytes_per_step': B * (3 * a.image_size * a.image_size * 4 + 8), 'd2h_bytes_per_step': 8,
                   'path': "reference's own loop: DataLoader(4 workers, PIL transforms) -> .cuda() -> step -> 2x .item()"},
           'gpu_launches': None}
    print(json.dumps(out), flush=True)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=30)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='native', choices=['native', 'reference', 'reference-bf16'])
    ap.add_argument('--model', default='res18')
    ap.add_argument('--per-gpu-batch', type=int, default=512)
    ap.add_argument('--image-size', type=int, default=32)
    ap.add_argument('--algo', default='auto')
    ap.add_argument('--bucket-mb', type=float, default=0.0)
    ap.add_argument('--no-graph', action='store_true')
    ap.add_argument('--host-dataset', action='store_true', help='e2e: copy every uint8 batch from pinned host memory')
    ap.add_argument('--no-exposed-comm', action='store_true', help='skip the collectives-stubbed timing runs')
    a = ap.parse_args()
    if a.impl in ('reference', 'reference-bf16'):
        run_reference(a)
    else:
        run_native(a)


if __name__ == '__main__':
    main()
