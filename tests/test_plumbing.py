"""Plumbing tier (BASELINE.json config 1): 2 processes, gloo, synthetic 32x32 -- identical weights
on every rank, rank-0-only logs in the reference format, reference-format checkpoint, resume, and
failure propagation from a dying rank."""
import os
import subprocess
import sys
import time

import pytest
import torch

from pmdist.parallel import launch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, timeout=900):
    env = dict(os.environ, OMP_NUM_THREADS='2', PMDIST_BIND_CPUS='0')
    return subprocess.run([sys.executable, os.path.join(ROOT, 'main.py')] + args, cwd=ROOT, env=env,
                          capture_output=True, text=True, timeout=timeout)


@pytest.fixture(scope='module')
def two_rank_run(tmp_path_factory):
    out = tmp_path_factory.mktemp('run')
    r = _run(['--world_size', '2', '--epochs', '2', '--batch_size', '16', '--data', 'synthetic',
              '--train-size', '48', '--test-size', '20', '--save_path', str(out) + '/', '--print-freq', '1',
              '--device', 'cpu'])
    assert r.returncode == 0, r.stdout + r.stderr
    return out, r.stdout


def test_files_and_log_format(two_rank_run):
    out, stdout = two_rank_run
    names = set(os.listdir(out))
    assert {'train.log', 'test.log', 'model_2.pth', 'state_2.pth', 'main.py', 'loss.png', 'test_accuracy.png'} <= names
    for f in ('train.log', 'test.log'):
        lines = open(out / f).read().splitlines()
        assert len(lines) == 2
        for i, ln in enumerate(lines, 1):
            cols = ln.split(' ')
            assert cols[0] == '%04d' % i and len(cols) == 4
            assert all(len(c.split('.')[1]) == 6 for c in cols[1:])
    # console formats of ref:main.py:119-127 / 162-170, printed by rank 0 only
    assert stdout.count('-------------------Make loader-------------------') == 1
    assert 'Epoch: [1][0/3]\tTime ' in stdout and '\tData ' in stdout and '\tPrec ' in stdout
    assert 'test : [0/2]\tTime ' in stdout and stdout.count('Accuracy ') == 2


def test_eval_counts_every_test_sample_once(two_rank_run):
    """D1/D5: global accuracy = hits / 20 samples, so it is a multiple of 5%."""
    out, _ = two_rank_run
    for ln in open(out / 'test.log').read().splitlines():
        acc = float(ln.split(' ')[2])
        assert abs(acc / 5.0 - round(acc / 5.0)) < 1e-4


def test_checkpoint_matches_reference_key_set(two_rank_run):
    out, _ = two_rank_run
    sd = torch.load(out / 'model_2.pth', weights_only=True)
    assert len(sd) == 74 and list(sd)[0] == 'module.conv1.weight'
    assert int(sd['module.bn1.num_batches_tracked']) == 6          # 3 steps x 2 epochs


def _equal_weights_worker(rank, world, tmp):
    import torch
    from pmdist.cli import parse_args
    from pmdist.data.loader import make_loaders
    from pmdist.data.sources import synthetic_source
    from pmdist.models.resnet import ResNet18
    from pmdist.parallel.comm import make_comm
    from pmdist.parallel.host_engine import HostEngine
    from pmdist.utils.schedule import SGDConfig
    torch.set_num_threads(2)
    torch.manual_seed(100 + rank)                      # different init per rank on purpose (D12)
    comm = make_comm(rank, world, 'gloo')
    eng = HostEngine(ResNet18(), comm, SGDConfig(), torch.device('cpu'))
    src = synthetic_source(32, 32, seed=0)
    tr, _ = make_loaders(src, src, 16, world, rank)
    fps = [eng.weights_fingerprint()]
    for b in list(tr)[:2]:
        eng.train_step(b, 0.1)
        fps.append(eng.weights_fingerprint())
    bn = eng.model.state_dict()['bn1.running_mean'].double().sum().item()
    torch.save({'fps': fps, 'bn': bn, 'order': list(eng.bucket_launch_order),
                'sizes': [b.nbytes() for b in eng.buckets]}, os.path.join(tmp, f'r{rank}.pt'))
    comm.close()


def test_ranks_hold_identical_weights_after_init_and_steps(tmp_path):
    launch.spawn(_equal_weights_worker, 2, args=(str(tmp_path),), timeout=600, env={'PMDIST_BIND_CPUS': '0'})
    a, b = torch.load(tmp_path / 'r0.pt'), torch.load(tmp_path / 'r1.pt')
    assert a['fps'] == b['fps']                        # bitwise-equal slabs after broadcast and after each step
    assert len(set(a['fps'])) == 3                     # and they actually moved
    assert a['bn'] == b['bn']                          # SyncBN: identical running stats without a buffer broadcast
    assert a['sizes'] == [9_990_184, 9_622_784] and a['order'] == [0, 1]


def test_resume_continues_epoch_numbering(two_rank_run, tmp_path):
    out, _ = two_rank_run
    r = _run(['--world_size', '1', '--epochs', '3', '--batch_size', '16', '--data', 'synthetic',
              '--train-size', '32', '--test-size', '16', '--save_path', str(out) + '/', '--device', 'cpu',
              '--resume', str(out), '--max-steps', '1'])
    assert r.returncode == 0, r.stdout + r.stderr
    lines = open(out / 'train.log').read().splitlines()
    assert [ln.split(' ')[0] for ln in lines] == ['0001', '0002', '0003']
    assert os.path.exists(out / 'model_3.pth')


def _dying_worker(rank, world):
    if rank == 1:
        time.sleep(0.5)
        raise ValueError('boom on rank 1')
    time.sleep(120)                                    # a healthy rank that would otherwise hang


def test_first_failure_is_reported_and_siblings_are_killed():
    t0 = time.time()
    with pytest.raises(launch.RankFailure) as ei:
        launch.spawn(_dying_worker, 2, timeout=60)
    assert ei.value.rank == 1 and 'boom on rank 1' in ei.value.child_traceback
    assert time.time() - t0 < 60


def _hanging_worker(rank, world):
    time.sleep(120)


def test_watchdog_timeout():
    with pytest.raises(TimeoutError):
        launch.spawn(_hanging_worker, 2, timeout=3)


def test_deterministic_flag_gives_bit_identical_checkpoints(tmp_path):
    """--deterministic: the same command twice -> the same bytes in every tensor of the checkpoint."""
    sds = []
    for i in range(2):
        out = str(tmp_path / f'run{i}') + '/'
        r = _run(['--world_size', '2', '--epochs', '1', '--batch_size', '16', '--data', 'synthetic', '--train-size', '32',
                  '--test-size', '16', '--save_path', out, '--device', 'cpu', '--deterministic'])
        assert r.returncode == 0, r.stdout + r.stderr
        sds.append(torch.load(out + 'model_1.pth', weights_only=True))
    assert sds[0].keys() == sds[1].keys()
    assert all(torch.equal(sds[0][k], sds[1][k]) for k in sds[0])
