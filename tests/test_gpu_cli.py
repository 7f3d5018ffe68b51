"""End-to-end tier on a B200: the reference's command line (python main.py ...) through the NATIVE engine --
spawned ranks, on-GPU augmentation, CUDA-graph steps, tail batches, evaluation with the duplicate mask,
reference-format logs + checkpoint, and --resume."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, timeout=600):
    return subprocess.run([sys.executable, os.path.join(ROOT, 'main.py')] + args, cwd=ROOT, capture_output=True,
                          text=True, timeout=timeout)


def _check_run_dir(out, epochs, stdout):
    names = set(os.listdir(out))
    assert {'train.log', 'test.log', f'model_{epochs}.pth', f'state_{epochs}.pth', 'main.py', 'loss.png',
            'test_accuracy.png'} <= names, names
    for f in ('train.log', 'test.log'):
        lines = open(os.path.join(out, f)).read().splitlines()
        assert [ln.split(' ')[0] for ln in lines] == ['%04d' % e for e in range(1, epochs + 1)]
        assert all(len(ln.split(' ')) == 4 for ln in lines)
    sd = torch.load(os.path.join(out, f'model_{epochs}.pth'), weights_only=True)
    assert len(sd) == 74 and list(sd)[0] == 'module.conv1.weight' and sd['module.conv1.weight'].shape == (64, 3, 3, 3)
    assert all(torch.isfinite(v.float()).all() for v in sd.values())
    assert 'Epoch: [1][0/' in stdout and 'Accuracy ' in stdout


@pytest.mark.gpu
def test_cli_native_single_gpu_train_eval_checkpoint_resume(tmp_path):
    out = str(tmp_path / 'run') + '/'
    base = ['--world_size', '1', '--batch_size', '64', '--data', 'synthetic', '--train-size', '200', '--test-size', '90',
            '--save_path', out, '--print-freq', '2', '--engine', 'native']
    r = _run(base + ['--epochs', '2'])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    _check_run_dir(out, 2, r.stdout)
    # 200 samples / 64 -> batches of 64,64,64,8 (tail batch), 90 test samples -> 64 + 26
    acc = [float(ln.split(' ')[2]) for ln in open(os.path.join(out, 'test.log')).read().splitlines()]
    assert all(abs(a * 90 / 100 - round(a * 90 / 100)) < 1e-3 for a in acc)          # hits / 90 exactly
    tr = [float(ln.split(' ')[1]) for ln in open(os.path.join(out, 'train.log')).read().splitlines()]
    assert all(t == t and t < 1e4 for t in tr)              # finite (lr 0.1 on 200 random images need not descend)
    r2 = _run(base + ['--epochs', '3', '--resume', out])
    assert r2.returncode == 0, r2.stdout[-2000:] + r2.stderr[-2000:]
    assert len(open(os.path.join(out, 'train.log')).read().splitlines()) == 3
    assert os.path.exists(os.path.join(out, 'model_3.pth'))


@pytest.mark.gpu
@pytest.mark.multigpu
def test_cli_native_two_gpus_p2p(tmp_path):
    out = str(tmp_path / 'run2') + '/'
    r = _run(['--world_size', '2', '--batch_size', '64', '--epochs', '2', '--data', 'synthetic', '--train-size', '256',
              '--test-size', '100', '--save_path', out, '--print-freq', '2'])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    _check_run_dir(out, 2, r.stdout)
    acc = [float(ln.split(' ')[2]) for ln in open(os.path.join(out, 'test.log')).read().splitlines()]
    assert all(abs(a - round(a)) < 1e-3 for a in acc)                                # global hits / 100 samples


@pytest.mark.gpu
@pytest.mark.multigpu
def test_cli_host_engine_nccl_baseline_two_gpus(tmp_path):
    """The NCCL + cuDNN baseline/oracle path (--engine host --backend nccl) keeps working."""
    out = str(tmp_path / 'run3') + '/'
    r = _run(['--world_size', '2', '--batch_size', '32', '--epochs', '1', '--data', 'synthetic', '--train-size', '64',
              '--test-size', '32', '--save_path', out, '--print-freq', '1', '--engine', 'host', '--backend', 'nccl'])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert os.path.exists(os.path.join(out, 'model_1.pth'))


@pytest.mark.gpu
def test_cli_deterministic_flag_gives_bit_identical_checkpoints_on_gpu(tmp_path):
    """--deterministic selects the torch-autograd engine under deterministic algorithms: two runs, identical bytes."""
    sds = []
    for i in range(2):
        out = str(tmp_path / f'det{i}') + '/'
        r = _run(['--world_size', '1', '--epochs', '1', '--batch_size', '32', '--data', 'synthetic', '--train-size', '96',
                  '--test-size', '32', '--save_path', out, '--deterministic'])
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        sds.append(torch.load(out + 'model_1.pth', weights_only=True))
    assert all(torch.equal(sds[0][k], sds[1][k]) for k in sds[0])
