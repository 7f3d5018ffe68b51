"""Unit tier: AverageMeter / Logger / accuracy / LR table / plot."""
import os

import pytest
import torch

from pmdist.utils.meters import AverageMeter, Logger, accuracy, accuracy_topk
from pmdist.utils.schedule import SGDConfig, lr_for_epoch
from pmdist.utils import plot_curves


def test_average_meter_is_sample_weighted(ref_modules):
    a, b = AverageMeter(), ref_modules['ref_utils'].AverageMeter()
    for v, n in [(1.0, 4), (3.0, 12), (0.5, 1)]:
        a.update(v, n); b.update(v, n)
    assert (a.val, a.sum, a.count, a.avg) == (b.val, b.sum, b.count, b.avg)
    assert a.avg == pytest.approx((4 + 36 + 0.5) / 17)


def test_logger_format_is_byte_exact(tmp_path, ref_modules):
    mine, ref = Logger(str(tmp_path / 'a.log')), ref_modules['ref_utils'].Logger(str(tmp_path / 'b.log'))
    rows = [[1, 2.3025851, 10.0], [2, 0.5, 99.123456789], [12345, 1e-7, 0.0]]
    for r in rows:
        mine.write(r); ref.write(r)
    assert open(tmp_path / 'a.log').read() == open(tmp_path / 'b.log').read()
    assert open(tmp_path / 'a.log').readline() == '0001 2.302585 10.000000\n'
    assert mine.read() == ref.read()
    assert len(mine) == 3


def test_logger_type_rules(tmp_path):
    lg = Logger(str(tmp_path / 'x.log'))
    lg.write([1, 2.0, 'abc'])
    with pytest.raises(AssertionError):
        lg.write([1, 2.0])
    strict = Logger(str(tmp_path / 'y.log'), strict=True)
    with pytest.raises(Exception, match='Not supported type'):
        strict.write([1, torch.tensor(2.0)])
    lenient = Logger(str(tmp_path / 'z.log'))
    lenient.write([3, torch.tensor(2.5), 7.0, 1.0])       # 4 columns, tensor accepted
    assert open(tmp_path / 'z.log').read() == '0003 2.500000 7.000000 1.000000\n'
    assert len(Logger(str(tmp_path / 'missing.log'))) == 0


def test_accuracy_matches_reference_top1(ref_modules):
    torch.manual_seed(0)
    out, tgt = torch.randn(64, 10), torch.randint(0, 10, (64,))
    p, c = accuracy(out, tgt)
    rp, rc = ref_modules['ref_utils'].accuracy(out, tgt)
    assert torch.equal(p, rp) and torch.equal(c, rc)


def test_accuracy_top5_works():
    torch.manual_seed(1)
    out, tgt = torch.randn(128, 10), torch.randint(0, 10, (128,))
    (p1, p5), correct = accuracy_topk(out, tgt, (1, 5))
    ranks = (out.argsort(1, descending=True) == tgt.view(-1, 1)).float().argmax(1)
    assert p1.item() == pytest.approx(100.0 * (ranks < 1).float().mean().item())
    assert p5.item() == pytest.approx(100.0 * (ranks < 5).float().mean().item())
    assert correct.shape == (5, 128)


def test_lr_table_matches_multisteplr_stepped_before_training():
    """Reference: scheduler.step() at the top of each epoch (ref:main.py:69-70) -> epochs 1-59 at
    0.1, 60-79 at 0.01, 80+ at 0.001 (SURVEY.md probe:6)."""
    import warnings
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([p], lr=0.1)
    sch = torch.optim.lr_scheduler.MultiStepLR(opt, milestones=[60, 80], gamma=0.1)
    cfg = SGDConfig()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        for epoch in range(1, 101):
            sch.step()
            assert cfg.lr_at(epoch) == pytest.approx(opt.param_groups[0]['lr'], rel=1e-12), epoch
    assert lr_for_epoch(59) == pytest.approx(0.1) and lr_for_epoch(60) == pytest.approx(0.01)
    assert lr_for_epoch(80) == pytest.approx(0.001)


@pytest.mark.parametrize('cols', [3, 4])
def test_plot_reads_3_and_4_column_logs(tmp_path, cols):
    tr, te = Logger(str(tmp_path / 'train.log')), Logger(str(tmp_path / 'test.log'))
    for e in range(1, 6):
        tr.write([e, 2.0 / e, 10.0 * e, 50.0 + e][:cols])
        te.write([e, 2.5 / e, 9.0 * e, 45.0 + e][:cols])
    plot_curves.draw_plot(str(tmp_path))
    for name in ('test_accuracy.png', 'loss.png'):
        data = open(tmp_path / name, 'rb').read()
        assert data[:8] == b'\x89PNG\r\n\x1a\n' and len(data) > 500


def test_reference_plotter_reads_our_3_column_logs(tmp_path, ref_modules):
    """--log-columns 3 must produce files ref plot_curves.py can unpack (ref:plot_curves.py:15-16)."""
    tr = Logger(str(tmp_path / 'train.log'))
    tr.write([1, 2.0, 10.0]); tr.write([2, 1.0, 20.0])
    rows = ref_modules['ref_utils'].Logger(str(tmp_path / 'train.log')).read()
    epoch, loss, acc = zip(*rows)
    assert epoch == (1.0, 2.0) and acc == (10.0, 20.0)
