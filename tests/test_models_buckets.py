"""Unit tier: ResNet shapes/keys vs the reference, bucket plan vs torch DDP, checkpoint format."""
import io

import pytest
import torch

from pmdist.models.resnet import (ResNet18, ResNet34, ResNet50, ResNet101, ResNet152, build_resnet,
                                  model_from_name, resnet_spec)
from pmdist.parallel.buckets import (TORCH_BUCKET_CAP_BYTES, TORCH_FIRST_BUCKET_BYTES, arrival_order,
                                     params_of, plan_buckets)
from pmdist.utils import checkpoint as ckpt

PARAMS = {18: 4_903_242, 34: 21_282_122, 50: 23_520_842, 101: 42_512_970, 152: 58_156_618}


@pytest.mark.parametrize('depth', [18, 34, 50, 101, 152])
def test_param_counts(depth):
    spec = resnet_spec(depth)
    assert spec.num_parameters() == PARAMS[depth]
    if depth in (18, 50):
        m = build_resnet(depth)
        assert sum(p.numel() for p in m.parameters()) == PARAMS[depth]


@pytest.mark.parametrize('depth,ctor', [(18, 'ResNet18'), (34, 'ResNet34'), (50, 'ResNet50')])
def test_state_dict_keys_and_shapes_match_reference(depth, ctor, ref_modules):
    ref = getattr(ref_modules['ref_resnet'], ctor)()
    mine = build_resnet(depth)
    rs, ms = ref.state_dict(), mine.state_dict()
    assert list(rs.keys()) == list(ms.keys())
    assert all(rs[k].shape == ms[k].shape and rs[k].dtype == ms[k].dtype for k in rs)


def test_forward_equals_reference_with_same_weights(ref_modules):
    torch.manual_seed(0)
    ref = ref_modules['ref_resnet'].ResNet18()
    mine = ResNet18()
    mine.load_state_dict(ref.state_dict())
    x = torch.randn(4, 3, 32, 32)
    ref.eval(); mine.eval()
    assert torch.allclose(ref(x), mine(x), atol=1e-5)
    ref.train(); mine.train()
    assert torch.allclose(ref(x), mine(x), atol=1e-5)


def test_224_input_is_legal_here_and_illegal_in_reference(ref_modules):
    x = torch.randn(1, 3, 64, 64)          # any size != 32 breaks the reference head
    with pytest.raises(RuntimeError):
        ref_modules['ref_resnet'].ResNet18().eval()(x)
    assert ResNet18().eval()(x).shape == (1, 10)


def test_flops_model():
    assert resnet_spec(18).train_flops_per_image(32) / 1e9 == pytest.approx(1.517, abs=0.01)
    assert resnet_spec(50).train_flops_per_image(32) / 1e9 == pytest.approx(7.783, abs=0.03)


def test_model_names():
    assert sum(p.numel() for p in model_from_name('res').parameters()) == PARAMS[18]
    with pytest.raises(ValueError):
        model_from_name('vgg')


def test_arrival_order_matches_torch_reducer_observation():
    order = [p.index for p in arrival_order(params_of(ResNet18()))]
    assert order[:8] == [37, 36, 34, 35, 33, 31, 32, 30]                 # SURVEY probe:1
    assert order[8:14] == [28, 29, 27, 25, 26, 24]
    assert sorted(order) == list(range(38))


def test_parity_bucket_sizes_match_torch_ddp():
    b18, total = plan_buckets(params_of(ResNet18()), TORCH_FIRST_BUCKET_BYTES, TORCH_BUCKET_CAP_BYTES)
    assert [b.nbytes() for b in b18] == [9_990_184, 9_622_784]
    assert total == PARAMS[18]
    b50, _ = plan_buckets(params_of(ResNet50()), TORCH_FIRST_BUCKET_BYTES, TORCH_BUCKET_CAP_BYTES)
    assert [b.nbytes() for b in b50] == [4_292_648, 27_291_648, 26_255_360, 26_550_272, 9_693_440]


def test_buckets_are_contiguous_aligned_ranges():
    buckets, total = plan_buckets(params_of(ResNet18()), 1 << 20, 4 << 20, align_elems=64)
    pos = 0
    for b in buckets:
        assert b.start >= pos and b.start % 64 == 0
        for e in b.entries:
            assert e.offset % 64 == 0 and b.start <= e.offset and e.offset + e.numel <= b.start + b.numel
        pos = b.start + b.numel
    assert total % 64 == 0 and total >= pos


def test_checkpoint_is_reference_format(tmp_path):
    m = ResNet18()
    path = ckpt.save_model(str(tmp_path), 2, m.state_dict())
    assert path.endswith('model_2.pth')
    sd = torch.load(path, weights_only=True)
    assert len(sd) == 74 and all(k.startswith('module.') for k in sd)                  # SURVEY probe:4
    assert list(sd)[0] == 'module.conv1.weight' and list(sd)[-1] == 'module.linear.bias'
    assert {v.dtype for v in sd.values()} == {torch.float32, torch.int64}
    fresh = ResNet18()
    fresh.load_state_dict(ckpt.load_model(path))
    assert all(torch.equal(a, b) for a, b in zip(m.state_dict().values(), fresh.state_dict().values()))


def test_reference_written_checkpoint_loads(tmp_path, ref_modules):
    ref = torch.nn.Sequential()
    ref.add_module('module', ref_modules['ref_resnet'].ResNet18())       # same keys as the DDP wrapper
    torch.save(ref.state_dict(), tmp_path / 'model_20.pth')
    mpth, spth, e = ckpt.find_resume(str(tmp_path))
    assert e == 20 and spth is None
    ResNet18().load_state_dict(ckpt.load_model(mpth))


def test_bucket_algorithm_selector_uses_measured_table(tmp_path, monkeypatch):
    import json
    from pmdist.parallel.tuning import AllreduceTable, choose_bucket_algos
    rows = [{'bytes': 4096, 'oneshot_b8_us': 5.0, 'twoshot_b8_us': 7.0, 'nvls_b8_us': 6.0, 'nccl_us': 20.0},
            {'bytes': 16 << 20, 'oneshot_b96_us': 90.0, 'twoshot_b96_us': 50.0, 'nvls_b96_us': 40.0, 'nccl_us': 55.0}]
    path = tmp_path / 'allreduce_w8.json'
    path.write_text(json.dumps({'world': 8, 'rows': rows}))
    monkeypatch.setenv('PMDIST_ALLREDUCE_TABLE', str(path))
    algos, blocks, why = choose_bucket_algos(8, [2048, 10 << 20], multicast=True)
    assert algos == ['oneshot', 'nvls'] and blocks == [8, 96] and 'measured' in why
    algos, _, _ = choose_bucket_algos(8, [2048, 10 << 20], multicast=False)        # no NVLS on this host
    assert algos == ['oneshot', 'twoshot']
    assert choose_bucket_algos(8, [1 << 20], True, forced='twoshot')[0] == ['twoshot']
    assert choose_bucket_algos(1, [1 << 20], True)[0] == ['oneshot']
    monkeypatch.delenv('PMDIST_ALLREDUCE_TABLE')
    assert AllreduceTable.load(7) is None
    assert choose_bucket_algos(7, [1024, 64 << 20], True)[0] == ['oneshot', 'nvls']     # heuristic fallback


def test_tail_bucket_split_keeps_layout_and_isolates_the_last_gradients():
    """tail_cap_bytes splits the last-arriving gradients (the stem) into a final small bucket for the push all-reduce:
    same slab layout (offsets, total), contiguous buckets, tail within the cap; a last bucket that already fits is left."""
    from pmdist.models.resnet import build_resnet
    from pmdist.parallel.buckets import params_of, plan_buckets
    MiB = 1 << 20
    for depth, expect_split in ((18, True), (34, True), (50, False)):
        infos = params_of(build_resnet(depth))
        base, total0 = plan_buckets(infos, MiB, 4 * MiB, align_elems=64)
        tail, total1 = plan_buckets(infos, MiB, 4 * MiB, align_elems=64, tail_cap_bytes=64 << 10)
        assert total0 == total1
        assert [(e.name, e.offset) for b in base for e in b.entries] == [(e.name, e.offset) for b in tail for e in b.entries]
        assert (len(tail) == len(base) + 1) == expect_split
        pos = tail[0].start
        for b in tail:
            assert b.start == pos
            pos = b.start + b.numel
        assert tail[-1].numel * 4 <= 64 << 10
        assert tail[-1].entries[-1].name == 'conv1.weight'          # the stem's weight gradient arrives last
