"""Device-resident dataset plumbing that runs without a GPU: the host reproduction of the in-kernel Philox draws
(known-answer vectors from the Random123 distribution), the distribution of crop offsets / flips
(ref:data.py:12-13: RandomCrop(32, padding=8) -> 17 offsets per axis, RandomHorizontalFlip p = 0.5), and the
index-only batches (encoding of sampler wrap-around duplicates, host-form oracle batch)."""
import numpy as np
import torch

from pmdist.data.loader import DeviceBatchLoader, augment_cpu
from pmdist.data.philox import AUG_STREAM, aug_params, philox4x32_10
from pmdist.data.sampler import ShardSampler
from pmdist.data.sources import synthetic_source


def test_philox_known_answers():
    z = philox4x32_10(np.zeros((1, 4), dtype=np.uint32), (0, 0))[0]
    assert [int(x) for x in z] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    f = philox4x32_10(np.full((1, 4), 0xffffffff, dtype=np.uint32), (0xffffffff, 0xffffffff))[0]
    assert [int(x) for x in f] == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    p = philox4x32_10(np.array([[0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344]], dtype=np.uint32),
                      (0xa4093822, 0x299f31d0))[0]
    assert [int(x) for x in p] == [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_aug_params_distribution_and_determinism():
    n = 200000
    a = aug_params(np.arange(n), seed=7, epoch=3, train=True)
    assert a.dtype == torch.int32 and tuple(a.shape) == (n, 3)
    assert int(a[:, :2].min()) == 0 and int(a[:, :2].max()) == 16 and set(a[:, 2].tolist()) == {0, 1}
    for col in (0, 1):
        counts = torch.bincount(a[:, col].long(), minlength=17).double()
        chi2 = float(((counts - n / 17) ** 2 / (n / 17)).sum())
        assert chi2 < 45.0, chi2                    # 16 dof: P(chi2 > 45) ~ 1e-4
    assert abs(float(a[:, 2].double().mean()) - 0.5) < 0.005
    assert abs(float(np.corrcoef(a[:, 0].numpy(), a[:, 1].numpy())[0, 1])) < 0.01
    assert torch.equal(a, aug_params(np.arange(n), 7, 3, True))              # pure function of (seed, epoch, index)
    assert not torch.equal(a, aug_params(np.arange(n), 7, 4, True))          # new epoch, new draws
    assert not torch.equal(a, aug_params(np.arange(n), 8, 3, True))
    # a sample's draw does not depend on which other samples share its batch (or which rank owns it)
    assert torch.equal(aug_params([5, 99999, 17], 7, 3, True), a[[5, 99999, 17]])
    e = aug_params(np.arange(10), 7, 3, False)
    assert torch.equal(e, torch.tensor([[8, 8, 0]] * 10, dtype=torch.int32))  # evaluation: centre crop, no flip


def test_device_batch_loader_encodes_duplicates_and_matches_sampler():
    src = synthetic_source(200, 32, seed=1)
    w = 3
    seen = []
    for rank in range(w):
        smp = ShardSampler(len(src), w, rank, False, 0)
        dl = DeviceBatchLoader(src, smp, 32, False, seed=11)
        dl.set_epoch(4)
        idx, valid = smp.indices_and_mask()
        got_idx, got_valid = [], []
        for b in dl:
            assert b.packed.dtype == torch.int32
            assert b.packed[:4].tolist() == [11, AUG_STREAM, 4, 0]
            got_idx.append(b.indices.clone())
            got_valid.append(b.valid.clone())
        assert torch.equal(torch.cat(got_idx), idx) and torch.equal(torch.cat(got_valid), valid)
        seen.append(torch.cat(got_idx)[torch.cat(got_valid)])
    assert sorted(torch.cat(seen).tolist()) == list(range(200))              # every sample counted exactly once


def test_index_batch_host_form_is_the_same_samples():
    src = synthetic_source(64, 32, seed=2)
    dl = DeviceBatchLoader(src, ShardSampler(64, 1, 0, True, 5), 16, True, seed=5)
    dl.set_epoch(1)
    b = next(iter(dl))
    hb = b.to_host_batch(src)
    assert torch.equal(hb.u8, src.images[b.indices]) and torch.equal(hb.labels, src.labels[b.indices])
    assert torch.equal(hb.aug, aug_params(b.indices.numpy(), 5, 1, True))
    x = augment_cpu(hb.u8, hb.aug)
    assert tuple(x.shape) == (16, 3, 32, 32) and float(x.min()) >= -1.0 and float(x.max()) <= 1.0
