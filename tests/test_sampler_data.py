"""Unit tier: sampler index parity, batch split, augmentation math."""
import pytest
import torch
from torch.utils.data.distributed import DistributedSampler

from pmdist.data.loader import BatchLoader, augment_cpu, draw_aug_params, make_loaders
from pmdist.data.sampler import ShardSampler, per_rank_batch
from pmdist.data.sources import synthetic_source


class _DS(torch.utils.data.Dataset):
    def __init__(self, n): self.n = n
    def __len__(self): return self.n
    def __getitem__(self, i): return i


@pytest.mark.parametrize('n,world', [(10, 2), (11, 2), (50, 3), (7, 8), (1000, 8), (3, 4)])
@pytest.mark.parametrize('shuffle', [True, False])
def test_indices_match_torch_distributed_sampler(n, world, shuffle):
    for rank in range(world):
        for epoch in (0, 3):
            ref = DistributedSampler(_DS(n), num_replicas=world, rank=rank, shuffle=shuffle, seed=5)
            ref.set_epoch(epoch)
            mine = ShardSampler(n, world, rank, shuffle, seed=5)
            mine.set_epoch(epoch)
            assert list(ref) == mine.indices().tolist()
            assert len(ref) == len(mine)


def test_valid_mask_marks_wraparound_duplicates():
    n, world = 11, 4
    seen, valid_total = [], 0
    for r in range(world):
        idx, valid = ShardSampler(n, world, r, shuffle=False).indices_and_mask()
        seen += idx[valid].tolist()
        valid_total += int(valid.sum())
    assert valid_total == n and sorted(seen) == list(range(n))


def test_per_rank_batch_is_reference_formula():
    assert per_rank_batch(64, 3) == 21 and per_rank_batch(128, 8) == 16 and per_rank_batch(128, 2) == 64


def test_ranks_see_disjoint_data_and_cover_global_batch():
    src = synthetic_source(64, 8, seed=1)
    src.labels = torch.arange(64)                      # label == sample id
    got = []
    for r in range(4):
        tr, _ = make_loaders(src, src, 32, 4, r, seed=0)
        tr.set_epoch(1)
        b = next(iter(tr))
        assert b.n == 8
        got += b.labels.tolist()
    assert len(set(got)) == 32                         # one global batch = 32 distinct samples


def test_eval_params_are_identity_and_train_params_are_seeded():
    p = draw_aug_params(5, False, None)
    assert p.tolist() == [[8, 8, 0]] * 5
    g1, g2 = torch.Generator().manual_seed(3), torch.Generator().manual_seed(3)
    assert torch.equal(draw_aug_params(16, True, g1), draw_aug_params(16, True, g2))


def test_augment_identity_is_totensor_normalize():
    src = synthetic_source(4, 32, seed=2)
    out = augment_cpu(src.images, draw_aug_params(4, False, None))
    ref = (src.images.permute(0, 3, 1, 2).float() / 255 - 0.5) / 0.5
    assert torch.allclose(out, ref)


def test_augment_matches_pad_crop_flip_reference():
    """Oracle: explicit zero-pad(8) -> crop -> flip -> ToTensor -> Normalize (ref:data.py:11-15)."""
    src = synthetic_source(6, 32, seed=3)
    aug = torch.tensor([[0, 0, 0], [16, 16, 1], [3, 11, 1], [8, 8, 1], [13, 2, 0], [16, 0, 0]], dtype=torch.int32)
    out = augment_cpu(src.images, aug)
    for i in range(6):
        img = src.images[i].permute(2, 0, 1)
        padded = torch.nn.functional.pad(img, (8, 8, 8, 8))
        ox, oy, fl = aug[i].tolist()
        crop = padded[:, oy:oy + 32, ox:ox + 32]
        if fl:
            crop = crop.flip(-1)
        ref = (crop.float() / 255 - 0.5) / 0.5
        assert torch.allclose(out[i], ref), i


def test_loader_reshuffles_per_epoch_and_counts_batches():
    src = synthetic_source(50, 8, seed=4)
    ld = BatchLoader(src, ShardSampler(50, 2, 0, True, 0), 8, True)
    assert len(ld) == 4                                 # ceil(25 / 8), no drop_last (ref:data.py:41-53)
    ld.set_epoch(1); a = [b.labels.clone() for b in ld]
    ld.set_epoch(2); b = [x.labels.clone() for x in ld]
    assert [x.numel() for x in a] == [8, 8, 8, 1]
    assert not all(torch.equal(x, y) for x, y in zip(a, b))
