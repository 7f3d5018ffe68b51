"""CIFAR-10 on-disk source: reads the python pickles torchvision's CIFAR10(download=False) expects
(ref:data.py:21-28), without torchvision."""
import os
import pickle

import numpy as np
import pytest
import torch

from pmdist.data.sources import cifar10_source, make_sources


def _write_fake_cifar(root, n_per_batch=20):
    base = os.path.join(root, 'cifar-10-batches-py')
    os.makedirs(base)
    rng = np.random.default_rng(0)
    for name in [f'data_batch_{i}' for i in range(1, 6)] + ['test_batch']:
        d = {'data': rng.integers(0, 256, size=(n_per_batch, 3072), dtype=np.uint8),
             'labels': rng.integers(0, 10, size=n_per_batch).tolist()}
        with open(os.path.join(base, name), 'wb') as f:
            pickle.dump(d, f)
    return base


def test_cifar10_pickles_become_nhwc_uint8(tmp_path):
    _write_fake_cifar(str(tmp_path))
    tr, te = cifar10_source(str(tmp_path), True), cifar10_source(str(tmp_path), False)
    assert tr.images.shape == (100, 32, 32, 3) and tr.images.dtype == torch.uint8 and len(te) == 20
    raw = pickle.load(open(tmp_path / 'cifar-10-batches-py' / 'data_batch_1', 'rb'))
    chw = torch.from_numpy(np.asarray(raw['data'][3]).reshape(3, 32, 32))
    assert torch.equal(tr.images[3].permute(2, 0, 1), chw)            # planar RGB rows -> HWC
    assert tr.labels[:20].tolist() == raw['labels']


def test_missing_dataset_is_a_clear_error_and_auto_falls_back(tmp_path):
    with pytest.raises(FileNotFoundError, match='download=False'):
        cifar10_source(str(tmp_path / 'nope'))
    tr, te = make_sources('auto', str(tmp_path / 'nope'), 32, 64, 32)
    assert tr.name.startswith('synthetic') and len(tr) == 64 and len(te) == 32
