import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a real B200 (run with -m gpu on the GPU box)')
    config.addinivalue_line('markers', 'multigpu: needs >= 2 B200s on one box')


def pytest_collection_modifyitems(config, items):
    import torch
    have = torch.cuda.is_available()
    ngpu = torch.cuda.device_count() if have else 0
    for item in items:
        if 'gpu' in item.keywords and not have:
            item.add_marker(pytest.mark.skip(reason='no CUDA device in this container'))
        if 'multigpu' in item.keywords and ngpu < 2:
            item.add_marker(pytest.mark.skip(reason='needs >= 2 GPUs'))


@pytest.fixture(scope='session')
def ref_modules():
    """The reference's own utils.py / model/resnet.py imported read-only as oracles, with the one
    shim they need on Python >= 3.10 (collections.Iterable)."""
    import collections
    import collections.abc
    import importlib.util
    if not os.path.isdir('/root/reference'):
        pytest.skip('reference checkout not mounted')
    if not hasattr(collections, 'Iterable'):
        collections.Iterable = collections.abc.Iterable
    out = {}
    for name, rel in (('ref_utils', 'utils.py'), ('ref_resnet', 'model/resnet.py')):
        spec = importlib.util.spec_from_file_location(name, os.path.join('/root/reference', rel))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        out[name] = mod
    return out
