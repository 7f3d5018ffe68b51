"""Import alias for the framework package.

The framework lives in ``pytorch_multiprocessing-distributed_b200/`` (the name the
project brief prescribes).  A hyphen is not a legal Python identifier, so this tiny
package points its ``__path__`` at that directory: ``import pmdist.models.resnet``
loads ``pytorch_multiprocessing-distributed_b200/models/resnet.py``.
"""
import os as _os

_REAL = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "pytorch_multiprocessing-distributed_b200")
__path__ = [_REAL]
REPO_ROOT = _os.path.dirname(_REAL)
__version__ = "0.1.0"
