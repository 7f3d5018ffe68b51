"""ctypes binding of libpmdist.so (C ABI declared in include/pmdist.h).

The library is built in-tree by tools/build_native.py (``__graft_entry__.build()`` calls it).  On a
machine with a GPU a missing or unloadable library is a hard error -- there is no silent PyTorch
fallback for the native ops.
"""
from __future__ import annotations

import ctypes as C
import os
import sys
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('PMDIST_LIB') or os.path.join(os.path.dirname(_HERE), '_lib', 'libpmdist.so')   # PMDIST_LIB: A/B builds
REPO_ROOT = os.path.dirname(os.path.dirname(_HERE))
_lib: Optional[C.CDLL] = None


class PmdConvShape(C.Structure):
    _fields_ = [('N', C.c_int), ('H', C.c_int), ('W', C.c_int), ('C', C.c_int), ('K', C.c_int),
                ('R', C.c_int), ('stride', C.c_int)]


class NativeError(RuntimeError):
    pass


def build(force: bool = False) -> str:
    """Builds the library in-tree (tools/build_native.py).  Serialised with a file lock: under torchrun all
    ranks of a fresh checkout arrive here at once and must not run nvcc over each other's objects."""
    import fcntl
    sys.path.insert(0, os.path.join(REPO_ROOT, 'tools'))
    os.makedirs(os.path.dirname(LIB_PATH), exist_ok=True)
    try:
        import build_native
        with open(os.path.join(os.path.dirname(LIB_PATH), '.build.lock'), 'w') as lock:
            fcntl.flock(lock, fcntl.LOCK_EX)
            try:
                if os.path.exists(LIB_PATH) and not force:
                    return build_native.build(force=False)      # up-to-date check only (another rank built it)
                return build_native.build(force=force)
            finally:
                fcntl.flock(lock, fcntl.LOCK_UN)
    finally:
        sys.path.pop(0)


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        _lib = C.CDLL(LIB_PATH)
        _lib.pmd_last_error.restype = C.c_char_p
        _lib.pmd_launch_count.restype = C.c_ulonglong
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise NativeError(lib().pmd_last_error().decode('utf-8', 'replace'))


def launch_count() -> int:
    return int(lib().pmd_launch_count())


def ptr(t) -> C.c_void_p:
    """Device (or host) pointer of a torch tensor, or NULL for None."""
    return C.c_void_p(0 if t is None else t.data_ptr())


def stream_ptr(stream=None) -> C.c_void_p:
    import torch
    s = torch.cuda.current_stream() if stream is None else stream
    return C.c_void_p(s.cuda_stream)
