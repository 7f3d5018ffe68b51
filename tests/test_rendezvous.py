"""Native rendezvous (csrc/runtime.cpp) on CPU: all-gather of blobs, barrier, and file-descriptor
exchange with SCM_RIGHTS (the mechanism that carries CUDA VMM handles between ranks), plus failure
detection when a peer dies."""
import ctypes as C
import os
import time

import pytest

from pmdist.parallel import launch


def _rdv_worker(rank, world, tmp):
    from pmdist.ops import _lib
    lib = _lib.lib()
    h = C.c_void_p()
    rid = os.environ['PMDIST_RUN_ID'].encode()
    assert lib.pmd_rdv_create(C.c_int(rank), C.c_int(world), rid, C.c_double(30.0), C.byref(h)) == 0, lib.pmd_last_error()
    # all-gather of 16-byte blobs
    mine = ('rank-%02d-payload!' % rank).encode()
    out = C.create_string_buffer(16 * world)
    assert lib.pmd_rdv_allgather(h, mine, C.c_int(16), out) == 0
    assert [out.raw[i * 16:(i + 1) * 16] for i in range(world)] == [('rank-%02d-payload!' % r).encode() for r in range(world)]
    assert lib.pmd_rdv_barrier(h) == 0
    # fd exchange: every rank shares a file it wrote; every rank must be able to read all of them
    path = os.path.join(tmp, f'secret{rank}')
    with open(path, 'w') as f:
        f.write(f'hello from {rank}')
    fd = os.open(path, os.O_RDONLY)
    os.unlink(path)                                   # only reachable through the descriptor now
    fds = (C.c_int * world)()
    assert lib.pmd_rdv_allgather_fds(h, C.c_int(fd), fds) == 0, lib.pmd_last_error()
    for r in range(world):
        # pread: descriptors received over SCM_RIGHTS share ONE file offset across all processes
        assert os.pread(fds[r], 64, 0).decode() == f'hello from {r}'
        os.close(fds[r])
    os.close(fd)
    assert lib.pmd_rdv_barrier(h) == 0
    lib.pmd_rdv_destroy(h)
    open(os.path.join(tmp, f'ok{rank}'), 'w').close()


@pytest.mark.parametrize('world', [2, 4])
def test_allgather_barrier_and_fd_passing(tmp_path, world):
    launch.spawn(_rdv_worker, world, args=(str(tmp_path),), timeout=120, env={'PMDIST_BIND_CPUS': '0'})
    assert all((tmp_path / f'ok{r}').exists() for r in range(world))


def _dying_peer(rank, world):
    from pmdist.ops import _lib
    lib = _lib.lib()
    h = C.c_void_p()
    rid = os.environ['PMDIST_RUN_ID'].encode()
    assert lib.pmd_rdv_create(C.c_int(rank), C.c_int(world), rid, C.c_double(20.0), C.byref(h)) == 0
    if rank == 1:
        os._exit(0)                                   # vanish without saying goodbye
    t0 = time.time()
    out = C.create_string_buffer(8 * world)
    rc = lib.pmd_rdv_allgather(h, b'12345678', C.c_int(8), out)
    assert rc != 0 and time.time() - t0 < 15          # detected as a closed socket, not a 20 s timeout
    assert b'peer' in lib.pmd_last_error() or b'allgather' in lib.pmd_last_error()


def test_dead_peer_is_detected_not_hung():
    launch.spawn(_dying_peer, 2, timeout=60, env={'PMDIST_BIND_CPUS': '0'})
