"""One process per GPU: spawn, bind, watch, and propagate the first failure.

Parity target: ``run_model`` / ``mp.spawn(main, args=(world_size,), nprocs=world_size, join=True)``
(ref:main.py:180-188) and what ``torch.multiprocessing.spawn`` does underneath
(torch:multiprocessing/spawn.py:79-96, 118-211): 'spawn' start method, child traceback shipped to
the parent, first failure terminates the siblings and is re-raised in the parent.

Differences by design: the rendezvous endpoint is unique per run (the reference hard-codes
127.0.0.1:20080, so two runs on one host collide -- SURVEY.md D9); children get their device and
CPU affinity bound before user code runs; a hung sibling is SIGTERMed then SIGKILLed after a grace
period; an optional watchdog timeout bounds the whole job.
"""
from __future__ import annotations

import multiprocessing as mp
import os
import signal
import socket
import sys
import time
import traceback
import uuid
from typing import Any, Callable, Optional, Sequence


class RankFailure(RuntimeError):
    """Raised in the parent when a rank died; carries the rank, exit code and child traceback."""

    def __init__(self, rank: int, exitcode: Optional[int], tb: str = '') -> None:
        self.rank, self.exitcode, self.child_traceback = rank, exitcode, tb
        how = f'exit code {exitcode}' if (exitcode is None or exitcode >= 0) else f'signal {-exitcode}'
        msg = f'rank {rank} failed ({how})'
        if tb:
            msg += '\n\n-- traceback from rank %d --\n%s' % (rank, tb)
        super().__init__(msg)


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(('127.0.0.1', 0))
        return int(s.getsockname()[1])


def _bind_cpu_affinity(rank: int, world_size: int) -> None:
    """Give each rank a contiguous slice of the allowed CPUs (keeps the launch thread and the
    pinned-memory staging close to one NUMA node on a 2-socket HGX board)."""
    try:
        cpus = sorted(os.sched_getaffinity(0))
        per = len(cpus) // world_size
        if per >= 2:
            os.sched_setaffinity(0, set(cpus[rank * per:(rank + 1) * per]))
    except (AttributeError, OSError):
        pass


def _child(fn: Callable, rank: int, world_size: int, args: Sequence[Any], env: dict, err_q) -> None:
    try:
        os.environ.update(env)
        os.environ['RANK'] = str(rank)
        os.environ['LOCAL_RANK'] = str(rank)
        os.environ['WORLD_SIZE'] = str(world_size)
        try:                                     # die with the parent (torch sets PDEATHSIG too)
            import ctypes
            ctypes.CDLL('libc.so.6', use_errno=True).prctl(1, signal.SIGTERM)
        except Exception:
            pass
        if os.environ.get('PMDIST_BIND_CPUS', '1') == '1':
            _bind_cpu_affinity(rank, world_size)
        fn(rank, world_size, *args)
    except KeyboardInterrupt:
        sys.exit(130)
    except BaseException:
        err_q.put((rank, traceback.format_exc()))
        sys.exit(1)


def spawn(fn: Callable, world_size: int, args: Sequence[Any] = (), timeout: Optional[float] = None,
          grace: float = 5.0, env: Optional[dict] = None) -> None:
    """Run ``fn(rank, world_size, *args)`` in ``world_size`` fresh processes and join them.

    Raises ``RankFailure`` naming the first rank that failed; the other ranks are terminated."""
    ctx = mp.get_context('spawn')
    run_env = {
        'MASTER_ADDR': '127.0.0.1',
        'MASTER_PORT': str(free_port()),
        'PMDIST_RUN_ID': uuid.uuid4().hex[:12],
    }
    run_env.update(env or {})
    err_q = ctx.SimpleQueue()
    procs = []
    for r in range(world_size):
        p = ctx.Process(target=_child, args=(fn, r, world_size, tuple(args), run_env, err_q), daemon=False)
        p.start()
        procs.append(p)

    deadline = None if timeout is None else time.time() + timeout
    failed: Optional[int] = None
    try:
        while True:
            alive = False
            for r, p in enumerate(procs):
                if p.exitcode is None:
                    alive = True
                elif p.exitcode != 0 and failed is None:
                    failed = r
            if failed is not None or not alive:
                break
            if deadline is not None and time.time() > deadline:
                failed = -1
                break
            time.sleep(0.05)
    finally:
        if failed is not None:
            for p in procs:
                if p.exitcode is None:
                    p.terminate()
            t_end = time.time() + grace
            for p in procs:
                p.join(max(0.0, t_end - time.time()))
                if p.exitcode is None:
                    p.kill()
        for p in procs:
            p.join()

    if failed is None:
        return
    if failed == -1:
        raise TimeoutError(f'job exceeded the {timeout}s watchdog; all ranks were terminated')
    tbs = {}
    while not err_q.empty():
        r, tb = err_q.get()
        tbs[r] = tb
    # report the rank whose traceback we hold if the first observed exit was collateral damage
    rank = failed if failed in tbs or not tbs else sorted(tbs)[0]
    raise RankFailure(rank, procs[rank].exitcode, tbs.get(rank, ''))
