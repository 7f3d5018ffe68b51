"""PeerComm: the product transport -- symmetric peer-mapped memory over NVLink 5 / NVSwitch.

Python face of csrc/runtime.cpp + csrc/comm.cu (C ABI in include/pmdist_comm.h).  Replaces
``dist.init_process_group('nccl')`` and the NCCL communicator (ref:main.py:190-193): ranks meet on
an abstract unix socket named after the run id, exchange CUDA VMM handles as file descriptors and
map each other's heaps; collectives are kernels that read and write peer heaps directly.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Sequence

import torch

from ..ops import _lib
from ..ops._lib import check
from .comm import Comm

MAX_RANKS = 8
MAX_SLOTS = 4096
ALGOS = {'auto': 0, 'oneshot': 1, 'twoshot': 2, 'nvls': 3, 'll': 4}
TRANSPORTS = {0: 'single', 1: 'vmm', 2: 'cuda-ipc'}


class PmdCommInfo(C.Structure):
    _fields_ = [('rank', C.c_int), ('world', C.c_int), ('device', C.c_int), ('transport', C.c_int),
                ('multicast_supported', C.c_int), ('multicast_mapped', C.c_int), ('p2p_ok', C.c_int), ('pad_', C.c_int),
                ('heap_bytes', C.c_ulonglong), ('local_base', C.c_ulonglong), ('mc_base', C.c_ulonglong),
                ('d_bases', C.c_ulonglong), ('note', C.c_char * 256)]


class PmdSym(C.Structure):
    _fields_ = [('rank', C.c_int), ('world', C.c_int), ('d_bases', C.c_void_p), ('mc_base', C.c_ulonglong)]


class _RawCuda:
    """Minimal __cuda_array_interface__ carrier so torch can alias memory the runtime allocated."""

    def __init__(self, ptr: int, nbytes: int) -> None:
        self.__cuda_array_interface__ = {'shape': (nbytes,), 'typestr': '|u1', 'data': (ptr, False), 'version': 3}


def alias_tensor(ptr: int, nbytes: int, device: torch.device) -> torch.Tensor:
    return torch.as_tensor(_RawCuda(ptr, nbytes), device=device)


def run_id_from_env() -> str:
    rid = os.environ.get('PMDIST_RUN_ID')
    if rid:
        return rid
    # torchrun: every rank sees the same MASTER_PORT (+ run id when set)
    return 'p%s_%s' % (os.environ.get('MASTER_PORT', '0'), os.environ.get('TORCHELASTIC_RUN_ID', 'none'))


class PeerComm(Comm):
    kind = 'p2p'

    def __init__(self, rank: int, world_size: int, device: torch.device, heap_bytes: int = 256 << 20,
                 run_id: str | None = None, force_ipc: bool = False, multicast: bool = True) -> None:
        if world_size > MAX_RANKS:
            raise ValueError(f'the peer transport is single-node: at most {MAX_RANKS} ranks')
        self.rank, self.world_size, self.device = rank, world_size, device
        self._lib = _lib.lib()
        self._h = C.c_void_p()
        flags = (1 if force_ipc or os.environ.get('PMDIST_FORCE_IPC') == '1' else 0) | (0 if multicast else 2)
        rid = (run_id or run_id_from_env()).encode()
        dev_index = device.index if device.index is not None else torch.cuda.current_device()
        check(self._lib.pmd_comm_create(C.c_int(rank), C.c_int(world_size), rid, C.c_int(dev_index),
                                        C.c_ulonglong(heap_bytes), C.c_int(flags), C.byref(self._h)))
        info = PmdCommInfo()
        check(self._lib.pmd_comm_info(self._h, C.byref(info)))
        self.info = info
        self.sym = PmdSym(info.rank, info.world, C.c_void_p(info.d_bases), info.mc_base)
        self.heap = alias_tensor(info.local_base, info.heap_bytes, device)
        self._scratch_off = None
        self._scratch_bytes = 0
        self._next_slot = MAX_SLOTS - 300      # slots [3796, 3896) for ad-hoc collectives issued from Python

    # ------------------------------------------------------------------ description
    @property
    def handle(self) -> C.c_void_p:
        return self._h

    def describe(self) -> dict:
        i = self.info
        return {'transport': TRANSPORTS.get(i.transport, '?'), 'world': i.world, 'p2p_ok': bool(i.p2p_ok),
                'multicast_supported': bool(i.multicast_supported), 'multicast_mapped': bool(i.multicast_mapped),
                'heap_mib': i.heap_bytes >> 20, 'note': i.note.decode('utf-8', 'replace')}

    # ------------------------------------------------------------------ heap
    def alloc(self, nbytes: int, align: int = 1024) -> int:
        lib = self._lib
        lib.pmd_comm_alloc.restype = C.c_longlong
        off = lib.pmd_comm_alloc(self._h, C.c_ulonglong(nbytes), C.c_ulonglong(align))
        if off < 0:
            check(-1)
        return int(off)

    def view(self, off: int, shape: Sequence[int], dtype: torch.dtype) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= s
        nbytes = n * torch.empty(0, dtype=dtype).element_size()
        return self.heap[off:off + nbytes].view(dtype).view(*shape)

    def peer_view(self, peer: int, off: int, shape: Sequence[int], dtype: torch.dtype) -> torch.Tensor:
        self._lib.pmd_comm_peer_base.restype = C.c_ulonglong
        base = self._lib.pmd_comm_peer_base(self._h, C.c_int(peer))
        n = 1
        for s in shape:
            n *= s
        nbytes = n * torch.empty(0, dtype=dtype).element_size()
        return alias_tensor(base + off, nbytes, self.device).view(dtype).view(*shape)

    def _scratch(self, nbytes: int) -> int:
        if self._scratch_off is None or nbytes > self._scratch_bytes:
            self._scratch_bytes = max(nbytes, 1 << 20)
            self._scratch_off = self.alloc(self._scratch_bytes)
        return self._scratch_off

    def _slot(self, blocks: int) -> int:
        return MAX_SLOTS - 300      # ad-hoc collectives are serialised on one stream: one slot range suffices

    # ------------------------------------------------------------------ Comm interface
    def barrier(self) -> None:
        torch.cuda.synchronize(self.device)
        check(self._lib.pmd_comm_host_barrier(self._h))

    def device_barrier(self, slot: int = MAX_SLOTS - 1) -> None:
        check(self._lib.pmd_dev_barrier(C.byref(self.sym), C.c_int(slot), _lib.stream_ptr()))

    def all_gather_bytes(self, payload: bytes) -> List[bytes]:
        w = self.world_size
        lens = (C.c_longlong * w)()
        mine = C.c_longlong(len(payload))
        check(self._lib.pmd_comm_allgather_host(self._h, C.byref(mine), C.c_int(8), lens))
        mx = max(int(x) for x in lens)
        buf = C.create_string_buffer(payload.ljust(mx, b'\0'), mx) if mx else C.create_string_buffer(1)
        out = C.create_string_buffer(max(mx * w, 1))
        if mx:
            check(self._lib.pmd_comm_allgather_host(self._h, buf, C.c_int(mx), out))
        return [out.raw[i * mx:i * mx + int(lens[i])] for i in range(w)]

    def broadcast_(self, t: torch.Tensor, src: int = 0) -> None:
        if self.world_size == 1:
            return
        flat = t.contiguous().view(-1).view(torch.uint8)
        nbytes = (flat.numel() + 15) // 16 * 16
        off = self._scratch(nbytes)
        stage = self.heap[off:off + flat.numel()]
        stage.copy_(flat)
        check(self._lib.pmd_broadcast(C.byref(self.sym), C.c_ulonglong(off), C.c_ulonglong(nbytes), C.c_int(src),
                                      C.c_int(32), C.c_int(self._slot(32)), _lib.stream_ptr()))
        t.copy_(stage.view(t.dtype).view(t.shape))

    def all_reduce_sum_(self, t: torch.Tensor, async_op: bool = False, algo: str = 'auto'):
        if self.world_size == 1:
            return None
        assert t.dtype == torch.float32, 'peer all-reduce is fp32'
        n = t.numel()
        n4 = (n + 3) // 4 * 4
        off = self._scratch(2 * n4 * 4 + 32)
        src = self.view(off, (n4,), torch.float32)
        dst_off = off + n4 * 4
        src.zero_()
        src[:n].copy_(t.reshape(-1))
        check(self._lib.pmd_allreduce_f32(C.byref(self.sym), C.c_ulonglong(off), C.c_ulonglong(dst_off), C.c_longlong(n4),
                                          C.c_int(ALGOS[algo]), C.c_int(32), C.c_int(self._slot(32)), _lib.stream_ptr()))
        t.copy_(self.view(dst_off, (n4,), torch.float32)[:n].view(t.shape))
        return None

    def reduce_metrics(self, values: Sequence[float]) -> List[float]:
        """Cross-rank SUM of [loss_sum, top1, top5, count] by the P2P metric kernel (the live version
        of the reference's dead reduce_tensor, ref:main.py:173-177)."""
        vals = torch.tensor(list(values), dtype=torch.float64)
        if self.world_size == 1:
            return vals.tolist()
        assert vals.numel() == 4
        off = self._scratch(1 << 20)
        pad = self.view(off, (4,), torch.float64)
        pad.copy_(vals.to(self.device))
        total = torch.empty(4, dtype=torch.float64, device=self.device)
        check(self._lib.pmd_metric_allreduce(C.byref(self.sym), C.c_int(MAX_SLOTS - 2), C.c_ulonglong(off),
                                             C.c_void_p(total.data_ptr()), _lib.stream_ptr()))
        return total.cpu().tolist()

    def close(self) -> None:
        if self._h:
            self.heap = None
            self._lib.pmd_comm_destroy(self._h)
            self._h = C.c_void_p()
