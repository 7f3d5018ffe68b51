"""Gradient bucket planner and flat-slab layout.

What the reference gets from ``DDP(model)`` (ref:main.py:44) is torch's C++ ``Reducer``: it
assigns parameters to buckets in *gradient-arrival* order with a 1 MiB first bucket and 25 MiB
later buckets (torch:include/torch/csrc/distributed/c10d/reducer.hpp:30-31, 593), which for the
reference's ResNet18 gives 9,990,184 B + 9,622,784 B (SURVEY.md probe:1).

Here the plan is explicit and drives memory layout: all parameters live in ONE flat slab ordered
by gradient arrival (last layer first), so every bucket is a contiguous byte range of the
gradient slab and the fused reduce+SGD kernel (allreduce_sgd_kernel, csrc/comm.cu) is launched on a range
as soon as backward has produced it.  ``parity`` caps reproduce torch's bucket sizes exactly
(tested); the default caps are smaller because over NVSwitch the cost model is launch latency +
bytes/900 GB/s, not a ring.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence, Tuple

MiB = 1 << 20
TORCH_FIRST_BUCKET_BYTES = 1 * MiB      # torch _DEFAULT_FIRST_BUCKET_BYTES
TORCH_BUCKET_CAP_BYTES = 25 * MiB       # torch bucket_cap_mb default


@dataclass(frozen=True)
class ParamInfo:
    name: str
    numel: int
    index: int            # position in model.named_parameters()
    is_bn_bias: bool = False
    is_bn_weight: bool = False


@dataclass(frozen=True)
class SlabEntry:
    name: str
    index: int
    offset: int           # element offset inside the flat slab
    numel: int


@dataclass(frozen=True)
class Bucket:
    start: int            # element offset of the bucket in the slab
    numel: int
    entries: Tuple[SlabEntry, ...]

    def nbytes(self, itemsize: int = 4) -> int:
        return sum(e.numel for e in self.entries) * itemsize


def arrival_order(params: Sequence[ParamInfo]) -> List[ParamInfo]:
    """Order in which autograd produces gradients for a feed-forward net: reverse registration
    order, except that a (weight, bias) pair of one BatchNorm arrives weight-first.  Matches the
    order torch's Reducer observed for the reference model: 37 36 34 35 33 31 32 30 | 28 29 27 ...
    (SURVEY.md probe:1)."""
    rev = list(reversed(params))
    out: List[ParamInfo] = []
    i = 0
    while i < len(rev):
        if i + 1 < len(rev) and rev[i].is_bn_bias and rev[i + 1].is_bn_weight:
            out.extend([rev[i + 1], rev[i]])
            i += 2
        else:
            out.append(rev[i])
            i += 1
    return out


def plan_buckets(params: Sequence[ParamInfo], first_cap_bytes: int, cap_bytes: int,
                 itemsize: int = 4, align_elems: int = 1, tail_cap_bytes: int = 0) -> Tuple[List[Bucket], int]:
    """Greedy assignment in arrival order: a bucket closes as soon as its size reaches its cap
    (the first bucket uses ``first_cap_bytes``).  ``align_elems`` pads every tensor's slab offset
    (the native path uses 64 elements = 256 B so vector loads and per-peer slices stay aligned).
    ``tail_cap_bytes`` > 0 splits the LAST gradients to arrive (the stem) off into a final bucket of at most that many
    bytes (slab span, alignment padding included): nothing is left to overlap the step's last all-reduce, so its
    latency is exposed, and a tiny range can use the barrier-free push all-reduce (csrc/comm.cu allreduce_sgd_ll_kernel).
    Returns (buckets, total slab elements)."""
    order = arrival_order(params)
    buckets: List[Bucket] = []
    cur: List[SlabEntry] = []
    cur_start = 0
    cur_payload = 0
    off = 0
    cap = first_cap_bytes
    for p in order:
        if align_elems > 1:
            off = (off + align_elems - 1) // align_elems * align_elems
        if not cur:
            cur_start = off
        cur.append(SlabEntry(p.name, p.index, off, p.numel))
        off += p.numel
        cur_payload += p.numel * itemsize
        if cur_payload >= cap:
            buckets.append(Bucket(cur_start, off - cur_start, tuple(cur)))
            cur, cur_payload, cap = [], 0, cap_bytes
    if cur:
        buckets.append(Bucket(cur_start, off - cur_start, tuple(cur)))
    if tail_cap_bytes > 0 and buckets and len(buckets[-1].entries) > 1:
        last = buckets[-1]
        end = last.start + last.numel
        if last.numel * itemsize > tail_cap_bytes:        # (a last bucket that already fits IS the tail)
            k = len(last.entries)                         # entries[k:] will form the tail bucket
            while k - 1 >= 1 and (end - last.entries[k - 1].offset) * itemsize <= tail_cap_bytes:
                k -= 1
            if k < len(last.entries):
                split_at = last.entries[k].offset
                buckets[-1] = Bucket(last.start, split_at - last.start, last.entries[:k])
                buckets.append(Bucket(split_at, end - split_at, last.entries[k:]))
    if align_elems > 1:
        off = (off + align_elems - 1) // align_elems * align_elems
    return buckets, off


def params_of(model) -> List[ParamInfo]:
    import torch.nn as nn
    bn_w, bn_b = set(), set()
    for mod_name, mod in model.named_modules():
        if isinstance(mod, nn.modules.batchnorm._BatchNorm):
            bn_w.add(f'{mod_name}.weight')
            bn_b.add(f'{mod_name}.bias')
    return [ParamInfo(n, p.numel(), i, n in bn_b, n in bn_w)
            for i, (n, p) in enumerate(model.named_parameters())]
