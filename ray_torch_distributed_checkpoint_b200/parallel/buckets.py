"""Gradient-bucket planning and flat arenas.

The bucket assignment reproduces what torch DDP computes for the reference's model (measured in
SURVEY.md §2.5-C3): parameters are taken in reverse registration order; a bucket is closed as soon as
its byte size reaches the current cap (first cap 1 MiB, then 25 MiB). For the 784-512-512-10 MLP this
gives bucket0 = {6.bias, 6.weight, 3.bias, 3.weight} = 1 071 144 B and bucket1 = {0.bias, 0.weight}
= 1 607 680 B in fp32 (reference: implicit in ``prepare_model`` at
``/root/reference/my_ray_module.py:135``).
"""
from __future__ import annotations

import dataclasses
from typing import Dict, Iterable, List, Sequence, Tuple

import torch

DEFAULT_FIRST_BUCKET_BYTES = 1024 * 1024
DEFAULT_BUCKET_CAP_BYTES = 25 * 1024 * 1024
ARENA_ALIGN_BYTES = 128      # every parameter starts on a 128-byte line inside the arena


@dataclasses.dataclass
class ParamSlot:
    name: str
    shape: Tuple[int, ...]
    numel: int
    offset: int            # element offset inside the arena
    bucket: int


@dataclasses.dataclass
class Bucket:
    index: int
    slots: List[ParamSlot]
    offset: int            # element offset of the bucket inside the arena
    numel: int             # padded element count (multiple of the alignment)
    payload_bytes: int     # unpadded bytes, as torch DDP would count them


def assign_buckets(sizes_bytes: Sequence[int], first_cap: int = DEFAULT_FIRST_BUCKET_BYTES,
                   cap: int = DEFAULT_BUCKET_CAP_BYTES) -> List[List[int]]:
    """Indices (into ``sizes_bytes``, which must already be in reduction order) grouped into buckets."""
    buckets: List[List[int]] = []
    cur: List[int] = []
    cur_bytes = 0
    limit = first_cap
    for i, nb in enumerate(sizes_bytes):
        cur.append(i)
        cur_bytes += nb
        if cur_bytes >= limit:
            buckets.append(cur)
            cur, cur_bytes, limit = [], 0, cap
    if cur:
        buckets.append(cur)
    return buckets


def plan(named_params: Iterable[Tuple[str, torch.Tensor]], first_cap: int = DEFAULT_FIRST_BUCKET_BYTES,
         cap: int = DEFAULT_BUCKET_CAP_BYTES, element_size: int = 4) -> Tuple[List[Bucket], int]:
    """Lay parameters out in one flat arena, bucket after bucket, in reverse registration order.

    Returns (buckets, total_elements). Offsets are in elements of ``element_size`` bytes.
    """
    items = [(n, p) for n, p in named_params if p.requires_grad]
    items.reverse()
    groups = assign_buckets([p.numel() * element_size for _, p in items], first_cap, cap)
    align = ARENA_ALIGN_BYTES // element_size
    buckets: List[Bucket] = []
    cursor = 0
    for b, idxs in enumerate(groups):
        start = cursor
        slots = []
        payload = 0
        for i in idxs:
            name, p = items[i]
            slots.append(ParamSlot(name, tuple(p.shape), p.numel(), cursor, b))
            payload += p.numel() * element_size
            cursor = (cursor + p.numel() + align - 1) // align * align
        buckets.append(Bucket(b, slots, start, cursor - start, payload))
    return buckets, cursor
