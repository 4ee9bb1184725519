"""Parameter-dimension sharding for multi-GPU aggregation (one process per GPU).

GPU g owns columns [lo_g, hi_g) of every client's update and the same slice of the
output.  Median, trimmed mean, FedAvg and the weighted combine need no communication;
Krum all-reduces its partial n x n FP64 Gram matrix and FLTrust / clipping all-reduce
2n+1 FP64 partial sums (+ n fp32 maxima) -- one NCCL collective each, issued through
torch.distributed by flagg.engine (``group=``).  Pure host logic: testable with gloo.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import torch

SHARD_ALIGN = 1024  # shard boundaries at multiples of 1024 columns; the last shard takes the tail


def shard_bounds(d: int, world: int, align: int = SHARD_ALIGN) -> List[Tuple[int, int]]:
    """Column range of every rank: equal multiples of ``align``, tail on the last rank."""
    if world < 1:
        raise ValueError("world must be >= 1")
    per = ((d + world - 1) // world + align - 1) // align * align
    out = []
    for r in range(world):
        lo = min(d, r * per)
        hi = d if r == world - 1 else min(d, (r + 1) * per)
        out.append((lo, hi))
    return out


def shard_update_lists(client_updates: Sequence[Sequence[torch.Tensor]], rank: int, world: int):
    """Cut every client's tensor list down to this rank's column range of the flattened
    parameter vector.  Returns (lists of 1-D views, (lo, hi)); no data is copied."""
    numel = [int(t.numel()) for t in client_updates[0]]
    d = sum(numel)
    lo, hi = shard_bounds(d, world)[rank]
    out = []
    for upd in client_updates:
        row, off = [], 0
        for t, m in zip(upd, numel):
            a, b = max(lo, off), min(hi, off + m)
            if b > a:
                row.append(t.reshape(-1)[a - off:b - off])
            off += m
        if not row:      # empty shard (lo == hi): one zero-length tensor, so the rank still joins the collectives with zeros
            row.append(upd[0].reshape(-1)[:0])
        out.append(row)
    return out, (lo, hi)
