"""Relocalization descriptor database sharded over the ranks of a torch.distributed group
(BASELINE.json config 4).

Rows are split into contiguous ranges in increasing global index, one per rank; queries are
replicated.  Each rank scans its shard on its GPU (orbm_top2_search); the per-shard
(best1, best2, idx) triples -- nq x 16 bytes per rank -- are exchanged with ONE all-gather (NCCL
over NVLink on GPUs; gloo in the CPU tests of the exchange + merge) and merged exactly
(orbm_merge_top2): lowest shard wins ties, i.e. the lowest global row, as one sequential scan
with the rule of src/ORBmatcher.cc:417-426 would.
"""
from __future__ import annotations

import numpy as np

from . import matcher
from ._lib import TOP2_DTYPE


def shard_range(n_rows: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous block partition [lo, hi) of the database rows for `rank`."""
    return n_rows * rank // world, n_rows * (rank + 1) // world


def exchange_top2(local: np.ndarray, group=None, device=None) -> np.ndarray:
    """All-gathers one rank's top-2 results; returns a (world, nq) TOP2_DTYPE array in rank order.
    `device`: torch device the collective runs on (cuda:N for NCCL, None/cpu for gloo)."""
    import torch
    import torch.distributed as dist
    local = np.ascontiguousarray(local, TOP2_DTYPE)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local.reshape(1, -1)
    world = dist.get_world_size(group)
    t = torch.from_numpy(local.view(np.int64).reshape(-1))  # 2 x int64 per query
    if device is not None:
        t = t.to(device)
    out = torch.empty(world * t.numel(), dtype=torch.int64, device=t.device)
    dist.all_gather_into_tensor(out, t, group=group)
    return out.cpu().numpy().view(TOP2_DTYPE).reshape(world, -1)


def merge_shards(gathered: np.ndarray, ratio: float = 0.75, th: int = matcher.TH_LOW):
    """(world, nq) per-shard results in row order -> (merged, accept) of the whole database."""
    return matcher.merge_top2(list(gathered), ratio, th)


class ShardedDescriptorDB:
    """This rank's shard of an N-row database plus the exchange/merge step."""

    def __init__(self, n_rows_total: int, rank: int, world: int, device: int = 0, group=None):
        self.rank, self.world, self.group, self.device = rank, world, group, device
        self.lo, self.hi = shard_range(n_rows_total, rank, world)
        self.shard = matcher.DescriptorDB(device, row_base=self.lo)

    def append(self, rows):
        self.shard.append(rows)

    def append_device(self, dev_ptr: int, n: int):
        self.shard.append_device(dev_ptr, n)

    def search(self, queries, ratio: float = 0.75, th: int = matcher.TH_LOW):
        """(merged, accept) of the whole database on every rank.  On GPUs the whole step runs inside liborbfront.so
        (matcher.ShardGroup: scan -> ncclAllGather -> device merge); the torch.distributed + host merge below is the
        form the CPU (gloo) tests of the exchange use."""
        import torch.distributed as dist
        if self.world > 1 and dist.is_initialized() and dist.get_backend(self.group) == "nccl":
            if getattr(self, "_group", None) is None:
                self._group = matcher.ShardGroup.from_torch(self.shard, self.group)
            return self._group.search(queries, ratio, th)
        local = self.shard.top2(queries)
        gathered = exchange_top2(local, self.group, None if self.world == 1 else f"cuda:{self.device}")
        return merge_shards(gathered, ratio, th)
