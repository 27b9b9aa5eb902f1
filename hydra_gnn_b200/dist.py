"""Graph-sharded data parallelism.

The graphs of a batch are independent connected components, so the forward shards by graph id
with NO exchange on the data path.  The only collective of the inference path is the gather of
the per-rank room logits (ragged: the room count differs per rank); it is issued on the
process group's backend -- NCCL over NVLink on the GPU box, gloo in the CPU tests.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(n_graphs: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous graph-id range ``[lo, hi)`` of ``rank``: sizes differ by at most one graph and
    concatenating the ranges in rank order gives ``0..n_graphs`` (so gathered logits are in graph order)."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    base, rem = divmod(n_graphs, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class LogitGather:
    """Ragged ``all_gather`` of ``[rows_r, width]`` fp32 logits.

    ``prepare(rows)`` exchanges the row counts once per batch layout (a host sync, outside the
    timed step); ``__call__(logits)`` is then ONE ``all_gather_into_tensor`` on a max-padded
    buffer with no host synchronisation, and ``unpad()`` returns the rows in rank (= graph) order.
    """

    def __init__(self, width: int, group: Optional[dist.ProcessGroup] = None):
        self.width, self.group = width, group
        self.world = dist.get_world_size(group)
        self.counts: List[int] = []
        self._pad = self._buf = None

    def prepare(self, rows: int, device) -> "LogitGather":
        mine = torch.tensor([rows], dtype=torch.int64, device=device)
        every = torch.empty(self.world, dtype=torch.int64, device=device)
        dist.all_gather_into_tensor(every, mine, group=self.group)
        self.counts = [int(c) for c in every.tolist()]
        self.rows, self.max_rows = rows, max(self.counts + [1])
        self._pad = torch.zeros(self.max_rows, self.width, dtype=torch.float32, device=device)
        self._buf = torch.empty(self.world * self.max_rows, self.width, dtype=torch.float32, device=device)
        return self

    def __call__(self, logits: torch.Tensor) -> torch.Tensor:
        if logits.size(0) != self.rows or logits.size(1) != self.width:
            raise ValueError("logits shape changed since prepare()")
        self._pad[: self.rows].copy_(logits)
        dist.all_gather_into_tensor(self._buf, self._pad, group=self.group)
        return self._buf

    def unpad(self) -> torch.Tensor:
        view = self._buf.view(self.world, self.max_rows, self.width)
        return torch.cat([view[r, :c] for r, c in enumerate(self.counts)], dim=0)
