"""Graph-sharded data parallelism.

The graphs of a batch are independent connected components, so the forward shards by graph id
with NO exchange on the data path.  The only collective of the inference path is the gather of
the per-rank room logits (ragged: the room count differs per rank); it is issued on the
process group's backend -- NCCL over NVLink on the GPU box, gloo in the CPU tests.
"""
from __future__ import annotations

import os
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


USE_PEER_COPIES = os.environ.get("HG_PEER_GATHER", "0") == "1"  # opt-in: pipelined gather by copy-engine pulls over symmetric (peer) memory instead of the NCCL kernel (measured on 8 x B200: 0.887-0.896 vs 0.878-0.881 ms per step, so NCCL stays the default)


class LogitGather:
    """Ragged ``all_gather`` of ``[rows_r, width]`` fp32 logits.

    ``prepare(rows)`` exchanges the row counts once per batch layout (a host sync, outside the
    timed step); ``__call__(logits)`` is then ONE ``all_gather_into_tensor`` on a max-padded
    buffer with no host synchronisation, and ``unpad()`` returns the rows in rank (= graph) order.

    Pipelined form (``slots = 2``): the collective is latency-bound (8 x 2 MB over NVLink: ~0.1 ms against a
    0.8 ms forward), so a serving / evaluation loop keeps it off the critical path -- ``stage(logits, slot)``
    copies the step's logits into the slot's padded buffer on the current stream (capturable in the step's CUDA
    graph), ``launch(slot)`` issues the all-gather on a side stream behind that copy, and the next step's forward
    (other slot) runs concurrently; ``wait(slot)`` joins before the slot is reused or read.
    """

    def __init__(self, width: int, group: Optional[dist.ProcessGroup] = None):
        self.width, self.group = width, group
        self.world = dist.get_world_size(group)
        self.counts: List[int] = []
        self._pads, self._bufs, self._done, self._side = [], [], [], None

    def prepare(self, rows: int, device, slots: int = 1) -> "LogitGather":
        mine = torch.tensor([rows], dtype=torch.int64, device=device)
        every = torch.empty(self.world, dtype=torch.int64, device=device)
        dist.all_gather_into_tensor(every, mine, group=self.group)
        self.counts = [int(c) for c in every.tolist()]
        self.rows, self.max_rows = rows, max(self.counts + [1])
        self._pads = [torch.zeros(self.max_rows, self.width, dtype=torch.float32, device=device) for _ in range(slots)]
        self._bufs = [torch.empty(self.world * self.max_rows, self.width, dtype=torch.float32, device=device)
                      for _ in range(slots)]
        self._peer = None
        if device.type == "cuda" and slots > 1 and self.world > 1 and USE_PEER_COPIES:
            self._peer = self._peer_setup(device, slots)
        if device.type == "cuda":
            self._side = torch.cuda.Stream(device)
            self._done = [torch.cuda.Event() for _ in range(slots)]
            self._staged = [torch.cuda.Event() for _ in range(slots)]
        self._last = 0
        return self

    def _peer_setup(self, device, slots):
        """Pipelined form over NVLink peer memory: the padded staging buffers live in symmetric memory, and the gather is
        `barrier -> one copy-engine copy per peer (pull) -> barrier` on the side stream -- no SM is taken from the
        persistent one-CTA-per-SM kernels of the forward it overlaps (an NCCL all-gather kernel occupies SMs for the
        whole exchange).  Returns None (NCCL is used) when symmetric memory cannot be set up on this system."""
        try:
            import torch.distributed._symmetric_memory as symm_mem

            group = self.group if self.group is not None else dist.group.WORLD
            pads, handles = [], []
            for _ in range(slots):
                t = symm_mem.empty(self.max_rows, self.width, dtype=torch.float32, device=device)
                t.zero_()
                handles.append(symm_mem.rendezvous(t, group.group_name))
                pads.append(t)
            self._pads = pads
            return handles
        except Exception as e:  # noqa: BLE001 -- any failure here means "no peer memory": fall back to NCCL
            import warnings

            warnings.warn(f"LogitGather: symmetric memory unavailable ({type(e).__name__}: {e}); using NCCL all_gather")
            return None

    def __call__(self, logits: torch.Tensor) -> torch.Tensor:
        self.stage(logits, 0)
        dist.all_gather_into_tensor(self._bufs[0], self._pads[0], group=self.group)
        self._last = 0
        return self._bufs[0]

    def stage(self, logits: torch.Tensor, slot: int = 0) -> None:
        if logits.size(0) != self.rows or logits.size(1) != self.width:
            raise ValueError("logits shape changed since prepare()")
        self._pads[slot][: self.rows].copy_(logits)

    def launch(self, slot: int = 0) -> None:
        """All-gather of the slot's staged logits on the side stream, ordered after everything queued so far on the
        current stream (the step that staged them)."""
        cur = torch.cuda.current_stream()
        self._staged[slot].record(cur)
        with torch.cuda.stream(self._side):
            self._side.wait_event(self._staged[slot])
            if self._peer is not None:
                h = self._peer[slot]
                shape = (self.max_rows, self.width)
                chunks = self._bufs[slot].view(self.world, self.max_rows, self.width)
                h.barrier()  # every rank has staged this slot
                for step in range(self.world):
                    r = (h.rank - step) % self.world
                    chunks[r].copy_(h.get_buffer(r, shape, torch.float32))  # device-to-device copy engine, pull over NVLink
                h.barrier()  # every rank has read this rank's slot: it may be staged again
            else:
                dist.all_gather_into_tensor(self._bufs[slot], self._pads[slot], group=self.group)
            self._done[slot].record(self._side)
        self._last = slot

    def wait(self, slot: int = 0) -> torch.Tensor:
        """The current stream waits for the slot's gather (before its buffers are overwritten or read)."""
        torch.cuda.current_stream().wait_event(self._done[slot])
        return self._bufs[slot]

    def unpad(self, slot: Optional[int] = None) -> torch.Tensor:
        buf = self._bufs[self._last if slot is None else slot]
        view = buf.view(self.world, self.max_rows, self.width)
        return torch.cat([view[r, :c] for r, c in enumerate(self.counts)], dim=0)
