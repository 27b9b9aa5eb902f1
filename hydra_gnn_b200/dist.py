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


USE_PEER_COPIES = os.environ.get("HG_PEER_GATHER", "1") != "0"  # pipelined gather by copy-engine pulls + stream memory operations over symmetric (peer) memory; 0: NCCL all_gather_into_tensor (8 x B200: 0.836 vs 0.880 ms per step)


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
            ok = torch.tensor([int(self._peer is not None)], device=device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)  # all ranks take the same path, or a wait never ends
            if int(ok) == 0:
                self._peer = None
                self._pads = [torch.zeros(self.max_rows, self.width, dtype=torch.float32, device=device) for _ in range(slots)]
        if device.type == "cuda":
            self._side = torch.cuda.Stream(device)
            self._done = [torch.cuda.Event() for _ in range(slots)]
            self._staged = [torch.cuda.Event() for _ in range(slots)]
        self._last = 0
        return self

    def _peer_setup(self, device, slots):
        """Pipelined form over NVLink peer memory with NO kernel: the padded staging buffers and two int32 flag vectors per
        slot live in symmetric memory; a gather is, on the side stream,
            write `t` into every peer's ready[me]      (stream memory operation, cuStreamWriteValue32, over NVLink)
            wait until my ready[r] >= t for every r     (cuStreamWaitValue32: the stream waits, no SM spins)
            pull every peer's buffer                    (device-to-device copy engines)
            write `t` into every peer's read[me]; wait until my read[r] >= t   (the slot may be staged again)
        An NCCL all-gather kernel (or a barrier kernel) holds SMs for the whole exchange INCLUDING the skew between the
        ranks, and every SM it holds delays one CTA of the persistent one-CTA-per-SM kernels of the forward it overlaps
        (8 GPUs: 0.807 ms/step without a gather, 0.876 with the NCCL one).  Returns None (NCCL is used) when symmetric
        memory or stream memory operations are not available."""
        try:
            import torch.distributed._symmetric_memory as symm_mem
            from cuda.bindings import driver

            group = self.group if self.group is not None else dist.group.WORLD
            pads, handles, flags, fh = [], [], [], []
            for _ in range(slots):
                t = symm_mem.empty(self.max_rows, self.width, dtype=torch.float32, device=device)
                t.zero_()
                handles.append(symm_mem.rendezvous(t, group.group_name))
                pads.append(t)
                f = symm_mem.empty(2 * self.world, dtype=torch.int32, device=device)  # ready[world] | read[world]
                f.zero_()
                fh.append(symm_mem.rendezvous(f, group.group_name))
                flags.append(f)
            torch.cuda.synchronize(device)
            dist.barrier(group=self.group)  # every rank's flags are zero before anyone writes into them
            self._pads, self._flags, self._flag_handles = pads, flags, fh
            self._turn = [0] * slots
            self._driver, self._symm = driver, symm_mem
            return handles
        except Exception as e:  # noqa: BLE001 -- any failure here means "no peer memory": fall back to NCCL
            import warnings

            warnings.warn(f"LogitGather: peer-memory gather unavailable ({type(e).__name__}: {e}); using NCCL all_gather")
            return None

    def _signal_and_wait(self, slot: int, base: int, t: int) -> None:
        """flags[base + me] := t on every rank (mine included), then the side stream waits for flags[base + r] >= t, all r."""
        drv, me = self._driver, self._peer[slot].rank
        fh = self._flag_handles[slot]
        stream = drv.CUstream(self._side.cuda_stream)
        for r in range(self.world):
            peer_flags = fh.get_buffer(r, (2 * self.world,), torch.int32)  # rank r's flag vector, mapped here
            (err,) = drv.cuStreamWriteValue32(stream, drv.CUdeviceptr(peer_flags.data_ptr() + 4 * (base + me)), t,
                                              drv.CUstreamWriteValue_flags.CU_STREAM_WRITE_VALUE_DEFAULT)
            if err != drv.CUresult.CUDA_SUCCESS:
                raise RuntimeError(f"cuStreamWriteValue32: {err}")
        addr0 = self._flags[slot].data_ptr()
        for r in range(self.world):
            (err,) = drv.cuStreamWaitValue32(stream, drv.CUdeviceptr(addr0 + 4 * (base + r)), t,
                                             drv.CUstreamWaitValue_flags.CU_STREAM_WAIT_VALUE_GEQ)
            if err != drv.CUresult.CUDA_SUCCESS:
                raise RuntimeError(f"cuStreamWaitValue32: {err}")

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
        if self._side is None:  # CPU process group (gloo, the host-side tests): no streams, the exchange is synchronous
            dist.all_gather_into_tensor(self._bufs[slot], self._pads[slot], group=self.group)
            self._last = slot
            return
        cur = torch.cuda.current_stream()
        self._staged[slot].record(cur)
        with torch.cuda.stream(self._side):
            self._side.wait_event(self._staged[slot])
            if self._peer is not None:
                h = self._peer[slot]
                shape = (self.max_rows, self.width)
                chunks = self._bufs[slot].view(self.world, self.max_rows, self.width)
                self._turn[slot] += 1
                t = self._turn[slot]
                self._signal_and_wait(slot, 0, t)  # every rank has staged this slot
                for step in range(self.world):
                    r = (h.rank - step) % self.world
                    chunks[r].copy_(h.get_buffer(r, shape, torch.float32))  # device-to-device copy engine, pull over NVLink
                self._signal_and_wait(slot, self.world, t)  # every rank has read this rank's slot: it may be staged again
            else:
                dist.all_gather_into_tensor(self._bufs[slot], self._pads[slot], group=self.group)
            self._done[slot].record(self._side)
        self._last = slot

    def wait(self, slot: int = 0) -> torch.Tensor:
        """The current stream waits for the slot's gather (before its buffers are overwritten or read)."""
        if self._side is not None:
            torch.cuda.current_stream().wait_event(self._done[slot])
        return self._bufs[slot]

    def unpad(self, slot: Optional[int] = None) -> torch.Tensor:
        buf = self._bufs[self._last if slot is None else slot]
        view = buf.view(self.world, self.max_rows, self.width)
        return torch.cat([view[r, :c] for r, c in enumerate(self.counts)], dim=0)
