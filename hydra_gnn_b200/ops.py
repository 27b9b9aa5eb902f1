"""Thin tensor-level wrappers over the C ABI (``include/hydra_gnn_b200.h``).

Every function launches hand-written sm_100a kernels on the current torch stream; torch
is used only to own device memory.  No function here has a CPU or eager fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import torch
from torch import Tensor

from . import _lib
from ._lib import ACT, GROUP, PREC, REDUCE, check, lib, ptr, stream_ptr

DEFAULT_PRECISION = "auto"  # "auto": tcgen05 TF32x3 for M >= TC_MIN_ROWS, fp32 CUDA cores below
TC_MIN_ROWS = 512
FP32_SMALL_ROWS = int(os.environ.get("HG_FP32_SMALL_ROWS", "0"))  # A/B knob, off by default
FP32_SMALL_K = int(os.environ.get("HG_FP32_SMALL_K", "64"))

# Optional per-launch timing (bench.py's roofline leg): a list that receives
# (abi_name, meta dict, start_event, end_event) for every ABI call while it is set.
_profile_sink = None


USE_TILE_WINDOW = True  # pass the plan's per-tile source windows to hg_gat_fused (False: the v6 global-gather kernel)


def set_profile_sink(sink):
    global _profile_sink
    _profile_sink = sink


# Generation of "parameters changed behind torch's back": bumped by every optimizer step that writes weights through raw
# pointers (train.FlatAdam -> hg_adam_step leaves data_ptr and _version untouched).  Part of every packed-weight cache key.
_param_generation = 0


def param_generation() -> int:
    return _param_generation


def bump_param_generation() -> None:
    global _param_generation
    _param_generation += 1


# NVTX ranges around every ABI call (HG_NVTX=1 or set_nvtx(True)): `ncu --nvtx --nvtx-include "hg_gat_fused/"` then
# selects the launches of one operator; off by default (a push / pop pair per call on the launch path).
_nvtx = os.environ.get("HG_NVTX", "0") == "1"


def set_nvtx(on: bool) -> None:
    global _nvtx
    _nvtx = bool(on)


def _call(name: str, meta, fn):
    if _nvtx:
        torch.cuda.nvtx.range_push(name)
        try:
            _call_inner(name, meta, fn)
        finally:
            torch.cuda.nvtx.range_pop()
        return
    _call_inner(name, meta, fn)


# While a profile sink is set, every timed call is preceded by a device-side delay of this many clock cycles (outside the
# event pair): the host then enqueues start event, launch and end event while the GPU is still busy, so the pair brackets
# the kernel alone and not the host's launch path (ctypes marshalling, tensor-map encoding) on an idle GPU.
PROFILE_PRESLEEP_CYCLES = 0


def _call_inner(name: str, meta, fn):
    if _profile_sink is None:
        check(fn(), name)
        return
    if PROFILE_PRESLEEP_CYCLES > 0:
        torch.cuda._sleep(PROFILE_PRESLEEP_CYCLES)
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    check(fn(), name)
    end.record()
    _profile_sink.append((name, meta, start, end))


def _require_no_grad(*tensors):
    if torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors):
        raise NotImplementedError(
            "this operator has no backward here: the raw ops are forward-only, gradients are recorded by "
            "hydra_gnn_b200.train (every conv block of the four model classes). "
            "Call under torch.no_grad() or model.eval() with frozen parameters")


def _f32(t: Tensor, name: str) -> Tensor:
    if t.dtype != torch.float32:
        raise TypeError(f"{name}: expected float32, got {t.dtype} (fp64 mode of the reference is not supported)")
    if not t.is_cuda:
        raise RuntimeError(f"{name}: expected a CUDA tensor; hydra_gnn_b200 has no CPU path")
    return t


def _rows(t: Tensor, name: str) -> Tensor:
    """2-D, unit stride along features (an arbitrary row stride is the leading dimension)."""
    _f32(t, name)
    if t.dim() != 2:
        raise ValueError(f"{name}: expected a 2-D tensor")
    if t.size(1) > 0 and t.stride(1) != 1:
        t = t.contiguous()
    return t


def _ld(t: Tensor) -> int:
    """Leading dimension in floats (any value works for a single row; keep it vector friendly)."""
    return t.stride(0) if t.size(0) > 1 else max((t.size(1) + 3) // 4 * 4, 4)


# ------------------------------------------------------------------------------ workspace
_workspaces = {}


def _workspace(device, nbytes: int) -> Tensor:
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


# ------------------------------------------------------------------------------ K1 plan
class _Join:
    """A plan built on the side stream: the first consumer of any of its CSRs makes the consuming
    stream wait for it (so K1 overlaps whatever was launched in between -- the first projection)."""
    __slots__ = ("side", "tensors", "done")

    def __init__(self, side, tensors):
        self.side, self.tensors, self.done = side, tensors, False

    def wait(self):
        cur = torch.cuda.current_stream(self.tensors[0].device)
        if self.done is False:
            self.done = set()
        if cur.cuda_stream in self.done:  # every consuming stream joins once
            return
        cur.wait_stream(self.side)
        for t in self.tensors:  # allocated while `side` was current, consumed on `cur`
            t.record_stream(cur)
        self.done.add(cur.cuda_stream)


_side_streams: Dict[tuple, "torch.cuda.Stream"] = {}
PARALLEL_BRANCHES = True  # independent per-destination-type chains of a hetero layer on concurrent streams


def _side_stream(device, k: int = 0) -> "torch.cuda.Stream":
    """Side stream k of the device (0: the plan; 1..: branches).  High priority: their kernels are the
    short ones that fill the gaps of the long kernel on the main stream."""
    s = _side_streams.get((device.index, k))
    if s is None:
        s = _side_streams[(device.index, k)] = torch.cuda.Stream(device, priority=int(os.environ.get("HG_SIDE_PRIORITY", "-1")))
    return s


def _tensors_of(obj):
    if isinstance(obj, Tensor):
        return [obj]
    if isinstance(obj, (list, tuple)):
        return [t for o in obj for t in _tensors_of(o)]
    if isinstance(obj, dict):
        return [t for o in obj.values() for t in _tensors_of(o)]
    return []


def parallel(fns, device):
    """Runs the zero-argument callables as concurrent branches: the first on the current stream, the
    others on side streams forked from it and joined back (also valid under CUDA-graph capture).
    Each branch may only read tensors that exist before the call and returns the tensors it produced."""
    if len(fns) <= 1 or not PARALLEL_BRANCHES:
        return [f() for f in fns]
    cur = torch.cuda.current_stream(device)
    results = [None] * len(fns)
    sides = []
    for k in range(1, len(fns)):
        s = _side_stream(device, k)
        s.wait_stream(cur)
        with torch.cuda.stream(s):
            results[k] = fns[k]()
        sides.append(s)
    results[0] = fns[0]()
    for k, s in enumerate(sides, 1):
        cur.wait_stream(s)
        for t in _tensors_of(results[k]):  # allocated on the side stream, consumed on the current one
            t.record_stream(cur)
    return results


@dataclass
class CSR:
    """Destination-sorted adjacency of one relation (device int32)."""
    rowptr: Tensor
    col: Tensor
    perm: Optional[Tensor]
    inv_sqrt_deg: Optional[Tensor]
    n_src: int
    n_dst: int
    cap: int  # allocated entries; the live count is rowptr[-1] (kept on the device)
    self_loops: bool
    join: Optional[_Join] = None
    edge_index: Optional[Tensor] = None  # kept when perm was requested: the training side is built from it
    train: Optional["TrainCSR"] = None   # reversed CSR + position map (ensure_train_csr)
    tile_win: Optional[Tensor] = None    # int32 [ceil(n_dst / 32), 2]: source-row window of every destination tile

    def ready(self) -> "CSR":
        """Call before handing the arrays to a kernel on the current stream."""
        if self.join is not None:
            self.join.wait()
        return self


@dataclass
class CSRSpec:
    edge_index: Tensor  # int64 [2, E] (row 0 = source, row 1 = target)
    n_src: int
    n_dst: int
    self_loops: bool = False
    want_perm: bool = False
    want_norm: bool = False
    transpose: bool = False  # build the CSR of the reversed relation (rows = sources)
    want_window: bool = False  # also record the source-row window of every 32-row destination tile (hg_gat_fused)


def build_csr(specs: Sequence[CSRSpec], overlap: bool = False):
    """All relations of a batch in one ``hg_plan_build`` call (four launches in total).

    ``overlap=True`` launches the plan on a side stream forked from the current one; the current
    stream joins it at the first ``CSR.ready()`` (every op below calls it), so independent work
    launched in between runs concurrently with K1.

    Returns (list of CSR, status) -- ``status`` is a device int32[1] counting edges whose
    endpoints were outside [0, n) (skipped, never dereferenced); read it with
    ``int(status)`` only when a sync is acceptable.
    """
    if len(specs) == 0:
        return [], None
    device = specs[0].edge_index.device
    if not specs[0].edge_index.is_cuda:
        raise RuntimeError("edge_index must be a CUDA tensor; hydra_gnn_b200 has no CPU path")
    if overlap:
        cur, side = torch.cuda.current_stream(device), _side_stream(device)
        side.wait_stream(cur)  # the edge lists are ready on the current stream
        with torch.cuda.stream(side):
            out, status = build_csr(specs)
        tensors = [t for c in out for t in (c.rowptr, c.col, c.perm, c.inv_sqrt_deg, c.tile_win) if t is not None] + [status]
        join = _Join(side, tensors)
        for c in out:
            c.join = join
        return out, status
    h = lib()
    out: List[CSR] = []
    status = torch.empty(1, dtype=torch.int32, device=device)
    for lo in range(0, len(specs), _lib.HG_MAX_PLAN_RELATIONS):
        chunk = specs[lo: lo + _lib.HG_MAX_PLAN_RELATIONS]
        rels = (_lib.PlanRel * len(chunk))()
        keep = []
        for k, s in enumerate(chunk):
            ei = s.edge_index
            if ei.dtype != torch.int64 or ei.dim() != 2 or ei.size(0) != 2:
                raise TypeError("edge_index must be int64 [2, E]")
            if not ei.is_cuda:
                raise RuntimeError("edge_index must be a CUDA tensor; hydra_gnn_b200 has no CPU path")
            if not ei.is_contiguous():
                ei = ei.contiguous()
            E = ei.size(1)
            n_src, n_dst = (s.n_dst, s.n_src) if s.transpose else (s.n_src, s.n_dst)
            loops = min(n_src, n_dst) if s.self_loops else 0
            cap = E + loops
            rowptr = torch.empty(n_dst + 1, dtype=torch.int32, device=device)
            col = torch.empty(max(cap, 1), dtype=torch.int32, device=device)
            perm = torch.empty(max(cap, 1), dtype=torch.int32, device=device) if s.want_perm else None
            norm = torch.empty(max(n_dst, 1), dtype=torch.float32, device=device) if s.want_norm else None
            n_tiles = (n_dst + _lib.HG_TILE_ROWS - 1) // _lib.HG_TILE_ROWS
            win = torch.empty(max(n_tiles, 1), 2, dtype=torch.int32, device=device) if s.want_window else None
            r = rels[k]
            src_row, dst_row = (1, 0) if s.transpose else (0, 1)
            r.src = ei.data_ptr() + src_row * E * 8
            r.dst = ei.data_ptr() + dst_row * E * 8
            r.n_edges, r.n_src, r.n_dst, r.self_loops = E, n_src, n_dst, int(s.self_loops)
            r.rowptr, r.col, r.perm, r.inv_sqrt_deg = ptr(rowptr), ptr(col), ptr(perm), ptr(norm)
            r.tile_win = ptr(win)
            keep.append(ei)
            out.append(CSR(rowptr, col, perm, norm, n_src, n_dst, cap, s.self_loops,
                           edge_index=ei if (s.want_perm and not s.transpose) else None, tile_win=win))
        nbytes = h.hg_plan_workspace_bytes(rels, len(chunk))
        if nbytes == 0:
            raise RuntimeError("hg_plan_workspace_bytes rejected the plan (sizes out of int32 range?)")
        ws = _workspace(device, nbytes)
        st = status if lo == 0 else torch.empty(1, dtype=torch.int32, device=device)
        meta = dict(n_rel=len(chunk), edges=[int(s.edge_index.size(1)) for s in chunk],
                    n_dst=[int(r.n_dst) for r in rels], caps=[c.cap for c in out[lo:]])
        _call("hg_plan_build", meta,
              lambda: h.hg_plan_build(rels, len(chunk), ptr(st), ptr(ws), ws.numel(), stream_ptr()))
        _lib.launch_count += 4  # count, scan, fill, rank (the tile windows ride on the count and scan passes)
        if lo != 0:
            status += st
    return out, status


# ------------------------------------------------------------------------------ K2 / K5
def seg_reduce(x: Tensor, csr: CSR, op: str = "sum", *, src_scale: Tensor = None, dst_scale: Tensor = None,
               row_map: Tensor = None, x_self: Tensor = None, self_coef: Tensor = None, self_coef_add: float = 0.0,
               bias: Tensor = None, act: str = None, out: Tensor = None) -> Tensor:
    _require_no_grad(x, x_self, bias)
    h = lib()
    x = _rows(x, "x")
    F = x.size(1)
    csr.ready()
    n_out = csr.n_dst if row_map is None else row_map.numel()
    if out is None:
        out = torch.empty(n_out, F, dtype=torch.float32, device=x.device)
    a = _lib.SegArgs()
    a.x, a.ldx = ptr(x), _ld(x)
    a.rowptr, a.col = ptr(csr.rowptr), ptr(csr.col)
    a.src_scale, a.dst_scale = ptr(src_scale), ptr(dst_scale)
    if row_map is not None and row_map.dtype != torch.int32:
        raise TypeError("row_map must be int32")
    a.row_map = ptr(row_map)
    if x_self is not None:
        x_self = _rows(x_self, "x_self")
        if x_self.size(1) != F:
            raise ValueError("x_self width mismatch")
        a.x_self, a.ldself = ptr(x_self), _ld(x_self)
        a.self_coef_ptr = ptr(self_coef)
        a.self_coef_add = float(self_coef_add)
    a.bias = ptr(bias)
    a.out, a.ldout = ptr(out), _ld(out)
    a.n_out, a.F = n_out, F
    a.op, a.act = REDUCE[op], ACT[act]
    meta = dict(n_src=x.size(0), n_out=n_out, F=F, cap=csr.cap, scaled=src_scale is not None, self=x_self is not None)
    _call("hg_seg_reduce", meta, lambda: h.hg_seg_reduce(C.byref(a), stream_ptr()))
    _lib.launch_count += 1
    return out


# ------------------------------------------------------------------------------ K3 + K4
@dataclass
class GatRelation:
    xs: Tensor            # [n_src, H*C] view (may be a column slice of the projection output)
    a_src: Tensor         # [n_src, H] view
    a_dst: Tensor         # [n_dst, H] view
    csr: CSR
    bias: Optional[Tensor] = None
    a_edge: Optional[Tensor] = None  # [cap, H] contiguous, CSR order
    alpha_out: Optional[Tensor] = None


def window_applies(rels: Sequence["GatRelation"], n_dst: int) -> bool:
    """Whether the plan's per-tile source windows are handed to ``hg_gat_fused`` (which then stages them in shared
    memory).  Only where the windows are short, judged from host-known sizes: every relation carries one and has at most
    9 edges per destination row on average (the kernel's edge pool per 32-row tile), and the relations together have at
    most 1.5 source rows per destination row -- all relations of a tile share ONE 64-row window pool, and a tile of 32
    destinations spans about 32 x n_src / n_dst source rows of each.  Objects of a scene graph (objects + their rooms:
    1.08) and rooms <- rooms qualify; a tile of rooms spans hundreds of object rows, and the 2- and 4-relation
    destination types of an H-tree (measured: 239 vs 215 us and 780 vs 689 us) stay on the global-gather kernel."""
    return USE_TILE_WINDOW and n_dst > 0 and all(
        r.csr.tile_win is not None and r.csr.cap <= 9 * n_dst for r in rels) and \
        sum(r.csr.n_src for r in rels) <= 1.5 * n_dst


def gat_fused(rels: Sequence[GatRelation], n_dst: int, H: int, Cc: int, concat: bool, *, group: str = "sum",
              act: str = None, neg_slope: float = 0.2, out: Tensor = None, addend: Tensor = None,
              src_shared: bool = False, dropout_p: float = 0.0, seed: int = 0, seed_dev: Tensor = None) -> Tensor:
    """Fused GAT aggregation of every relation targeting one node type (``hg_gat_fused``).

    ``src_shared=True`` is the aggregate-first form: ``xs`` is ``[n_src, C]`` (one row shared by all
    heads) and the result is ``[n_dst, H*C]`` with ``out[i, h] = sum_e w_eh xs[col[e]]``.
    ``addend`` ``[n_dst, width]`` is added after the relation group, before the activation."""
    h = lib()
    if src_shared:
        concat = True
    if not 1 <= len(rels) <= _lib.HG_MAX_RELATIONS:
        raise ValueError("1..8 relations per destination type")
    device = rels[0].xs.device
    width = H * Cc if concat else Cc
    if out is None:
        out = torch.empty(n_dst, width, dtype=torch.float32, device=device)
    a = _lib.GatArgs()
    use_window = window_applies(rels, n_dst)
    for k, r in enumerate(rels):
        _require_no_grad(r.xs, r.a_src, r.a_dst, r.bias)
        for t, nm in ((r.xs, "xs"), (r.a_src, "a_src"), (r.a_dst, "a_dst")):
            _f32(t, nm)
            if t.dim() != 2 or (t.size(1) > 1 and t.stride(1) != 1):
                raise ValueError(f"{nm}: expected 2-D with unit feature stride")
        if r.xs.size(1) != (Cc if src_shared else H * Cc) or r.a_src.size(1) != H or r.a_dst.size(1) != H:
            raise ValueError("GAT operand widths do not match H, C")
        r.csr.ready()
        if r.csr.n_dst != n_dst:
            raise ValueError("relation CSR has a different destination count")
        g = a.rel[k]
        g.xs, g.ldxs = ptr(r.xs), _ld(r.xs)
        g.a_src, g.lda_src = ptr(r.a_src), _ld(r.a_src)
        g.a_dst, g.lda_dst = ptr(r.a_dst), _ld(r.a_dst)
        g.a_edge = ptr(r.a_edge)
        g.rowptr, g.col, g.bias = ptr(r.csr.rowptr), ptr(r.csr.col), ptr(r.bias)
        g.tile_win = ptr(r.csr.tile_win) if use_window else None
        g.n_src = int(r.xs.size(0))
        a.alpha_saved[k] = ptr(r.alpha_out)
    a.n_rel, a.n_dst, a.H, a.C = len(rels), n_dst, H, Cc
    a.concat, a.group, a.act, a.neg_slope = int(concat), GROUP[group], ACT[act], float(neg_slope)
    a.out, a.ldout = ptr(out), _ld(out)
    if addend is not None:
        _require_no_grad(addend)
        addend = _rows(addend, "addend")
        if addend.shape != (n_dst, width):
            raise ValueError("addend shape mismatch")
        a.addend, a.ldadd = ptr(addend), _ld(addend)
    a.src_shared = int(src_shared)
    a.dropout_p, a.dropout_seed = float(dropout_p), int(seed) & 0xFFFFFFFFFFFFFFFF
    if seed_dev is not None:
        if seed_dev.dtype != torch.int64 or seed_dev.numel() != 1 or not seed_dev.is_cuda:
            raise TypeError("seed_dev must be a CUDA int64 scalar")
        a.dropout_seed_dev = ptr(seed_dev)
    meta = dict(n_dst=n_dst, H=H, C=Cc, concat=bool(concat), width=width, src_width=Cc if src_shared else H * Cc,
                addend=addend is not None,
                rels=[dict(n_src=r.xs.size(0), cap=r.csr.cap, edge=r.a_edge is not None) for r in rels])
    _call("hg_gat_fused", meta, lambda: h.hg_gat_fused(C.byref(a), stream_ptr()))
    _lib.launch_count += 1
    return out


# ------------------------------------------------------------------------------ K6
def linear(segments: Sequence, N: int, *, bias: Tensor = None, scale: Tensor = None, shift: Tensor = None,
           act: str = None, precision: str = None, out: Tensor = None, pack_cache: dict = None) -> Tensor:
    """``y = act((sum_s x_s @ W_s^T + bias) * scale + shift)``; ``segments`` = [(x_s, W_s), ...].

    ``pack_cache``: a dict owned by whoever owns the weights (a layer module).  The tensor-core path
    reads a split + swizzled image of W; with a cache the image is built once per weight version
    (``hg_linear_pack_weights``) and every later call is a single kernel launch."""
    h = lib()
    if not 1 <= len(segments) <= _lib.HG_MAX_SEGMENTS:
        raise ValueError("1..8 K-segments")
    a = _lib.LinearArgs()
    M = segments[0][0].size(0)
    device = segments[0][0].device
    keep = []
    for k, (x, w) in enumerate(segments):
        _require_no_grad(x, w)
        x, w = _rows(x, "x"), _rows(w, "W")
        if x.size(0) != M or w.size(0) != N or x.size(1) != w.size(1):
            raise ValueError(f"segment {k}: shapes x{tuple(x.shape)} W{tuple(w.shape)} do not match M={M}, N={N}")
        s = a.seg[k]
        s.x, s.ldx, s.w, s.ldw, s.K = ptr(x), _ld(x), ptr(w), _ld(w), x.size(1)
        keep.append((x, w))
    _require_no_grad(bias, scale, shift)
    if out is None:
        out = torch.empty(M, N, dtype=torch.float32, device=device)
    a.n_seg, a.M, a.N = len(segments), M, N
    a.bias, a.scale, a.shift = ptr(bias), ptr(scale), ptr(shift)
    a.act = ACT[act]
    precision = precision or DEFAULT_PRECISION
    if precision == "auto":
        precision = "tf32x3" if M >= TC_MIN_ROWS else "fp32"
        if FP32_SMALL_ROWS and M < FP32_SMALL_ROWS and sum(x.size(1) for x, _ in keep) <= FP32_SMALL_K:
            precision = "fp32"  # A/B: short, shallow projections on the CUDA cores (no TMEM / barrier set-up)
    a.precision = PREC[precision]
    a.y, a.ldy = ptr(out), _ld(out)
    meta = dict(M=M, N=N, K=[x.size(1) for x, _ in keep], precision=precision)
    if M == 0 or N == 0:
        return out
    nbytes = h.hg_linear_workspace_bytes(C.byref(a))
    ws = None
    if nbytes and pack_cache is not None:
        key = (N, _param_generation, tuple((w.data_ptr(), w._version, w.size(1), _ld(w)) for _, w in keep))
        ws = pack_cache.get(key)
        if ws is None:
            pack_cache.clear()  # one live image per owner: older weight versions are dead
            ws = torch.empty(nbytes, dtype=torch.uint8, device=device)
            _call("hg_linear_pack_weights", dict(bytes=nbytes),
                  lambda: h.hg_linear_pack_weights(C.byref(a), ptr(ws), nbytes, stream_ptr()))
            _lib.launch_count += 1
            pack_cache[key] = ws
        a.w_prepacked = 1
        launches = 1
    else:
        ws = _workspace(device, nbytes) if nbytes else None
        launches = 2 if nbytes else 1
    _call("hg_linear", meta, lambda: h.hg_linear(C.byref(a), ptr(ws), nbytes, stream_ptr()))
    _lib.launch_count += launches
    return out


# ------------------------------------------------------------------------------ edge helpers
def edge_logits(edge_attr: Tensor, csr: CSR, v: Tensor, H: int, fill: Tensor = None) -> Tensor:
    """``a_edge[p, h] = <edge_attr[perm[p]], v[h]>`` in CSR order (added loops: ``<fill, v[h]>`` or 0)."""
    h = lib()
    csr.ready()
    if csr.perm is None:
        raise ValueError("edge_logits needs a CSR built with want_perm=True")
    edge_attr = _rows(edge_attr, "edge_attr")
    v = _rows(v, "v")
    D = edge_attr.size(1)
    out = torch.empty(max(csr.cap, 1), H, dtype=torch.float32, device=v.device)
    last = csr.rowptr[-1:]
    _call("hg_edge_logits", dict(cap=csr.cap, H=H, D=D),
          lambda: h.hg_edge_logits(ptr(edge_attr), _ld(edge_attr), D, ptr(csr.perm), last.data_ptr(), csr.cap,
                                   ptr(v), ptr(fill), H, ptr(out), stream_ptr()))
    _lib.launch_count += 1
    return out


def relative_pos(pos_src: Tensor, pos_dst: Tensor, edge_index: Tensor, status: Tensor = None) -> Tensor:
    """``edge_attr = pos_dst[edge_index[1]] - pos_src[edge_index[0]]`` (mp3d_dataset.py:316-319), bit-identical to
    the torch expression.  ``status``: optional CUDA int32[1] that receives the number of edges with an endpoint out of
    range (their rows are NaN; torch would raise) -- read it only where a sync is acceptable."""
    h = lib()
    pos_src, pos_dst = _f32(pos_src, "pos_src").contiguous(), _f32(pos_dst, "pos_dst").contiguous()
    if pos_src.dim() != 2 or pos_src.size(1) != 3 or pos_dst.dim() != 2 or pos_dst.size(1) != 3:
        raise ValueError("pos must be [n, 3]")
    if edge_index.dtype != torch.int64 or edge_index.dim() != 2 or edge_index.size(0) != 2 or not edge_index.is_cuda:
        raise TypeError("edge_index must be a CUDA int64 [2, E] tensor")
    ei = edge_index.contiguous()
    E = ei.size(1)
    out = torch.empty(E, 3, dtype=torch.float32, device=ei.device)
    _call("hg_relative_pos", dict(E=E),
          lambda: h.hg_relative_pos(ptr(pos_src), pos_src.size(0), ptr(pos_dst), pos_dst.size(0), ei.data_ptr(),
                                    ei.data_ptr() + 8 * E, E, ptr(out), ptr(status), stream_ptr()))
    _lib.launch_count += 1
    return out


def gather_rows(x: Tensor, idx: Tensor, act: str = None) -> Tensor:
    """``act(x[idx])`` with ``idx`` int32 (the models' ``x[room_mask, :]``)."""
    h = lib()
    x = _rows(x, "x")
    if idx.dtype != torch.int32:
        raise TypeError("idx must be int32")
    out = torch.empty(idx.numel(), x.size(1), dtype=torch.float32, device=x.device)
    _call("hg_gather_rows", dict(n=idx.numel(), F=x.size(1)),
          lambda: h.hg_gather_rows(ptr(x), _ld(x), ptr(idx), idx.numel(), x.size(1), ACT[act], ptr(out), _ld(out),
                                   stream_ptr()))
    _lib.launch_count += 1
    return out


# ------------------------------------------------------------------------------ collate (integer, bit-exact)
def collate_offsets_(index: Tensor, edge_ptr: Tensor, src_ptr: Tensor, dst_ptr: Tensor) -> Tensor:
    """In place: ``index[0, e] += src_ptr[g(e)]``, ``index[1, e] += dst_ptr[g(e)]`` where ``g(e)`` is the
    graph whose column range ``edge_ptr[g]:edge_ptr[g+1]`` holds ``e`` (all device int64)."""
    h = lib()
    if index.dtype != torch.int64 or index.dim() != 2 or index.size(0) != 2 or not index.is_contiguous():
        raise TypeError("index must be a contiguous int64 [2, E] tensor")
    for t in (edge_ptr, src_ptr, dst_ptr):
        if t.dtype != torch.int64 or not t.is_cuda or t.numel() != edge_ptr.numel():
            raise TypeError("ptr arrays must be CUDA int64 of the same length (n_graphs + 1)")
    if not index.is_cuda:
        raise RuntimeError("index must be a CUDA tensor; hydra_gnn_b200 has no CPU path")
    _call("hg_collate_offsets", dict(E=index.size(1)),
          lambda: h.hg_collate_offsets(ptr(index), index.size(1), ptr(edge_ptr), ptr(src_ptr), ptr(dst_ptr),
                                       edge_ptr.numel() - 1, stream_ptr()))
    _lib.launch_count += 1
    return index


def batch_vector(ptr_: Tensor, n_nodes: int) -> Tensor:
    """``batch[n] = g`` for ``ptr[g] <= n < ptr[g+1]`` (device int64), PyG's ``batch`` vector."""
    h = lib()
    if ptr_.dtype != torch.int64 or not ptr_.is_cuda:
        raise TypeError("ptr must be a CUDA int64 tensor")
    out = torch.empty(n_nodes, dtype=torch.int64, device=ptr_.device)
    _call("hg_batch_vector", dict(N=n_nodes),
          lambda: h.hg_batch_vector(ptr(out), n_nodes, ptr(ptr_), ptr_.numel() - 1, stream_ptr()))
    _lib.launch_count += 1
    return out


# ------------------------------------------------------------------------------ training step (configs[4])
@dataclass
class TrainCSR:
    """Source-major side of a relation for the backward: CSR of the reversed relation and, per entry, its
    position in the destination-major CSR (where alpha / dlogit live)."""
    rowptr_t: Tensor
    col_t: Tensor
    pos_t: Tensor


def ensure_train_csr(csrs: Sequence[CSR]) -> None:
    """Builds (one plan call for all of them) the reversed CSRs of the relations that lack one."""
    todo = [c for c in {id(c): c for c in csrs}.values() if c.train is None]
    if not todo:
        return
    h = lib()
    for c in todo:
        if c.perm is None or c.edge_index is None:
            raise RuntimeError("training needs a plan built with perm (build it with grad mode enabled)")
        c.ready()
    specs = [CSRSpec(c.edge_index, c.n_src, c.n_dst, c.self_loops, True, False, transpose=True) for c in todo]
    rev, _ = build_csr(specs)
    for c, t in zip(todo, rev):
        E = c.edge_index.size(1)
        n_loop = min(c.n_src, c.n_dst) if c.self_loops else 0
        scratch = torch.empty(max(E + n_loop, 1), dtype=torch.int32, device=c.col.device)
        pos_t = torch.empty(max(c.cap, 1), dtype=torch.int32, device=c.col.device)
        # rowptr[-1] (the live entry count) stays on the device
        _call("hg_csr_transpose_map", dict(cap=c.cap),
              lambda: h.hg_csr_transpose_map(ptr(c.perm), ptr(c.col), c.rowptr.data_ptr() + 4 * c.n_dst, c.cap,
                                             ptr(t.perm), ptr(t.col), t.rowptr.data_ptr() + 4 * t.n_dst, E, n_loop,
                                             ptr(scratch), ptr(pos_t), stream_ptr()))
        _lib.launch_count += 2
        c.train = TrainCSR(t.rowptr, t.col, pos_t)


def linear_wgrad(dy: Tensor, x: Tensor, out: Tensor = None, accumulate: bool = False) -> Tensor:
    """``dW[N, K] = dy[M, N]^T @ x[M, K]`` (``hg_linear_wgrad``)."""
    h = lib()
    dy, x = _rows(dy, "dy"), _rows(x, "x")
    M, N, K = dy.size(0), dy.size(1), x.size(1)
    if x.size(0) != M:
        raise ValueError("dy / x row counts differ")
    if out is None:
        out = torch.empty(N, K, dtype=torch.float32, device=x.device)
        accumulate = False
    meta = dict(M=M, N=N, K=K)
    _call("hg_linear_wgrad", meta,
          lambda: h.hg_linear_wgrad(ptr(dy), _ld(dy), ptr(x), _ld(x), M, N, K, ptr(out), _ld(out), int(accumulate),
                                    stream_ptr()))
    _lib.launch_count += 1 if accumulate else 2
    return out


def adam_step(param: Tensor, grad: Tensor, exp_avg: Tensor, exp_avg_sq: Tensor, step: int, lr: float, beta1: float,
              beta2: float, eps: float, weight_decay: float, grad_scale: float = 1.0, step_dev: Tensor = None) -> None:
    """``step_dev``: CUDA int32 scalar holding t (>= 1), read by the kernel instead of the host value ``step``."""
    h = lib()
    for t in (param, grad, exp_avg, exp_avg_sq):
        _f32(t, "adam operand")
        if not t.is_contiguous() or t.numel() != param.numel():
            raise ValueError("adam operands must be contiguous and of equal size")
    _call("hg_adam_step", dict(n=param.numel()),
          lambda: h.hg_adam_step(ptr(param), ptr(grad), ptr(exp_avg), ptr(exp_avg_sq), param.numel(), lr, beta1, beta2,
                                 eps, weight_decay, grad_scale, step, ptr(step_dev), stream_ptr()))
    _lib.launch_count += 1


@dataclass
class GatTrainRelation:
    """One relation of a training-form GAT group.  ``y_src`` / ``y_dst`` are the projection outputs of the source /
    destination node type; the operands are column ranges of them (same layout for the gradients)."""
    y_src: Tensor
    y_dst: Tensor
    xs_off: int
    as_off: int
    ad_off: int
    csr: CSR
    bias: Optional[Tensor] = None
    alpha: Optional[Tensor] = None
    a_edge: Optional[Tensor] = None  # [cap, H] contiguous, CSR order (GAT_edge): backward only, forward = gat_fused


def _gat_train_args(rels: Sequence[GatTrainRelation], n_dst: int, H: int, Cc: int, concat: bool, group: str, act: str,
                    neg_slope: float, dropout_p: float, seed: int, out: Tensor):
    if not 1 <= len(rels) <= _lib.HG_MAX_RELATIONS:
        raise ValueError("1..8 relations per destination type")
    a = _lib.GatTrainArgs()
    for k, r in enumerate(rels):
        for t in (r.y_src, r.y_dst):
            _f32(t, "y")
            if t.dim() != 2 or (t.size(1) > 1 and t.stride(1) != 1):
                raise ValueError("projection outputs must be 2-D with unit feature stride")
        r.csr.ready()
        if r.csr.n_dst != n_dst or r.y_dst.size(0) != n_dst or r.y_src.size(0) != r.csr.n_src:
            raise ValueError("relation shapes do not match the CSR")
        g = a.rel[k]
        es = r.y_src.element_size()
        g.xs, g.ldxs = r.y_src.data_ptr() + es * r.xs_off, _ld(r.y_src)
        g.a_src, g.lda_src = r.y_src.data_ptr() + es * r.as_off, _ld(r.y_src)
        g.a_dst, g.lda_dst = r.y_dst.data_ptr() + es * r.ad_off, _ld(r.y_dst)
        g.rowptr, g.col, g.bias, g.alpha = ptr(r.csr.rowptr), ptr(r.csr.col), ptr(r.bias), ptr(r.alpha)
        g.n_src = r.csr.n_src
    a.n_rel, a.n_dst, a.H, a.C = len(rels), n_dst, H, Cc
    a.concat, a.group, a.act, a.neg_slope = int(concat), GROUP[group], ACT[act], float(neg_slope)
    a.dropout_p, a.seed = float(dropout_p), int(seed) & 0xFFFFFFFFFFFFFFFF
    a.out, a.ldout = ptr(out), _ld(out)
    return a


def gat_train_fwd(rels: Sequence[GatTrainRelation], n_dst: int, H: int, Cc: int, concat: bool, *, group: str = "sum",
                  act: str = None, neg_slope: float = 0.2, dropout_p: float = 0.0, seed: int = 0) -> Tensor:
    """Training-form GAT aggregation (``hg_gat_train_fwd``); fills ``r.alpha`` ([cap, H]) of every relation."""
    h = lib()
    device = rels[0].y_src.device
    out = torch.empty(n_dst, H * Cc if concat else Cc, dtype=torch.float32, device=device)
    for r in rels:
        if r.alpha is None:
            r.alpha = torch.empty(max(r.csr.cap, 1), H, dtype=torch.float32, device=device)
    a = _gat_train_args(rels, n_dst, H, Cc, concat, group, act, neg_slope, dropout_p, seed, out)
    meta = dict(n_dst=n_dst, H=H, C=Cc, width=out.size(1), rels=[dict(n_src=r.csr.n_src, cap=r.csr.cap) for r in rels])
    _call("hg_gat_train_fwd", meta, lambda: h.hg_gat_train_fwd(C.byref(a), stream_ptr()))
    _lib.launch_count += 1
    return out


def gat_train_bwd(rels: Sequence[GatTrainRelation], dys: Sequence[tuple], n_dst: int, H: int, Cc: int, concat: bool,
                  out: Tensor, dout: Tensor, *, group: str = "sum", act: str = None, neg_slope: float = 0.2,
                  dropout_p: float = 0.0, seed: int = 0, dbias: Tensor = None) -> None:
    """Backward of :func:`gat_train_fwd`.  ``dys[k] = (dY_src, dY_dst)``: gradient buffers with the layout of the
    projection outputs; the relation's columns of them are overwritten.  Returns the per-relation ``dlogit`` buffers
    ([cap, H], CSR order): the gradient of the raw scores, i.e. of ``a_edge`` for a relation that carries one (rows past
    the live entry count are zero for those)."""
    h = lib()
    device = out.device
    ensure_train_csr([r.csr for r in rels])
    a = _gat_train_args(rels, n_dst, H, Cc, concat, group, act, neg_slope, dropout_p, seed, out)
    dout = _rows(dout, "dout")
    gbuf = torch.empty(n_dst, out.size(1), dtype=torch.float32, device=device)
    keep = []
    for k, (r, (dys_, dyd_)) in enumerate(zip(rels, dys)):
        g = a.rel[k]
        alloc = torch.zeros if r.a_edge is not None else torch.empty
        dl = alloc(max(r.csr.cap, 1), H, dtype=torch.float32, device=device)
        keep.append(dl)
        if r.a_edge is not None:
            if r.a_edge.dtype != torch.float32 or not r.a_edge.is_contiguous() or r.a_edge.shape != dl.shape:
                raise ValueError("a_edge must be contiguous float32 [cap, H]")
            g.a_edge = ptr(r.a_edge)
        t = r.csr.train
        g.dlogit, g.rowptr_t, g.col_t, g.pos_t = ptr(dl), ptr(t.rowptr_t), ptr(t.col_t), ptr(t.pos_t)
        g.dxs, g.lddxs = dys_.data_ptr() + 4 * r.xs_off, _ld(dys_)
        g.da_src, g.ldda_src = dys_.data_ptr() + 4 * r.as_off, _ld(dys_)
        g.da_dst, g.ldda_dst = dyd_.data_ptr() + 4 * r.ad_off, _ld(dyd_)
    a.dout, a.lddout = ptr(dout), _ld(dout)
    a.g, a.ldg = ptr(gbuf), _ld(gbuf)
    a.dbias = ptr(dbias)
    meta = dict(n_dst=n_dst, H=H, C=Cc, width=out.size(1), rels=[dict(n_src=r.csr.n_src, cap=r.csr.cap) for r in rels])
    _call("hg_gat_train_bwd", meta, lambda: h.hg_gat_train_bwd(C.byref(a), stream_ptr()))
    _lib.launch_count += 2
    return keep


def act_bwd(dy: Tensor, y: Tensor, act: str = None, want_g: bool = True, want_dbias: bool = False):
    """``g = dy * act'(y)`` (``y`` = the activation's output) and/or ``dbias = g.sum(0)`` (``hg_act_bwd``).
    Returns ``(g, dbias)``; with no activation ``g`` is ``dy`` itself."""
    h = lib()
    dy = _rows(dy, "dy")
    M, N = dy.shape
    code = ACT[act]
    g = dy
    if code != 0 and want_g:
        g = torch.empty(M, N, dtype=torch.float32, device=dy.device)
    dbias = torch.zeros(N, dtype=torch.float32, device=dy.device) if want_dbias else None
    if code != 0 or want_dbias:
        y = _rows(y, "y") if code != 0 else None
        _call("hg_act_bwd", dict(M=M, N=N),
              lambda: h.hg_act_bwd(ptr(dy), _ld(dy), ptr(y), _ld(y) if y is not None else 0, M, N, code,
                                   ptr(g) if (code != 0 and want_g) else None, _ld(g), ptr(dbias), stream_ptr()))
        _lib.launch_count += 1
    return g, dbias


def reversed_csr(csr: CSR) -> CSR:
    """The by-source CSR of a relation (rows = its sources, col = destination ids) as a CSR object for
    :func:`seg_reduce` -- the scatter direction of every aggregation backward."""
    ensure_train_csr([csr])
    t = csr.train
    return CSR(t.rowptr_t, t.col_t, None, None, csr.n_dst, csr.n_src, csr.cap, csr.self_loops)
