"""Deterministic synthetic "MP3D-shaped" scene graphs.

Shape statistics follow the one real Hydra MP3D graph bundled with the reference
(tests/test_data/x8F5xyUWy9e_0_gt_partial_dsg_1447.json: 5 rooms, ~12 objects per room,
sparse room graph) and the feature layout of preprocess_dsgs.py:392-421
(x = pos(3) | bbox size(3) | word2vec(300)); relation names / order are ``EDGE_TYPES``.

Every random draw is a pure function ``hash(seed, graph_id, stream, index)`` (a splitmix64
counter RNG), so graph ``g`` is identical whatever batch, rank or shard it is generated in:
sharding 1 M graphs over 8 GPUs needs no coordination.  Generation is vectorised numpy on
the host; the model inputs are the tensors of the returned ``HeteroData`` batch.
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch

from .data import HeteroData
from .schema import EDGE_TYPES

SEED = 20240229
MAX_ROOMS, MAX_OBJ = 12, 40
N_OBJECT_LABELS, N_ROOM_LABELS, W2V_DIM = 28, 26, 300

_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def _mix(z: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser (uint64 arrays, wrap-around arithmetic)."""
    z = (z + np.uint64(0x9E3779B97F4A7C15)) & _M
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M
    return z ^ (z >> np.uint64(31))


def _u01(seed: int, gid: np.ndarray, stream: int, idx: np.ndarray) -> np.ndarray:
    """Uniform [0, 1) float64 from (seed, graph id, stream, index), broadcast over gid/idx."""
    with np.errstate(over="ignore"):
        k = _mix(np.uint64(seed) ^ _mix(gid.astype(np.uint64)))
        k = _mix(k ^ (np.uint64(stream) * np.uint64(0xD1B54A32D192ED03)))
        k = _mix(k ^ _mix(idx.astype(np.uint64)))
    return (k >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def _poisson_cdf(lam: float, kmax: int) -> np.ndarray:
    """Cumulative Poisson(lam) table over 0..kmax in float64 (the device generator is handed the same array)."""
    k = np.arange(kmax + 1)
    logp = -lam + k * np.log(lam) - np.cumsum(np.concatenate([[0.0], np.log(np.arange(1, kmax + 1))]))
    return np.cumsum(np.exp(logp))


def _poisson_icdf(u: np.ndarray, lam: float, kmax: int) -> np.ndarray:
    return np.searchsorted(_poisson_cdf(lam, kmax), u, side="right")


def _normal(seed, gid, stream, idx):
    u1 = _u01(seed, gid, stream, 2 * idx)
    u2 = _u01(seed, gid, stream, 2 * idx + 1)
    return np.sqrt(-2.0 * np.log(1.0 - u1)) * np.cos(2.0 * np.pi * u2)


def word2vec_table(seed: int = SEED) -> np.ndarray:
    """Fixed 28 x 300 table standing in for the word2vec rows of the 28 object labels."""
    lab = np.repeat(np.arange(N_OBJECT_LABELS), W2V_DIM)
    d = np.tile(np.arange(W2V_DIM), N_OBJECT_LABELS)
    return (0.15 * _normal(seed, np.full_like(lab, 0xFFFFFF), 99, lab * W2V_DIM + d)).reshape(
        N_OBJECT_LABELS, W2V_DIM).astype(np.float32)


def synth_mp3d_batch(graph_ids, seed: int = SEED, with_word2vec: bool = True, relative_pos: bool = False,
                     device=None, pin: bool = False) -> HeteroData:
    """Disjoint-union batch of the graphs ``graph_ids`` (any iterable of ints, any order).

    Node order: objects grouped by room in room order, graphs in the given order.  Edges are
    emitted relation by relation in ``EDGE_TYPES`` order, in generation order (NOT sorted by
    destination).  ``relative_pos=True`` gives the GAT_edge variant (mp3d_dataset.py:298-319):
    the 3 position columns are dropped and ``edge_attr = pos_dst - pos_src`` is attached.
    """
    gid = np.asarray(list(graph_ids), dtype=np.int64)
    G = gid.size
    # ---- rooms per graph, objects per room
    R = np.clip(_poisson_icdf(_u01(seed, gid, 1, np.zeros(G, np.int64)), 5.0, 64), 1, MAX_ROOMS)  # [G]
    room_graph = np.repeat(np.arange(G), R)                       # graph slot of each room
    room_start = np.concatenate([[0], np.cumsum(R)])              # [G+1]
    room_local = np.arange(R.sum()) - room_start[room_graph]      # room index inside its graph
    NR = room_graph.size
    nobj = np.clip(_poisson_icdf(_u01(seed, gid[room_graph], 2, room_local), 12.0, 128), 1, MAX_OBJ)  # [NR]
    obj_room = np.repeat(np.arange(NR), nobj)                     # global room of each object
    obj_start = np.concatenate([[0], np.cumsum(nobj)])            # [NR+1]
    obj_local = np.arange(nobj.sum()) - obj_start[obj_room]       # object index inside its room
    NO = obj_room.size
    obj_graph = room_graph[obj_room]
    obj_in_graph = np.arange(NO) - obj_start[room_start[obj_graph]]  # object index inside its graph

    # ---- rooms <-> rooms: spanning link to a random earlier room w.p. 0.8 + other pairs w.p. 0.05
    a_idx, b_idx = np.triu_indices(MAX_ROOMS, k=1)               # a < b
    pg = np.repeat(np.arange(G), a_idx.size)
    pa, pb = np.tile(a_idx, G), np.tile(b_idx, G)
    live = pb < R[pg]
    pg, pa, pb = pg[live], pa[live], pb[live]
    parent = np.floor(_u01(seed, gid[pg], 3, pb) * pb).astype(np.int64)       # in [0, b)
    tree = (_u01(seed, gid[pg], 4, pb) < 0.8) & (pa == parent)
    extra = _u01(seed, gid[pg], 5, pa * MAX_ROOMS + pb) < 0.05
    sel = tree | extra
    ra, rb = room_start[pg[sel]] + pa[sel], room_start[pg[sel]] + pb[sel]
    rr = np.stack([np.stack([ra, rb], 1).reshape(-1), np.stack([rb, ra], 1).reshape(-1)])  # a->b, b->a interleaved

    # ---- objects <-> objects inside a room: each pair w.p. min(1, 4 / (n - 1)), both directions
    oa_idx, ob_idx = np.triu_indices(MAX_OBJ, k=1)
    chunks = []
    step = max(1, (1 << 22) // oa_idx.size)                       # bound the temporary arrays
    for lo in range(0, NR, step):
        rooms = np.arange(lo, min(NR, lo + step))
        qr = np.repeat(rooms, oa_idx.size)
        qa, qb = np.tile(oa_idx, rooms.size), np.tile(ob_idx, rooms.size)
        ok = qb < nobj[qr]
        qr, qa, qb = qr[ok], qa[ok], qb[ok]
        p = np.minimum(1.0, 4.0 / np.maximum(nobj[qr] - 1, 1))
        hit = _u01(seed, gid[room_graph[qr]], 6, (room_local[qr] * MAX_OBJ + qa) * MAX_OBJ + qb) < p
        ga, gb = obj_start[qr[hit]] + qa[hit], obj_start[qr[hit]] + qb[hit]
        chunks.append(np.stack([np.stack([ga, gb], 1).reshape(-1), np.stack([gb, ga], 1).reshape(-1)]))
    oo = np.concatenate(chunks, axis=1) if chunks else np.zeros((2, 0), np.int64)

    # ---- objects <-> rooms
    o_ids = np.arange(NO)
    o2r = np.stack([o_ids, obj_room])
    r2o = np.stack([obj_room, o_ids])

    # ---- features
    rl3 = (room_local[:, None] * 3 + np.arange(3)[None, :])
    room_pos = -10.0 + 20.0 * _u01(seed, gid[room_graph][:, None], 7, rl3)
    room_size = 2.0 + 6.0 * _u01(seed, gid[room_graph][:, None], 8, rl3)
    ol3 = (obj_in_graph[:, None] * 3 + np.arange(3)[None, :])
    obj_pos = room_pos[obj_room] + 1.5 * _normal(seed, gid[obj_graph][:, None], 9, ol3)
    obj_size = 0.1 + 1.9 * _u01(seed, gid[obj_graph][:, None], 10, ol3)
    obj_label = np.floor(_u01(seed, gid[obj_graph], 11, obj_in_graph) * N_OBJECT_LABELS).astype(np.int64)
    room_label = np.floor(_u01(seed, gid[room_graph], 12, room_local) * N_ROOM_LABELS).astype(np.int64)

    x_room = np.concatenate([room_pos, room_size], 1).astype(np.float32)
    parts = [obj_pos, obj_size]
    x_obj = np.concatenate(parts, 1).astype(np.float32)
    if with_word2vec:
        x_obj = np.concatenate([x_obj, word2vec_table(seed)[obj_label]], 1)

    data = HeteroData()
    data["objects"].x = torch.from_numpy(np.ascontiguousarray(x_obj))
    data["objects"].pos = torch.from_numpy(obj_pos.astype(np.float32))
    data["objects"].y = torch.from_numpy(obj_label)
    data["objects"].batch = torch.from_numpy(obj_graph.astype(np.int64))
    data["rooms"].x = torch.from_numpy(np.ascontiguousarray(x_room))
    data["rooms"].pos = torch.from_numpy(room_pos.astype(np.float32))
    data["rooms"].y = torch.from_numpy(room_label)
    data["rooms"].batch = torch.from_numpy(room_graph.astype(np.int64))
    for et, ei in zip(EDGE_TYPES, (oo, rr, o2r, r2o)):
        data[et].edge_index = torch.from_numpy(np.ascontiguousarray(ei.astype(np.int64)))
    data.num_graphs = G
    from .convert import compute_relative_pos

    def rel_pos():  # GAT_edge variant (mp3d_dataset.py:298-319); the sliced x made contiguous for the projections
        compute_relative_pos(data)
        for nt in ("objects", "rooms"):
            data[nt].x = data[nt].x.contiguous()

    if relative_pos and device is None:
        rel_pos()
    if pin:
        data.pin_memory()
    if device is not None:
        data.to(device)
        if relative_pos:  # the batch lives on the device: hg_relative_pos, bit-identical to the host expression
            rel_pos()
    return data


class DeviceSynth:
    """``synth_mp3d_batch`` generated on the GPU by ``hg_synth_mp3d_counts`` / ``hg_synth_mp3d_fill`` (csrc/synth.cu):
    same counter hash, same tables -- counts, edge lists, labels and batch vectors are bit-identical to the host
    generator, features too except that the Box-Muller object positions may differ in the last float32 bit where
    libm and CUDA round a double differently.

    ``sizes(first, n)`` launches the counting pass of a range of graph ids and returns its totals (ONE host sync;
    call it for every micro-batch of a job up front), ``batch(first, n)`` then fills a ``HeteroData`` without any
    further synchronisation."""

    def __init__(self, device, seed: int = SEED, with_word2vec: bool = True):
        self.device, self.seed, self.with_word2vec = torch.device(device), seed, with_word2vec
        self.cdf_rooms = torch.from_numpy(_poisson_cdf(5.0, 64)).to(self.device)
        self.cdf_objects = torch.from_numpy(_poisson_cdf(12.0, 128)).to(self.device)
        self.w2v = torch.from_numpy(word2vec_table(seed)).to(self.device) if with_word2vec else None
        self._prefix = {}

    def _args(self, first: int, n: int, prefix: torch.Tensor):
        from . import _lib

        a = _lib.SynthArgs()
        a.graph_id0, a.n_graphs, a.seed = first, n, self.seed
        a.cdf_rooms, a.cdf_objects, a.prefix = self.cdf_rooms.data_ptr(), self.cdf_objects.data_ptr(), prefix.data_ptr()
        return a

    def counts_async(self, first: int, n: int) -> torch.Tensor:
        """Launch the counting pass; returns the device tensor ``[n + 1, 4]`` of exclusive prefix sums (+ totals)."""
        import ctypes as C

        from . import _lib
        from .ops import _call

        h = _lib.lib()
        prefix = torch.empty(n + 1, 4, dtype=torch.int64, device=self.device)
        a = self._args(first, n, prefix)
        with torch.cuda.device(self.device):
            _call("hg_synth_mp3d_counts", dict(n_graphs=n), lambda: h.hg_synth_mp3d_counts(C.byref(a), _lib.stream_ptr()))
        _lib.launch_count += 2
        self._prefix[(first, n)] = prefix
        return prefix

    def sizes(self, first: int, n: int):
        """Totals ``(rooms, objects, room pairs, object pairs)`` of graphs ``first .. first + n - 1`` (host sync)."""
        prefix = self._prefix.get((first, n))
        if prefix is None:
            prefix = self.counts_async(first, n)
        return tuple(int(v) for v in prefix[n].tolist())

    def batch(self, first: int, n: int, sizes=None, labels: bool = True) -> HeteroData:
        import ctypes as C

        from . import _lib
        from .ops import _call

        h = _lib.lib()
        if sizes is None:
            sizes = self.sizes(first, n)
        prefix = self._prefix.pop((first, n), None)
        if prefix is None:
            prefix = self.counts_async(first, n)
            self._prefix.pop((first, n), None)
        NR, NO, Prr, Poo = sizes
        dev, i64 = self.device, torch.int64
        width = 6 + (W2V_DIM if self.with_word2vec else 0)
        data = HeteroData()
        x_room = torch.empty(NR, 6, dtype=torch.float32, device=dev)
        x_obj = torch.empty(NO, width, dtype=torch.float32, device=dev)
        rr, oo = torch.empty(2, 2 * Prr, dtype=i64, device=dev), torch.empty(2, 2 * Poo, dtype=i64, device=dev)
        o2r, r2o = torch.empty(2, NO, dtype=i64, device=dev), torch.empty(2, NO, dtype=i64, device=dev)
        status = torch.empty(1, dtype=torch.int32, device=dev)
        a = self._args(first, n, prefix)
        a.n_rooms, a.n_objects, a.n_rr_pairs, a.n_oo_pairs = NR, NO, Prr, Poo
        a.word2vec = self.w2v.data_ptr() if self.with_word2vec else None
        a.x_rooms, a.x_objects, a.ld_objects = x_room.data_ptr(), x_obj.data_ptr(), width
        a.rooms_to_rooms, a.objects_to_objects = rr.data_ptr(), oo.data_ptr()
        a.objects_to_rooms, a.rooms_to_objects, a.status = o2r.data_ptr(), r2o.data_ptr(), status.data_ptr()
        if labels:
            data["rooms"].y = torch.empty(NR, dtype=i64, device=dev)
            data["objects"].y = torch.empty(NO, dtype=i64, device=dev)
            data["rooms"].batch = torch.empty(NR, dtype=i64, device=dev)
            data["objects"].batch = torch.empty(NO, dtype=i64, device=dev)
            a.room_label, a.object_label = data["rooms"].y.data_ptr(), data["objects"].y.data_ptr()
            a.room_batch, a.object_batch = data["rooms"].batch.data_ptr(), data["objects"].batch.data_ptr()
        with torch.cuda.device(dev):
            _call("hg_synth_mp3d_fill", dict(n_graphs=n, objects=NO), lambda: h.hg_synth_mp3d_fill(C.byref(a), _lib.stream_ptr()))
        _lib.launch_count += 1
        data["objects"].x, data["rooms"].x = x_obj, x_room
        data["objects"].pos, data["rooms"].pos = x_obj[:, :3], x_room[:, :3]
        for et, ei in zip(EDGE_TYPES, (oo, rr, o2r, r2o)):
            data[et].edge_index = ei
        data.num_graphs = n
        data.synth_status = status  # device int32[1]: 1 when `sizes` disagree with the counting pass
        return data


def random_state_dict_(model: torch.nn.Module, seed: int = 0, bias_std: float = 0.1) -> torch.nn.Module:
    """Re-draw every parameter (biases included -- PyG zero-inits them, which would hide bias
    bugs) from a seeded generator; used so oracle and product share non-trivial weights."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        seen = set()
        for name, p in model.named_parameters():
            if id(p) in seen:
                continue
            seen.add(id(p))
            if p.dim() >= 2:
                fan = p.size(-1) + p.size(-2)
                p.copy_((torch.rand(p.shape, generator=g) * 2 - 1) * (6.0 / fan) ** 0.5)
            elif name.endswith("eps"):
                p.copy_(torch.rand(p.shape, generator=g) * 0.2)
            else:
                p.copy_(torch.randn(p.shape, generator=g) * bias_std)
        for name, b in model.named_buffers():
            if name.endswith("running_mean"):
                b.copy_(torch.randn(b.shape, generator=g) * 0.1)
            elif name.endswith("running_var"):
                b.copy_(torch.rand(b.shape, generator=g) + 0.5)
    return model
