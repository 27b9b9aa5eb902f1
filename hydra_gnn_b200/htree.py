"""H-tree construction: host-side mirror of ``hydra_gnn.neural_tree.construct`` on the native builder.

Reference entry points (neural_tree/construct.py):

    generate_htree(dsg_torch)            :241-310   HeteroData -> networkx H-tree
    add_virtual_nodes_to_htree(htree_nx) :313-371   + virtual nodes, pooling and init edges
    nx_htree_to_torch(htree_aug)         :374-483   -> HeteroData over HTREE_NODE_TYPES + HTREE_VIRTUAL_NODE_TYPES

Here the three steps are one call, :func:`generate_htree` (one graph) / :func:`generate_htree_list` (many graphs,
built in parallel by ``libhydra_htree.so``, include/hydra_htree.h), returning the same HeteroData
``nx_htree_to_torch(add_virtual_nodes_to_htree(generate_htree(d)))`` returns: same node numbering, same edge
order, same features.  There is no networkx fallback: a missing library raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Sequence

import numpy as np
import torch

from .data import HeteroData
from .schema import (HTREE_EDGE_TYPES, HTREE_INIT_EDGE_TYPES, HTREE_NODE_TYPES, HTREE_POOL_EDGE_TYPES,
                     HTREE_VIRTUAL_NODE_TYPES)

NODE_TYPE_NAMES = HTREE_NODE_TYPES + HTREE_VIRTUAL_NODE_TYPES
EDGE_TYPE_NAMES = HTREE_EDGE_TYPES + HTREE_POOL_EDGE_TYPES + HTREE_INIT_EDGE_TYPES

_LIB = None
_vp, _i64, _i32 = C.c_void_p, C.c_int64, C.c_int32
_SYMBOLS = {
    "hg_htree_last_error": (C.c_char_p, []),
    "hg_htree_build": (C.c_int, [_i32, _vp, _vp, _vp, _vp, _i32, _vp]),
    "hg_htree_sizes": (C.c_int, [_vp, _vp, _vp, _vp]),
    "hg_htree_export": (C.c_int, [_vp] * 11),
    "hg_htree_free": (None, [_vp]),
    "hg_pyset_trace": (_i64, [_vp, _vp, _vp, _i64, _vp, _i64]),
    "hg_pytuple_hash": (_i64, [_vp, _i64]),
}


def library_path() -> str:
    return os.environ.get("HYDRA_HTREE_SO", os.path.join(os.path.dirname(os.path.abspath(__file__)), "libhydra_htree.so"))


def load_shared_object():
    """dlopen libhydra_htree.so and bind every symbol of include/hydra_htree.h (raises if one is missing)."""
    global _LIB
    if _LIB is None:
        path = library_path()
        if not os.path.exists(path):
            raise RuntimeError(f"{path} not built: run `python __graft_entry__.py` (make -C hydra_gnn_b200/csrc)")
        lib = C.CDLL(path)
        for name, (res, args) in _SYMBOLS.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _LIB = lib
    return _LIB


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def build_raw(graphs: Sequence, n_threads: int = 0) -> dict:
    """graphs: sequence of ``(is_room uint8 [N], edge_index int64 [2, E])`` in to_homogeneous() order.
    Returns the arrays of ``hg_htree_export`` (numpy, concatenated over the batch)."""
    lib = load_shared_object()
    G = len(graphs)
    node_ptr = np.zeros(G + 1, np.int64)
    edge_ptr = np.zeros(G + 1, np.int64)
    keep, eptrs = [], (C.c_void_p * max(G, 1))()
    rooms = []
    for g, (is_room, ei) in enumerate(graphs):
        is_room = np.ascontiguousarray(np.asarray(is_room, dtype=np.uint8))
        ei = np.ascontiguousarray(np.asarray(ei, dtype=np.int64).reshape(2, -1))
        node_ptr[g + 1] = node_ptr[g] + is_room.shape[0]
        edge_ptr[g + 1] = edge_ptr[g] + ei.shape[1]
        rooms.append(is_room)
        keep.append(ei)
        eptrs[g] = ei.ctypes.data
    is_room_all = np.concatenate(rooms) if rooms else np.zeros(0, np.uint8)
    handle = C.c_void_p()
    rc = lib.hg_htree_build(G, _ptr(node_ptr), _ptr(is_room_all), _ptr(edge_ptr), C.cast(eptrs, C.c_void_p), n_threads,
                            C.byref(handle))
    if rc != 0:
        raise RuntimeError(f"hg_htree_build failed ({rc}): {lib.hg_htree_last_error().decode()}")
    try:
        n, e, l = _i64(), _i64(), _i64()
        lib.hg_htree_sizes(handle, C.byref(n), C.byref(e), C.byref(l))
        out = dict(node_ptr=np.zeros(G + 1, np.int64), node_type=np.zeros(n.value, np.uint8),
                   clique_has=np.zeros(n.value, np.int32), node_order=np.zeros(n.value, np.int32),
                   edge_ptr=np.zeros(G + 1, np.int64), edge_src=np.zeros(e.value, np.int32),
                   edge_dst=np.zeros(e.value, np.int32), edge_type=np.zeros(e.value, np.uint8),
                   list_ptr=np.zeros(n.value + 1, np.int64), list_items=np.zeros(l.value, np.int32))
        rc = lib.hg_htree_export(handle, *[_ptr(out[k]) for k in ("node_ptr", "node_type", "clique_has", "node_order",
                                                                  "edge_ptr", "edge_src", "edge_dst", "edge_type",
                                                                  "list_ptr", "list_items")])
        if rc != 0:
            raise RuntimeError(f"hg_htree_export failed ({rc})")
    finally:
        lib.hg_htree_free(handle)
    return out


def signature(raw: dict, g: int) -> str:
    """``repr([(n, d["node_type"], d["clique_has"]) for n, d in aug.nodes.items()]) + repr(list(aug.edges))`` of
    tree ``g`` -- the string SURVEY.md section 8c hashes (SHA-256) over the reference's networkx output."""
    n0, n1 = int(raw["node_ptr"][g]), int(raw["node_ptr"][g + 1])
    e0, e1 = int(raw["edge_ptr"][g]), int(raw["edge_ptr"][g + 1])
    nodes = []
    for key in raw["node_order"][n0:n1].tolist():
        k = n0 + key
        t = int(raw["node_type"][k])
        if t in (2, 3):
            has = raw["list_items"][int(raw["list_ptr"][k]):int(raw["list_ptr"][k + 1])].tolist()
        else:
            has = int(raw["clique_has"][k])
        nodes.append((key, NODE_TYPE_NAMES[t], has))
    edges = list(zip(raw["edge_src"][e0:e1].tolist(), raw["edge_dst"][e0:e1].tolist()))
    return repr(nodes) + repr(edges)


# ------------------------------------------------------------------------------ HeteroData in, HeteroData out
_ROOM_TYPE, _OBJECT_TYPE = "rooms", "objects"  # construct.py:57 strips the final "s" to name the H-tree leaf types


def _graph_slices(data: HeteroData):
    """Per node type ``ptr`` [G+1] of a (possibly batched) room-object HeteroData."""
    ptrs, G = {}, None
    for nt in data.node_types:
        st = data[nt]
        if "ptr" in st:
            p = st.ptr.cpu().numpy().astype(np.int64)
        elif "batch" in st:
            b = st.batch.cpu().numpy()
            if b.size and np.any(np.diff(b) < 0):
                raise ValueError(f"batch vector of '{nt}' is not sorted")
            n_g = int(getattr(data, "num_graphs", 0) or (int(b.max()) + 1 if b.size else 1))
            p = np.concatenate([[0], np.cumsum(np.bincount(b, minlength=n_g))]).astype(np.int64)
        else:
            p = np.array([0, st.num_nodes], np.int64)
        ptrs[nt] = p
        G = len(p) - 1 if G is None else G
        if len(p) - 1 != G:
            raise ValueError("node types disagree on the number of graphs")
    return ptrs, G


def _native_inputs(data: HeteroData):
    """to_homogeneous() per graph (construct.py:46): node types in store order, relations in store order, columns
    in their given order.  Returns (graph tuples for build_raw, per-type ptr, per-type offset inside each graph)."""
    if set(data.node_types) != {_ROOM_TYPE, _OBJECT_TYPE}:
        raise ValueError(f"expected node types 'objects' and 'rooms', got {data.node_types}")
    ptrs, G = _graph_slices(data)
    counts = {nt: np.diff(ptrs[nt]) for nt in data.node_types}
    type_off, cum = {}, np.zeros(G, np.int64)
    for nt in data.node_types:
        type_off[nt] = cum.copy()
        cum = cum + counts[nt]
    node_ptr = np.concatenate([[0], np.cumsum(cum)]).astype(np.int64)
    is_room = np.zeros(int(node_ptr[-1]), np.uint8)
    gr = np.repeat(np.arange(G), counts[_ROOM_TYPE])
    is_room[node_ptr[gr] + type_off[_ROOM_TYPE][gr] + (np.arange(gr.size) - ptrs[_ROOM_TYPE][gr])] = 1
    srcs, dsts, graphs, rels = [], [], [], []
    for r, et in enumerate(data.edge_types):
        ei = data[et].edge_index.cpu().numpy().astype(np.int64)
        s_t, d_t = et[0], et[2]
        g = np.searchsorted(ptrs[s_t], ei[0], side="right") - 1
        if ei.shape[1] and np.any(g != np.searchsorted(ptrs[d_t], ei[1], side="right") - 1):
            raise ValueError(f"relation {et} has an edge between two graphs")
        srcs.append(ei[0] - ptrs[s_t][g] + type_off[s_t][g])
        dsts.append(ei[1] - ptrs[d_t][g] + type_off[d_t][g])
        graphs.append(g)
        rels.append(np.full(g.size, r, np.int64))
    if srcs:
        src, dst, g, rel = map(np.concatenate, (srcs, dsts, graphs, rels))
        order = np.argsort(g * len(data.edge_types) + rel, kind="stable")
        src, dst, g = src[order], dst[order], g[order]
    else:
        src = dst = g = np.zeros(0, np.int64)
    edge_ptr = np.concatenate([[0], np.cumsum(np.bincount(g, minlength=G))]).astype(np.int64)
    tuples = [(is_room[node_ptr[k]:node_ptr[k + 1]], np.stack([src[edge_ptr[k]:edge_ptr[k + 1]],
                                                                dst[edge_ptr[k]:edge_ptr[k + 1]]])) for k in range(G)]
    return tuples, ptrs, type_off, G


def _leaf_attr(data: HeteroData, nt: str, key: str):
    st = data[nt]
    if key == "pos":
        return st.pos if "pos" in st else st.x[:, :3]
    if key == "label":  # the reference's datasets carry `label`; the synthetic generator names it `y`
        return st.label if "label" in st else (st.y if "y" in st else None)
    return st[key] if key in st else None


def generate_htree_batch(data: HeteroData, clique_dim: int | None = None, n_threads: int = 0,
                         return_raw: bool = False):
    """Augmented H-trees of every graph of a (batched) room-object HeteroData, as ONE batched HeteroData: what
    ``Batch.from_data_list([nx_htree_to_torch(add_virtual_nodes_to_htree(generate_htree(g))) for g in graphs])``
    gives with the reference's functions (construct.py:241-483), followed by the truncation of the zero semantic
    columns of the clique features (mp3d_dataset.py:447-455): clique features are ``[pos, 0...]`` of width
    ``clique_dim`` (default: the room feature width).  Leaf and virtual nodes carry the source node's
    ``x`` / ``pos`` / ``label``; cliques carry label = clique_has = -1 (construct.py:428-430)."""
    tuples, ptrs, type_off, G = _native_inputs(data)
    raw = build_raw(tuples, n_threads=n_threads)
    node_ptr, ntype, has = raw["node_ptr"], raw["node_type"].astype(np.int64), raw["clique_has"].astype(np.int64)
    n_total = int(node_ptr[-1])
    graph_of = np.repeat(np.arange(G), np.diff(node_ptr))
    # per-type index of every H-tree node = its rank among the nodes of its type (to_heterogeneous + collate)
    rank = np.zeros(n_total, np.int64)
    type_nodes = {}
    for t, name in enumerate(NODE_TYPE_NAMES):
        idx = np.flatnonzero(ntype == t)
        rank[idx] = np.arange(idx.size)
        type_nodes[name] = idx
    obj_first = type_off[_OBJECT_TYPE]  # offset of the object block inside each graph's homogeneous numbering
    room_first = type_off[_ROOM_TYPE]
    room_dim = data[_ROOM_TYPE].x.size(1)
    clique_dim = room_dim if clique_dim is None else clique_dim
    out = HeteroData()
    src_rows = {}
    for name, src_type, first in (("object", _OBJECT_TYPE, obj_first), ("room", _ROOM_TYPE, room_first),
                                  ("object_virtual", _OBJECT_TYPE, obj_first), ("room_virtual", _ROOM_TYPE, room_first)):
        idx = type_nodes[name]
        g = graph_of[idx]
        rows = torch.from_numpy(ptrs[src_type][g] + has[idx] - first[g])
        src_rows[name] = rows
        st = out[name]
        st.x = data[src_type].x.index_select(0, rows)
        st.pos = _leaf_attr(data, src_type, "pos").index_select(0, rows)
        lab = _leaf_attr(data, src_type, "label")
        if lab is not None:
            st.label = lab.index_select(0, rows).to(torch.int32)
        st.clique_has = torch.from_numpy(has[idx].astype(np.int32))
        if "node_ids" in data[src_type]:
            st.node_ids = data[src_type].node_ids.index_select(0, rows)
    # edges: keys -> per-type ranks, split by relation in list(edges) order
    e_graph = np.repeat(np.arange(G), np.diff(raw["edge_ptr"]))
    e_src = node_ptr[e_graph] + raw["edge_src"]
    e_dst = node_ptr[e_graph] + raw["edge_dst"]
    e_type = raw["edge_type"]
    for r, et in enumerate(EDGE_TYPE_NAMES):
        sel = np.flatnonzero(e_type == r)
        out[et].edge_index = torch.from_numpy(np.stack([rank[e_src[sel]], rank[e_dst[sel]]]))
    # clique position = mean position of the room_virtual predecessors (construct.py:414-424), float64 mean -> f32
    rv_pos = out["room_virtual"].pos.double()
    for name, rel in (("object-room", ("room_virtual", "rv_to_or", "object-room")),
                      ("room-room", ("room_virtual", "rv_to_rr", "room-room"))):
        n = type_nodes[name].size
        ei = out[rel].edge_index
        acc = torch.zeros(n, 3, dtype=torch.float64).index_add_(0, ei[1], rv_pos.index_select(0, ei[0]))
        cnt = torch.zeros(n, dtype=torch.float64).index_add_(0, ei[1], torch.ones(ei.size(1), dtype=torch.float64))
        pos = torch.where(cnt[:, None] > 0, acc / cnt.clamp(min=1)[:, None], torch.zeros_like(acc)).float()
        st = out[name]
        st.x = torch.cat([pos, torch.zeros(n, max(clique_dim - 3, 0))], dim=1)[:, :clique_dim]
        st.pos = pos
        st.label = torch.full((n,), -1, dtype=torch.int32)
        st.clique_has = torch.full((n,), -1, dtype=torch.int32)
    for name in NODE_TYPE_NAMES:
        g = graph_of[type_nodes[name]]
        out[name].batch = torch.from_numpy(g)
        out[name].ptr = torch.from_numpy(np.concatenate([[0], np.cumsum(np.bincount(g, minlength=G))]).astype(np.int64))
    out.num_graphs = G
    return (out, raw) if return_raw else out


def generate_htree(data: HeteroData, clique_dim: int | None = None) -> HeteroData:
    """One room-object graph -> its augmented H-tree (the three reference calls of the module docstring)."""
    out = generate_htree_batch(data, clique_dim=clique_dim, n_threads=1)
    for name in NODE_TYPE_NAMES:
        del out[name]["batch"], out[name]["ptr"]
    return out
