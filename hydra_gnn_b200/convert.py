"""Heterogeneous -> homogeneous graph conversions (integer index work, bit-exact with the reference).

Host-side mirror of ``heterogeneous_data_to_homogeneous`` (mp3d_dataset.py:29-69) and
``heterogeneous_htree_to_homogeneous`` (mp3d_dataset.py:72-119), which the reference applies to
every graph when ``network.type == homogeneous`` (scripts/train_mp3d.py:134-135;
bin/room_classification_server:259-260).  The underlying ``HeteroData.to_homogeneous`` rule
(SURVEY.md section 9.7): nodes concatenated in node-type order, edges concatenated in edge-type
order with the per-type node offsets added, a ``node_type`` vector attached.

These are data-format functions (they run wherever the tensors live, like the reference's);
they are not part of the timed forward.
"""
from __future__ import annotations

from typing import List, Tuple

import torch

from .data import Data, HeteroData


def to_homogeneous(data: HeteroData, add_edge_type: bool = False) -> Data:
    """PyG ``HeteroData.to_homogeneous(add_edge_type=...)`` for stores whose attributes agree."""
    node_types = data.node_types
    offsets, cum = {}, 0
    for t in node_types:
        offsets[t] = cum
        cum += int(data[t].num_nodes)
    out = Data()
    eis, ets = [], []
    for k, et in enumerate(data.edge_types):
        if "edge_index" not in data[et]:
            continue
        ei = data[et].edge_index.long().clone()
        ei[0] += offsets[et[0]]
        ei[1] += offsets[et[2]]
        eis.append(ei)
        ets.append(torch.full((ei.size(1),), k, dtype=torch.long, device=ei.device))
    if eis:
        out.edge_index = torch.cat(eis, dim=1)
    # node attributes present in EVERY node store with matching trailing shape
    first = data[node_types[0]]
    for key, v in first.items():
        if not torch.is_tensor(v) or v.dim() < 1 or key in ("batch", "ptr"):
            continue
        vals = []
        for t in node_types:
            w = data[t].get(key)
            if not torch.is_tensor(w) or w.shape[1:] != v.shape[1:]:
                vals = None
                break
            vals.append(w)
        if vals is not None:
            out[key] = torch.cat(vals, dim=0)
    edge_stores = [data[et] for et in data.edge_types if "edge_index" in data[et]]
    if edge_stores and all("edge_attr" in s for s in edge_stores):
        out.edge_attr = torch.cat([s.edge_attr for s in edge_stores], dim=0)
    sizes = torch.tensor([int(data[t].num_nodes) for t in node_types], dtype=torch.long)
    out.node_type = torch.arange(len(node_types)).repeat_interleave(sizes)
    if "x" in out:
        out.node_type = out.node_type.to(out.x.device)
    if add_edge_type and eis:
        out.edge_type = torch.cat(ets)
    return out


def heterogeneous_data_to_homogeneous(data: HeteroData) -> Tuple[Data, List[str]]:
    """mp3d_dataset.py:29-69 (``removed_edge_type=None`` branch): zero-pad every node type's features
    to the widest one, then ``to_homogeneous``.  Returns ``(data, node_types)``; the caller adds
    ``room_mask = node_type == node_types.index("rooms")`` (mp3d_dataset.py:346-348)."""
    width = max(int(data[t].num_node_features) for t in data.x_dict)
    for t in data.x_dict:
        x = data[t].x
        if x.size(0) == 0:
            data[t].x = torch.empty((0, width), dtype=x.dtype, device=x.device)
        else:
            data[t].x = torch.nn.functional.pad(x, (0, width - x.size(1), 0, 0), mode="constant", value=0)
    node_types = list(data.x_dict.keys())
    return to_homogeneous(data, add_edge_type=False), node_types


def dsg_to_homogeneous(data: HeteroData) -> Data:
    """``Hydra_mp3d_data.to_homogeneous`` (mp3d_dataset.py:339-348): conversion + ``room_mask``."""
    out, node_types = heterogeneous_data_to_homogeneous(data)
    out.room_mask = out.node_type == node_types.index("rooms")
    return out


def heterogeneous_htree_to_homogeneous(data: HeteroData) -> Data:
    """mp3d_dataset.py:72-119: flat augmented H-tree with ``init_edge_index`` (source is a virtual
    node), ``pool_edge_index`` (target is a virtual node), the remaining tree edges in
    ``edge_index``, and ``room_mask`` / ``object_mask`` marking the virtual (classification) nodes."""
    assert "object_virtual" in data.x_dict and "room_virtual" in data.x_dict
    if "y" in data["room_virtual"]:
        y = data["room_virtual"].y
        for t in data.x_dict:
            if "y" not in data[t]:
                data[t].y = -torch.ones(int(data[t].num_nodes), dtype=y.dtype, device=y.device)
    out, node_types = heterogeneous_data_to_homogeneous(data)
    ov, rv = node_types.index("object_virtual"), node_types.index("room_virtual")
    src_t = out.node_type[out.edge_index[0]]
    dst_t = out.node_type[out.edge_index[1]]
    init = (src_t == ov) | (src_t == rv)
    pool = (dst_t == ov) | (dst_t == rv)
    out.init_edge_index = out.edge_index[:, init]
    out.pool_edge_index = out.edge_index[:, pool]
    tree = ~(init | pool)
    if "edge_attr" in out:
        out.edge_attr = out.edge_attr[tree, :]
    out.edge_index = out.edge_index[:, tree]
    out.room_mask = out.node_type == rv
    out.object_mask = out.node_type == ov
    return out


def batched_htree_to_homogeneous(batch: HeteroData) -> Data:
    """``Batch.from_data_list([heterogeneous_htree_to_homogeneous(t) for t in trees])`` computed from
    ``batch = Batch.from_data_list(trees)`` (or ``htree.generate_htree_batch``) in one pass of tensor ops
    instead of a Python loop over graphs (mp3d_dataset.py:72-119 applied per graph, then PyG's collate):
    graph g's nodes are its node types in store order, its edges the relations in store order, split into
    init / pool / tree by the endpoint types.  Bit-identical to the per-graph route (tests/test_host.py)."""
    node_types = [t for t in batch.node_types if "x" in batch[t]]
    assert "object_virtual" in node_types and "room_virtual" in node_types
    dev = batch[node_types[0]].x.device
    if dev.type == "cuda" and USE_SPLIT_KERNEL:
        out = _batched_htree_to_homogeneous_device(batch, node_types)
        if out is not None:
            return out
    ptrs = {}
    for t in node_types:
        st = batch[t]
        if "ptr" in st:
            ptrs[t] = st.ptr.long().to(dev)
        else:
            ptrs[t] = torch.tensor([0, int(st.num_nodes)], dtype=torch.long, device=dev)
    G = int(ptrs[node_types[0]].numel()) - 1
    type_off, cum = {}, torch.zeros(G, dtype=torch.long, device=dev)
    for t in node_types:
        type_off[t] = cum
        cum = cum + (ptrs[t][1:] - ptrs[t][:-1])
    ptr = torch.zeros(G + 1, dtype=torch.long, device=dev)
    ptr[1:] = cum.cumsum(0)
    N = int(ptr[-1])
    # homogeneous index of every node of every type
    graph_of, place = {}, {}
    for t in node_types:
        n = int(ptrs[t][-1])
        g = torch.repeat_interleave(torch.arange(G, device=dev), ptrs[t][1:] - ptrs[t][:-1], output_size=n)
        graph_of[t] = g
        place[t] = ptr[:-1][g] + type_off[t][g] + (torch.arange(n, device=dev) - ptrs[t][g])
    out = Data()
    # ---- edges: per graph, relations in store order (a stable sort by (graph, relation))
    rels = [et for et in batch.edge_types if "edge_index" in batch[et]]
    K = len(batch.edge_types)
    virtual = ("object_virtual", "room_virtual")
    src, dst, keys, init, pool = [], [], [], [], []
    for k, et in enumerate(batch.edge_types):
        if et not in rels:
            continue
        ei = batch[et].edge_index.long()
        src.append(place[et[0]][ei[0]])
        dst.append(place[et[2]][ei[1]])
        keys.append(graph_of[et[0]][ei[0]] * K + k)
        init.append(torch.full((ei.size(1),), et[0] in virtual, dtype=torch.bool, device=dev))
        pool.append(torch.full((ei.size(1),), et[2] in virtual, dtype=torch.bool, device=dev))
    tree = None
    if rels:
        order = torch.sort(torch.cat(keys), stable=True).indices
        edges = torch.stack([torch.cat(src), torch.cat(dst)])[:, order]
        init, pool = torch.cat(init)[order], torch.cat(pool)[order]
        tree = ~(init | pool)
        out.edge_index = edges[:, tree]
    # ---- node attributes present in every store with one trailing shape (x zero-padded to the widest)
    width = max(int(batch[t].x.size(1)) for t in node_types)
    first = batch[node_types[0]]
    names = [k for k, v in first.items() if torch.is_tensor(v) and v.dim() >= 1 and k not in ("batch", "ptr")]
    if "y" in batch["room_virtual"] and "y" not in names:
        names.append("y")
    for key in names:
        vals = []
        for t in node_types:
            w = batch[t].get(key)
            if key == "y" and w is None and "y" in batch["room_virtual"]:
                y = batch["room_virtual"].y
                w = -torch.ones(int(ptrs[t][-1]), dtype=y.dtype, device=dev)
            if not torch.is_tensor(w):
                vals = None
                break
            vals.append(w)
        if vals is None:
            continue
        trail = (width,) if key == "x" else tuple(vals[0].shape[1:])
        if key != "x" and any(tuple(w.shape[1:]) != trail for w in vals):
            continue
        buf = torch.zeros((N,) + trail, dtype=vals[0].dtype, device=dev)
        for t, w in zip(node_types, vals):
            if key == "x":
                buf[place[t], : w.size(1)] = w
            else:
                buf[place[t]] = w
        out[key] = buf
    if rels and all("edge_attr" in batch[et] for et in rels):
        out.edge_attr = torch.cat([batch[et].edge_attr for et in rels], dim=0)[order][tree, :]
    node_type = torch.empty(N, dtype=torch.long, device=dev)
    for i, t in enumerate(node_types):
        node_type[place[t]] = i
    out.node_type = node_type
    if rels:
        out.init_edge_index = edges[:, init]
        out.pool_edge_index = edges[:, pool]
    out.room_mask = node_type == node_types.index("room_virtual")
    out.object_mask = node_type == node_types.index("object_virtual")
    out.batch = torch.repeat_interleave(torch.arange(G, device=dev), ptr[1:] - ptr[:-1], output_size=N)
    out.ptr = ptr
    out.num_graphs = G
    return out


last_split_status = None  # status word of the most recent hg_htree_split call (device int32[1])
USE_SPLIT_KERNEL = True  # device-resident batches: the integer half runs in hg_htree_split (False: the tensor-op pass)


def _batched_htree_to_homogeneous_device(batch: HeteroData, node_types: List[str]):
    """``batched_htree_to_homogeneous`` for a CUDA-resident batch: node placement, ``node_type`` / ``batch`` / ``ptr`` /
    masks and the tree / init / pool edge lists come from ONE ``hg_htree_split`` call (six launches, no host
    synchronisation: every output size is host-known); only the feature rows are moved with tensor indexing.
    Returns None for the shapes the kernel does not cover (edge attributes, a virtual -> virtual relation)."""
    import ctypes as C

    from . import _lib
    from .ops import _call, _workspace

    rels = [et for et in batch.edge_types if "edge_index" in batch[et]]
    virtual = ("object_virtual", "room_virtual")
    if len(node_types) > _lib.HG_MAX_HTREE_TYPES or len(rels) > _lib.HG_MAX_PLAN_RELATIONS:
        return None
    if any("edge_attr" in batch[et] for et in rels) or any(et[0] in virtual and et[2] in virtual for et in rels):
        return None
    if any(et[0] not in node_types or et[2] not in node_types for et in rels):
        return None
    h = _lib.lib()
    dev = batch[node_types[0]].x.device
    a = _lib.HtreeSplitArgs()
    keep, place, N = [], {}, 0
    G = None
    for i, t in enumerate(node_types):
        st = batch[t]
        n = int(st.x.size(0))
        ptr_t = st.ptr.long().to(dev).contiguous() if "ptr" in st else torch.tensor([0, n], dtype=torch.long, device=dev)
        if G is None:
            G = int(ptr_t.numel()) - 1
        elif int(ptr_t.numel()) - 1 != G:
            return None
        place[t] = torch.empty(n, dtype=torch.long, device=dev)
        keep.append(ptr_t)
        a.type[i].ptr, a.type[i].n_nodes, a.type[i].place = ptr_t.data_ptr(), n, place[t].data_ptr()
        a.type[i].is_room_virtual, a.type[i].is_object_virtual = int(t == "room_virtual"), int(t == "object_virtual")
        N += n
    sizes = [0, 0, 0]
    for k, et in enumerate(rels):
        ei = batch[et].edge_index
        if ei.dtype != torch.int64 or not ei.is_contiguous():
            ei = ei.long().contiguous()
        keep.append(ei)
        E = int(ei.size(1))
        a.rel[k].src, a.rel[k].dst, a.rel[k].n_edges = ei.data_ptr(), ei.data_ptr() + 8 * E, E
        a.rel[k].src_type, a.rel[k].dst_type = node_types.index(et[0]), node_types.index(et[2])
        sizes[1 if et[0] in virtual else (2 if et[2] in virtual else 0)] += E
    a.n_types, a.n_rel, a.n_graphs = len(node_types), len(rels), G
    out = Data()
    out.ptr = torch.empty(G + 1, dtype=torch.long, device=dev)
    out.node_type = torch.empty(N, dtype=torch.long, device=dev)
    out.batch = torch.empty(N, dtype=torch.long, device=dev)
    out.room_mask = torch.empty(N, dtype=torch.bool, device=dev)
    out.object_mask = torch.empty(N, dtype=torch.bool, device=dev)
    edges = [torch.empty(2, n, dtype=torch.long, device=dev) for n in sizes]
    status = torch.empty(1, dtype=torch.int32, device=dev)
    a.ptr, a.node_type, a.batch = out.ptr.data_ptr(), out.node_type.data_ptr(), out.batch.data_ptr()
    a.room_mask, a.object_mask, a.status = out.room_mask.data_ptr(), out.object_mask.data_ptr(), status.data_ptr()
    for c in range(3):
        a.edges[c] = edges[c].data_ptr()
    nbytes = h.hg_htree_split_workspace_bytes(C.byref(a))
    if nbytes == 0:
        return None
    ws = _workspace(dev, nbytes)
    rc = [0]

    def run():
        rc[0] = h.hg_htree_split(C.byref(a), ws.data_ptr(), ws.numel(), _lib.stream_ptr())
        return 0 if rc[0] == -6 else rc[0]  # HG_EUNSUPPORTED: the caller takes the tensor-op pass

    _call("hg_htree_split", dict(n_graphs=G, nodes=N, edges=sizes), run)
    if rc[0] == -6:
        return None
    _lib.launch_count += 6
    global last_split_status
    last_split_status = status  # device int32[1]: edges out of graph order / range (0 for collated batches)
    if rels:
        out.edge_index, out.init_edge_index, out.pool_edge_index = edges
    # ---- node attributes present in every store with one trailing shape (x zero-padded to the widest)
    width = max(int(batch[t].x.size(1)) for t in node_types)
    first = batch[node_types[0]]
    names = [k for k, v in first.items() if torch.is_tensor(v) and v.dim() >= 1 and k not in ("batch", "ptr")]
    if "y" in batch["room_virtual"] and "y" not in names:
        names.append("y")
    for key in names:
        vals = []
        for t in node_types:
            w = batch[t].get(key)
            if key == "y" and w is None and "y" in batch["room_virtual"]:
                y = batch["room_virtual"].y
                w = -torch.ones(int(batch[t].x.size(0)), dtype=y.dtype, device=dev)
            if not torch.is_tensor(w):
                vals = None
                break
            vals.append(w)
        if vals is None:
            continue
        trail = (width,) if key == "x" else tuple(vals[0].shape[1:])
        if key != "x" and any(tuple(w.shape[1:]) != trail for w in vals):
            continue
        buf = torch.zeros((N,) + trail, dtype=vals[0].dtype, device=dev)
        for t, w in zip(node_types, vals):
            if key == "x":
                buf[place[t], : w.size(1)] = w
            else:
                buf[place[t]] = w
        out[key] = buf
    out.num_graphs = G
    return out


def compute_relative_pos(data: HeteroData) -> HeteroData:
    """``Hydra_mp3d_data.compute_relative_pos`` (mp3d_dataset.py:298-319), in place: the three position columns are
    dropped from every node type's ``x`` and every relation gets ``edge_attr = pos_dst[edge_index[1]] -
    pos_src[edge_index[0]]``.  Device-resident data (the batches the models consume) goes through ``hg_relative_pos``
    -- one launch per relation, bit-identical to the indexing expression; host tensors (dataset preparation, as in the
    reference) are indexed in place where they live."""
    if data.edge_attr_dict:
        raise Warning("Cannot compute relative pos as edge_attr -- edge_attr is not empty.")  # mp3d_dataset.py:304-307
    for t in data.x_dict:
        data[t].x = data[t].x[:, 3:]
    for et, ei in data.edge_index_dict.items():
        ps, pd = data[et[0]].pos, data[et[2]].pos
        if ei.is_cuda:
            from . import ops
            data[et].edge_attr = ops.relative_pos(ps, pd, ei)
        else:
            data[et].edge_attr = pd[ei[1]] - ps[ei[0]]
    return data
