"""Restatement of the ``torch_geometric.utils`` functions on the reference's path.

ORACLE -- test infrastructure only (see ``oracle/__init__.py``).

Follows the published PyG 2.3.1 algorithms (third-party, absent from
``/root/reference``; call sites in the reference: mp3d_dataset.py:16,116,
neural_tree/construct.py:6,52 and, inside the PyG layers, every ``GATConv`` /
``GCNConv`` built by models/utils.py).  Convention everywhere:
``edge_index[0]`` = source j, ``edge_index[1]`` = target i.
"""
from __future__ import annotations

import torch
from torch import Tensor


def maybe_num_nodes(edge_index: Tensor, num_nodes=None) -> int:
    if num_nodes is not None:
        return int(num_nodes)
    return int(edge_index.max()) + 1 if edge_index.numel() > 0 else 0


def scatter(src: Tensor, index: Tensor, dim: int = 0, dim_size=None, reduce: str = "sum") -> Tensor:
    """``torch_geometric.utils.scatter`` for dim=0 (the only use on this path).

    sum / mean via ``index_add_``; max / min via ``scatter_reduce_(include_self=False)``
    with untouched rows left at 0 -- PyG's semantics for empty segments.
    """
    assert dim == 0
    if dim_size is None:
        dim_size = int(index.max()) + 1 if index.numel() > 0 else 0
    size = (dim_size,) + tuple(src.shape[1:])
    if reduce in ("sum", "add"):
        return src.new_zeros(size).index_add_(0, index, src)
    if reduce == "mean":
        count = src.new_zeros(dim_size).index_add_(0, index, src.new_ones(src.size(0)))
        count = count.clamp_(min=1)
        out = src.new_zeros(size).index_add_(0, index, src)
        return out / count.view((-1,) + (1,) * (src.dim() - 1))
    if reduce in ("max", "min"):
        idx = index.view((-1,) + (1,) * (src.dim() - 1)).expand_as(src)
        return src.new_zeros(size).scatter_reduce_(0, idx, src, reduce="a" + reduce, include_self=False)
    raise ValueError(reduce)


def softmax(src: Tensor, index: Tensor, ptr=None, num_nodes=None, dim: int = 0) -> Tensor:
    """Segment softmax: max-shift, exp, ``sum + 1e-16``, divide (PyG 2.3.1 utils/softmax.py)."""
    N = maybe_num_nodes(index, num_nodes)
    src_max = scatter(src.detach(), index, 0, dim_size=N, reduce="max")
    out = (src - src_max.index_select(0, index)).exp()
    out_sum = scatter(out, index, 0, dim_size=N, reduce="sum") + 1e-16
    return out / out_sum.index_select(0, index)


def remove_self_loops(edge_index: Tensor, edge_attr=None):
    mask = edge_index[0] != edge_index[1]
    edge_index = edge_index[:, mask]
    if edge_attr is None:
        return edge_index, None
    return edge_index, edge_attr[mask]


def add_self_loops(edge_index: Tensor, edge_attr=None, fill_value=None, num_nodes=None):
    """Append ``(k, k), k = 0..N-1`` AFTER the real edges (PyG 2.3.1 utils/loop.py).

    ``fill_value``: float | Tensor | 'mean' etc. (per-target reduce of incoming edge_attr).
    """
    N = maybe_num_nodes(edge_index, num_nodes)
    loop_index = torch.arange(0, N, dtype=torch.long, device=edge_index.device)
    loop_index = loop_index.unsqueeze(0).repeat(2, 1)
    if edge_attr is not None:
        if fill_value is None:
            loop_attr = edge_attr.new_full((N,) + tuple(edge_attr.shape[1:]), 1.0)
        elif isinstance(fill_value, (int, float)):
            loop_attr = edge_attr.new_full((N,) + tuple(edge_attr.shape[1:]), fill_value)
        elif isinstance(fill_value, Tensor):
            loop_attr = fill_value.to(edge_attr.device, edge_attr.dtype)
            if edge_attr.dim() != loop_attr.dim():
                loop_attr = loop_attr.unsqueeze(0)
            sizes = [N] + [1] * (loop_attr.dim() - 1)
            loop_attr = loop_attr.repeat(*sizes)
        elif isinstance(fill_value, str):
            loop_attr = scatter(edge_attr, edge_index[1], 0, dim_size=N, reduce=fill_value)
        else:
            raise AttributeError("No valid 'fill_value' provided")
        edge_attr = torch.cat([edge_attr, loop_attr], dim=0)
    edge_index = torch.cat([edge_index, loop_index], dim=1)
    return edge_index, edge_attr


def add_remaining_self_loops(edge_index: Tensor, edge_attr=None, fill_value=1.0, num_nodes=None):
    """Drop existing loops, append one loop per node that inherits the weight of an
    existing loop when there was one (PyG 2.3.1 utils/loop.py)."""
    N = maybe_num_nodes(edge_index, num_nodes)
    mask = edge_index[0] != edge_index[1]
    loop_index = torch.arange(0, N, dtype=torch.long, device=edge_index.device)
    loop_index = loop_index.unsqueeze(0).repeat(2, 1)
    if edge_attr is not None:
        loop_attr = edge_attr.new_full((N,) + tuple(edge_attr.shape[1:]), fill_value)
        inv_mask = ~mask
        loop_attr[edge_index[0][inv_mask]] = edge_attr[inv_mask]
        edge_attr = torch.cat([edge_attr[mask], loop_attr], dim=0)
    edge_index = torch.cat([edge_index[:, mask], loop_index], dim=1)
    return edge_index, edge_attr


def degree(index: Tensor, num_nodes=None, dtype=None) -> Tensor:
    N = maybe_num_nodes(index, num_nodes)
    out = torch.zeros((N,), dtype=dtype or torch.get_default_dtype(), device=index.device)
    return out.index_add_(0, index, out.new_ones(index.size(0)))


def is_undirected(edge_index: Tensor, edge_attr=None, num_nodes=None) -> bool:
    """True iff the multiset of (j, i) pairs equals the multiset of (i, j) pairs."""
    N = maybe_num_nodes(edge_index, num_nodes)
    a = (edge_index[0] * N + edge_index[1]).sort().values
    b = (edge_index[1] * N + edge_index[0]).sort().values
    return bool(torch.equal(a, b))


def to_networkx(data, node_attrs=None, edge_attrs=None, to_undirected=False, remove_self_loops=False):
    """``Data`` -> ``nx.DiGraph``; nodes 0..N-1 inserted in order, then edges in
    ``edge_index`` column order; tensor attributes become Python lists / scalars
    (PyG 2.3.1 utils/convert.py).  Used by neural_tree/construct.py:52."""
    import networkx as nx

    G = nx.Graph() if to_undirected else nx.DiGraph()
    G.add_nodes_from(range(data.num_nodes))
    node_attrs = node_attrs or []
    values = {}
    for key in node_attrs:
        v = getattr(data, key)
        values[key] = v.tolist() if isinstance(v, Tensor) else v
    for u, v in data.edge_index.t().tolist():
        if to_undirected and v > u:
            continue
        if remove_self_loops and u == v:
            continue
        G.add_edge(u, v)
    for key in node_attrs:
        for i, feat_dict in G.nodes(data=True):
            feat_dict.update({key: values[key][i]})
    return G
