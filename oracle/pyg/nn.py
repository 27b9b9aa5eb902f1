"""Restatement of the ``torch_geometric.nn`` operators the reference calls.

ORACLE -- test infrastructure only (see ``oracle/__init__.py``).

Third-party dependency restated: ``torch_geometric`` (PyG), unpinned in the
reference (setup.cfg:14); this file follows the published algorithms of
**PyG 2.3.1** (newest version README.md:53-56 lists as tested).  Reference call
sites: ``SAGEConv`` models/utils.py:14; ``GCNConv`` :16; ``GINConv`` :18;
``GATConv`` :49,63,76 and heterogeneous_neural_tree_network.py:95,
homogeneous_neural_tree_network.py:63; ``HeteroConv`` models/utils.py:100,132,
heterogeneous_neural_tree_network.py:103; ``MessagePassing`` (as ``LeafPool``)
heterogeneous_neural_tree_network.py:18-31; ``BatchNorm`` homogeneous_network.py:9,97.

Every operator is gather (``index_select``) -> elementwise -> scatter
(``index_add_`` / ``scatter_reduce_``), like PyG's own CPU path, so that timing it
is a fair stand-in for "the reference PyG CPU path".
"""
from __future__ import annotations

import inspect
import math
from collections import defaultdict
from typing import Dict, Optional, Tuple, Union

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch import Tensor

from .utils import (
    add_remaining_self_loops,
    add_self_loops as _add_self_loops,
    remove_self_loops as _remove_self_loops,
    scatter,
    softmax,
)


def glorot(t: Tensor):
    if t is not None:
        stdv = math.sqrt(6.0 / (t.size(-2) + t.size(-1)))
        t.data.uniform_(-stdv, stdv)


class Linear(nn.Module):
    """``torch_geometric.nn.dense.linear.Linear``: weight ``[out, in]``; glorot or
    torch-default (kaiming-uniform a=sqrt(5)) initialisation."""

    def __init__(self, in_channels, out_channels, bias=True, weight_initializer=None, bias_initializer=None):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.weight_initializer, self.bias_initializer = weight_initializer, bias_initializer
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels))
        if bias:
            self.bias = nn.Parameter(torch.empty(out_channels))
        else:
            self.register_parameter("bias", None)
        self.reset_parameters()

    def reset_parameters(self):
        if self.weight_initializer == "glorot":
            glorot(self.weight)
        else:
            nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            if self.bias_initializer == "zeros":
                nn.init.zeros_(self.bias)
            else:
                bound = 1.0 / math.sqrt(self.in_channels) if self.in_channels > 0 else 0.0
                nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, x):
        return F.linear(x, self.weight, self.bias)


class MessagePassing(nn.Module):
    """Minimal ``MessagePassing``: ``propagate`` = gather ``*_j`` / ``*_i`` arguments of
    ``message`` -> ``message`` -> scatter with ``aggr`` over targets -> ``update``.

    A single-Tensor argument ``x`` implies ``size = (x.size(0), x.size(0))`` -- the
    output has as many rows as ``x`` (this is what makes ``LeafPool`` return
    leaf-count rows, heterogeneous_neural_tree_network.py:181-185).
    """

    def __init__(self, aggr="add", flow="source_to_target", node_dim=0, **kwargs):
        super().__init__()
        self.aggr = aggr
        assert flow == "source_to_target" and node_dim == 0
        self._msg_params = None

    def _params(self):
        if self._msg_params is None:
            self._msg_params = [p for p in inspect.signature(self.message).parameters]
        return self._msg_params

    def propagate(self, edge_index: Tensor, size=None, **kwargs):
        size = list(size) if size is not None else [None, None]
        j, i = edge_index[0], edge_index[1]
        names = self._params()
        # size inference from the arguments `message` gathers (PyG `_collect`/`_set_size`)
        for name in names:
            if name[-2:] not in ("_j", "_i"):
                continue
            data = kwargs.get(name[:-2], None)
            if isinstance(data, (tuple, list)):
                for d in (0, 1):
                    if isinstance(data[d], Tensor) and size[d] is None:
                        size[d] = data[d].size(0)
            elif isinstance(data, Tensor):
                for d in (0, 1):
                    if size[d] is None:
                        size[d] = data.size(0)
        if size[1] is None:
            size[1] = size[0]
        msg_kwargs = {}
        for name in names:
            if name[-2:] in ("_j", "_i"):
                data = kwargs.get(name[:-2], None)
                idx, d = (j, 0) if name[-1] == "j" else (i, 1)
                if isinstance(data, (tuple, list)):
                    data = data[d]
                msg_kwargs[name] = None if data is None else data.index_select(0, idx)
            elif name == "index":
                msg_kwargs[name] = i
            elif name == "size_i":
                msg_kwargs[name] = size[1]
            elif name == "ptr":
                msg_kwargs[name] = None
            else:
                msg_kwargs[name] = kwargs.get(name, None)
        out = self.message(**msg_kwargs)
        reduce = "sum" if self.aggr in ("add", "sum") else self.aggr
        out = scatter(out, i, 0, dim_size=size[1], reduce=reduce)
        return self.update(out)

    def message(self, x_j):
        return x_j

    def update(self, out):
        return out


class SAGEConv(MessagePassing):
    """``out_i = lin_l(mean_{j->i} x_j) + lin_r(x_i)`` -- aggregate RAW features first,
    no L2-normalise (models/utils.py:14: normalize=False, bias=True)."""

    def __init__(self, in_channels, out_channels, aggr="mean", normalize=False, root_weight=True, bias=True, **kw):
        super().__init__(aggr=aggr)
        if isinstance(in_channels, int):
            in_channels = (in_channels, in_channels)
        self.in_channels, self.out_channels = in_channels, out_channels
        self.normalize, self.root_weight = normalize, root_weight
        self.lin_l = Linear(in_channels[0], out_channels, bias=bias)
        if root_weight:
            self.lin_r = Linear(in_channels[1], out_channels, bias=False)

    def forward(self, x, edge_index, size=None):
        if isinstance(x, Tensor):
            x = (x, x)
        out = self.propagate(edge_index, x=x, size=size)
        out = self.lin_l(out)
        if self.root_weight and x[1] is not None:
            out = out + self.lin_r(x[1])
        if self.normalize:
            out = F.normalize(out, p=2.0, dim=-1)
        return out


def gcn_norm(edge_index, edge_weight=None, num_nodes=None, improved=False, add_self_loops=True, dtype=None):
    """``D^-1/2 (A + I) D^-1/2`` with degrees over TARGETS (PyG 2.3.1 conv/gcn_conv.py)."""
    fill_value = 2.0 if improved else 1.0
    if edge_weight is None:
        edge_weight = torch.ones((edge_index.size(1),), dtype=dtype, device=edge_index.device)
    if add_self_loops:
        edge_index, edge_weight = add_remaining_self_loops(edge_index, edge_weight, fill_value, num_nodes)
    row, col = edge_index[0], edge_index[1]
    deg = scatter(edge_weight, col, 0, dim_size=num_nodes, reduce="sum")
    deg_inv_sqrt = deg.pow(-0.5)
    deg_inv_sqrt.masked_fill_(deg_inv_sqrt == float("inf"), 0)
    return edge_index, deg_inv_sqrt[row] * edge_weight * deg_inv_sqrt[col]


class GCNConv(MessagePassing):
    def __init__(self, in_channels, out_channels, improved=False, cached=False, add_self_loops=True,
                 normalize=True, bias=True, **kw):
        super().__init__(aggr="add")
        assert isinstance(in_channels, int), "GCNConv is not bipartite"
        self.in_channels, self.out_channels = in_channels, out_channels
        self.improved, self.add_self_loops, self.normalize = improved, add_self_loops, normalize
        self.lin = Linear(in_channels, out_channels, bias=False, weight_initializer="glorot")
        if bias:
            self.bias = nn.Parameter(torch.zeros(out_channels))
        else:
            self.register_parameter("bias", None)

    def forward(self, x, edge_index, edge_weight=None):
        if self.normalize:
            edge_index, edge_weight = gcn_norm(edge_index, edge_weight, x.size(0), self.improved,
                                               self.add_self_loops, dtype=x.dtype)
        x = self.lin(x)
        out = self.propagate(edge_index, x=x, edge_weight=edge_weight)
        if self.bias is not None:
            out = out + self.bias
        return out

    def message(self, x_j, edge_weight):
        return x_j if edge_weight is None else edge_weight.view(-1, 1) * x_j


class GINConv(MessagePassing):
    """``out = nn((1 + eps) x_i + sum_j x_j)``; ``eps`` a learnable ``[1]`` tensor
    (models/utils.py:17-26: eps=0, train_eps=True)."""

    def __init__(self, nn_module, eps=0.0, train_eps=False, **kw):
        super().__init__(aggr="add")
        self.nn = nn_module
        self.initial_eps = eps
        if train_eps:
            self.eps = nn.Parameter(torch.Tensor([eps]))
        else:
            self.register_buffer("eps", torch.Tensor([eps]))

    def forward(self, x, edge_index, size=None):
        if isinstance(x, Tensor):
            x = (x, x)
        out = self.propagate(edge_index, x=x, size=size)
        if x[1] is not None:
            out = out + (1 + self.eps) * x[1]
        return self.nn(out)


class GATConv(MessagePassing):
    """PyG 2.3.1 ``GATConv``.

    int ``in_channels`` -> ``lin_dst is lin_src`` (one shared weight, both keys in the
    state_dict); tuple -> two weights.  Single-Tensor ``x`` -> ``x_dst = x_src =
    lin_src(x)`` (``lin_dst`` unused).  Self loops only if ``add_self_loops``: existing
    loops removed, ``(k,k), k < min(N_src, N_dst)`` appended after the real edges.
    """

    def __init__(self, in_channels, out_channels, heads=1, concat=True, negative_slope=0.2, dropout=0.0,
                 add_self_loops=True, edge_dim=None, fill_value="mean", bias=True, **kw):
        super().__init__(aggr="add")
        self.in_channels, self.out_channels, self.heads = in_channels, out_channels, heads
        self.concat, self.negative_slope, self.dropout = concat, negative_slope, dropout
        self.add_self_loops, self.edge_dim, self.fill_value = add_self_loops, edge_dim, fill_value
        H, C = heads, out_channels
        if isinstance(in_channels, int):
            self.lin_src = Linear(in_channels, H * C, bias=False, weight_initializer="glorot")
            self.lin_dst = self.lin_src
        else:
            self.lin_src = Linear(in_channels[0], H * C, bias=False, weight_initializer="glorot")
            self.lin_dst = Linear(in_channels[1], H * C, bias=False, weight_initializer="glorot")
        self.att_src = nn.Parameter(torch.empty(1, H, C))
        self.att_dst = nn.Parameter(torch.empty(1, H, C))
        if edge_dim is not None:
            self.lin_edge = Linear(edge_dim, H * C, bias=False, weight_initializer="glorot")
            self.att_edge = nn.Parameter(torch.empty(1, H, C))
        else:
            self.lin_edge = None
            self.register_parameter("att_edge", None)
        if bias and concat:
            self.bias = nn.Parameter(torch.zeros(H * C))
        elif bias:
            self.bias = nn.Parameter(torch.zeros(C))
        else:
            self.register_parameter("bias", None)
        glorot(self.att_src)
        glorot(self.att_dst)
        glorot(self.att_edge)

    def forward(self, x, edge_index, edge_attr=None, size=None):
        H, C = self.heads, self.out_channels
        if isinstance(x, Tensor):
            x_src = x_dst = self.lin_src(x).view(-1, H, C)
        else:
            x_src, x_dst = x
            x_src = self.lin_src(x_src).view(-1, H, C)
            if x_dst is not None:
                x_dst = self.lin_dst(x_dst).view(-1, H, C)
        alpha_src = (x_src * self.att_src).sum(dim=-1)
        alpha_dst = None if x_dst is None else (x_dst * self.att_dst).sum(dim=-1)
        if self.add_self_loops:
            num_nodes = x_src.size(0)
            if x_dst is not None:
                num_nodes = min(num_nodes, x_dst.size(0))
            num_nodes = min(size) if size is not None else num_nodes
            edge_index, edge_attr = _remove_self_loops(edge_index, edge_attr)
            edge_index, edge_attr = _add_self_loops(edge_index, edge_attr, fill_value=self.fill_value,
                                                    num_nodes=num_nodes)
        n_dst = x_dst.size(0) if x_dst is not None else x_src.size(0)
        j, i = edge_index[0], edge_index[1]
        # edge_updater
        alpha = alpha_src.index_select(0, j)
        if alpha_dst is not None:
            alpha = alpha + alpha_dst.index_select(0, i)
        if i.numel() > 0:
            if edge_attr is not None and self.lin_edge is not None:
                if edge_attr.dim() == 1:
                    edge_attr = edge_attr.view(-1, 1)
                e = self.lin_edge(edge_attr).view(-1, H, C)
                alpha = alpha + (e * self.att_edge).sum(dim=-1)
            alpha = F.leaky_relu(alpha, self.negative_slope)
            alpha = softmax(alpha, i, None, n_dst)
            alpha = F.dropout(alpha, p=self.dropout, training=self.training)
        # propagate: message = alpha.unsqueeze(-1) * x_j ; aggr = add over targets
        msg = alpha.unsqueeze(-1) * x_src.index_select(0, j)
        out = scatter(msg, i, 0, dim_size=n_dst, reduce="sum")
        out = out.view(-1, H * C) if self.concat else out.mean(dim=1)
        if self.bias is not None:
            out = out + self.bias
        return out


def group(xs, aggr):
    if len(xs) == 0:
        return None
    if aggr is None:
        return torch.stack(xs, dim=1)
    if len(xs) == 1:
        return xs[0]
    if aggr == "cat":
        return torch.cat(xs, dim=-1)
    out = torch.stack(xs, dim=0)
    out = getattr(torch, aggr)(out, dim=0)
    return out[0] if isinstance(out, tuple) else out


class HeteroConv(nn.Module):
    """PyG 2.3.1 ``HeteroConv``: iterate ``edge_index_dict`` in ITS
    order, skip relations with no registered conv; ``src == dst`` -> single Tensor,
    else ``(x_src, x_dst)``; extra positional dicts looked up by edge type; outputs
    grouped per dst type with ``aggr``.  Module keys ``'__'.join(edge_type)``."""

    def __init__(self, convs: Dict[Tuple[str, str, str], nn.Module], aggr: Optional[str] = "sum"):
        super().__init__()
        self.convs = nn.ModuleDict({"__".join(k): v for k, v in convs.items()})
        self.aggr = aggr

    def forward(self, x_dict, edge_index_dict, *args_dict, **kwargs_dict):
        out_dict = defaultdict(list)
        for edge_type, edge_index in edge_index_dict.items():
            src, rel, dst = edge_type
            str_edge_type = "__".join(edge_type)
            if str_edge_type not in self.convs:
                continue
            args = []
            for value_dict in args_dict:
                if edge_type in value_dict:
                    args.append(value_dict[edge_type])
                elif src == dst and src in value_dict:
                    args.append(value_dict[src])
                elif src in value_dict or dst in value_dict:
                    args.append((value_dict.get(src, None), value_dict.get(dst, None)))
            conv = self.convs[str_edge_type]
            if src == dst:
                out = conv(x_dict[src], edge_index, *args)
            else:
                out = conv((x_dict[src], x_dict[dst]), edge_index, *args)
            out_dict[dst].append(out)
        return {key: group(value, self.aggr) for key, value in out_dict.items()}


class BatchNorm(nn.Module):
    """PyG ``BatchNorm``: thin wrapper, parameters live under ``.module.``."""

    def __init__(self, in_channels, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True, **kw):
        super().__init__()
        self.module = nn.BatchNorm1d(in_channels, eps, momentum, affine, track_running_stats)

    def forward(self, x):
        return self.module(x)
