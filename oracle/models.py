"""Restatement of the control flow of the reference's four model classes.

ORACLE -- test infrastructure only (see ``oracle/__init__.py``).

Hosted on ``oracle.pyg`` (the PyG 2.3.1 restatement).  Each class cites the
reference lines it follows; parameter names / module tree are identical to the
reference so ``state_dict``s interchange with it and with the product classes.
Validated in the build container against the reference's *own* files imported
through ``oracle.ref_import`` (tests/golden/make_golden.py; tests/test_oracle.py).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .pyg import nn as pyg_nn
from .pyg.nn import BatchNorm, HeteroConv, MessagePassing

# mp3d_dataset.py:21-26
EDGE_TYPES = [
    ("objects", "objects_to_objects", "objects"),
    ("rooms", "rooms_to_rooms", "rooms"),
    ("objects", "objects_to_rooms", "rooms"),
    ("rooms", "rooms_to_objects", "objects"),
]
# neural_tree/construct.py:14-38 (the "or_ro_rr"/"rr_ro_or" spellings are the reference's)
HTREE_NODE_TYPES = ["object", "room", "object-room", "room-room"]
HTREE_EDGE_TYPES = [
    ("object", "o_to_or", "object-room"),
    ("object-room", "or_to_o", "object"),
    ("room", "r_to_or", "object-room"),
    ("object-room", "or_to_r", "room"),
    ("room", "r_to_rr", "room-room"),
    ("room-room", "rr_to_r", "room"),
    ("object-room", "or_ro_rr", "room-room"),
    ("room-room", "rr_ro_or", "object-room"),
    ("object-room", "or_to_or", "object-room"),
    ("room-room", "rr_to_rr", "room-room"),
]
HTREE_VIRTUAL_NODE_TYPES = ["object_virtual", "room_virtual"]
HTREE_INIT_EDGE_TYPES = [
    ("object_virtual", "ov_to_or", "object-room"),
    ("room_virtual", "rv_to_or", "object-room"),
    ("room_virtual", "rv_to_rr", "room-room"),
]
HTREE_POOL_EDGE_TYPES = [
    ("object", "o_to_ov", "object_virtual"),
    ("room", "r_to_rv", "room_virtual"),
]


# ----------------------------------------------------------------------------- factories
def make_conv(kind, d_in, d_out):
    """models/utils.py:9-28."""
    if kind == "GraphSAGE":
        return pyg_nn.SAGEConv(d_in, d_out, normalize=False, bias=True)
    if kind == "GCN":
        return pyg_nn.GCNConv(d_in, d_out, add_self_loops=True)
    if kind == "GIN":
        mlp = nn.Sequential(nn.Linear(d_in, d_out), nn.ReLU(), nn.Linear(d_out, d_out))
        return pyg_nn.GINConv(mlp, eps=0.0, train_eps=True)
    return NotImplemented


def make_gat_stack(d_in, dims, heads, concats, dropout=0.0, add_self_loop=True, edge_dim=None, fill_value="mean"):
    """models/utils.py:31-87: layer l's input width is dims[l-1]*heads[l-1] if
    concats[l-1] else dims[l-1]."""
    assert len(dims) == len(heads) == len(concats)
    stack, width = nn.ModuleList(), d_in
    for l, (c, h, cat) in enumerate(zip(dims, heads, concats)):
        stack.append(pyg_nn.GATConv(width, c, heads=h, concat=cat, dropout=dropout, add_self_loops=add_self_loop,
                                    edge_dim=edge_dim, fill_value=fill_value))
        width = c * h if cat else c
    return stack


def make_hetero(kind, edge_types, din, dout, aggr="sum"):
    """models/utils.py:90-100."""
    return HeteroConv({(s, r, t): make_conv(kind, (din[s], din[t]), dout[t]) for s, r, t in edge_types}, aggr=aggr)


def make_gat_hetero(edge_types, din, dout, dims, heads, concats, dropout, aggr="sum", edge_dim=None, fill_value="mean"):
    """models/utils.py:103-140: per relation a GAT stack with bipartite input dims and
    self loops only when src == dst; layer l of every relation wrapped in one HeteroConv."""
    stacks = {
        (s, r, t): make_gat_stack((din[s], din[t]), dims + [dout[t]], heads, concats, dropout,
                                  add_self_loop=(s == t), edge_dim=edge_dim, fill_value=fill_value)
        for s, r, t in edge_types
    }
    return nn.ModuleList(HeteroConv({et: st[l] for et, st in stacks.items()}, aggr=aggr) for l in range(len(heads)))


def cross_entropy_loss(pred, label, mask=None):
    """models/utils.py:143-161."""
    if isinstance(pred, torch.Tensor):
        return F.cross_entropy(pred, label) if mask is None else F.cross_entropy(pred[mask, :], label[mask])
    if mask is None:
        mask = [torch.ones_like(l, dtype=torch.bool) for l in label]
    total = sum(F.cross_entropy(p[m, :], l[m], reduction="sum") for p, l, m in zip(pred, label, mask))
    return total / sum(int(m.sum()) for m in mask)


class _Common(nn.Module):
    def _act(self, x):
        return F.elu(x) if self.conv_block[:3] == "GAT" else F.relu(x)

    def _act_drop(self, x):
        return F.dropout(self._act(x), p=self.dropout, training=self.training)

    def loss(self, pred, label, mask=None):
        return cross_entropy_loss(pred, label, mask)


# ----------------------------------------------------------------------------- hetero
def _hetero_layers(self, edge_types, din, dout, hidden_dim, num_layers, dims, heads, concats, dropout, node_types):
    """heterogeneous_network.py:55-97 / heterogeneous_neural_tree_network.py:106-149."""
    if self.conv_block in ("GAT", "GAT_edge"):
        extra = {}
        if self.conv_block == "GAT_edge":
            extra = dict(edge_dim=3, fill_value=torch.zeros(3, dtype=torch.float64))
        return make_gat_hetero(edge_types, din, dout, dims, heads, concats, dropout, **extra)
    hid = {t: hidden_dim for t in node_types}
    widths = [din] + [hid] * (num_layers - 1) + [dout]
    return nn.ModuleList(make_hetero(self.conv_block, edge_types, widths[l], widths[l + 1]) for l in range(num_layers))


class HeterogeneousNetwork(_Common):
    """heterogeneous_network.py:12-136."""

    def __init__(self, input_dim_dict, output_dim=None, output_dim_dict=None, conv_block="GraphSAGE",
                 hidden_dim=None, num_layers=None, GAT_hidden_dims=None, GAT_heads=None, GAT_concats=None,
                 dropout=0.25, **kwargs):
        super().__init__()
        assert conv_block in ["GraphSAGE", "GAT", "GAT_edge"]  # :41
        self.conv_block = conv_block
        if output_dim is not None:  # :43-50
            assert output_dim_dict is None
            self.classification_task = "room"
            output_dim_dict = {"rooms": output_dim, "objects": output_dim}
        else:
            assert output_dim_dict is not None
            self.classification_task = "all"
        self.num_layers = num_layers if conv_block[:3] != "GAT" else len(GAT_heads)
        self.dropout = dropout
        self.convs = _hetero_layers(self, EDGE_TYPES, input_dim_dict, output_dim_dict, hidden_dim, self.num_layers,
                                    GAT_hidden_dims, GAT_heads, GAT_concats, dropout, ["rooms", "objects"])

    def forward(self, data):  # :99-133
        x, ei = data.x_dict, data.edge_index_dict
        for l, conv in enumerate(self.convs):
            x = conv(x, ei, data.edge_attr_dict) if self.conv_block == "GAT_edge" else conv(x, ei)
            if l != self.num_layers - 1:
                x = {k: self._act_drop(v) for k, v in x.items()}
        if self.classification_task == "room":
            return x["rooms"]
        x = {k: self._act_drop(v) for k, v in x.items()}
        return x["rooms"], x["objects"]


class LeafPool(MessagePassing):
    """heterogeneous_neural_tree_network.py:18-31: identity message, mean aggregation;
    single-Tensor input => output has x.size(0) rows."""

    def __init__(self, aggr="mean"):
        super().__init__(aggr=aggr)

    def forward(self, x, edge_index):
        return self.propagate(edge_index, x=x)

    def message(self, x_j):
        return x_j


class HeterogeneousNeuralTreeNetwork(_Common):
    """heterogeneous_neural_tree_network.py:34-205."""

    def __init__(self, input_dim_dict, output_dim=None, output_dim_dict=None, conv_block="GraphSAGE",
                 disable_initialization=False, hidden_dim=None, num_layers=None, GAT_hidden_dims=None,
                 GAT_heads=None, GAT_concats=None, dropout=0.25, **kwargs):
        super().__init__()
        assert conv_block in ["GraphSAGE", "GAT", "GAT_edge"]  # :71
        self.conv_block = conv_block
        if output_dim is not None:
            assert output_dim_dict is None
            self.classification_task = "room"
            output_dim_dict = {t: output_dim for t in HTREE_NODE_TYPES}
        else:
            assert output_dim_dict is not None
            self.classification_task = "all"
        self.num_layers = num_layers if conv_block[:3] != "GAT" else len(GAT_heads)
        self.dropout = dropout
        assert input_dim_dict["object"] == input_dim_dict["object_virtual"]  # :84-86
        assert input_dim_dict["room"] == input_dim_dict["room_virtual"]
        assert input_dim_dict["object-room"] == input_dim_dict["room-room"]
        if disable_initialization:  # :89-103
            self.pre_mp = None
        else:
            self.pre_mp = HeteroConv({
                (s, r, t): pyg_nn.GATConv((input_dim_dict[s], input_dim_dict[t]), input_dim_dict[t], heads=1,
                                          concat=False, dropout=0.0, add_self_loops=False)
                for s, r, t in HTREE_INIT_EDGE_TYPES}, aggr="mean")
        din = {t: input_dim_dict[t] for t in HTREE_NODE_TYPES}
        self.convs = _hetero_layers(self, HTREE_EDGE_TYPES, din, output_dim_dict, hidden_dim, self.num_layers,
                                    GAT_hidden_dims, GAT_heads, GAT_concats, dropout, HTREE_NODE_TYPES)
        self.post_mp = LeafPool(aggr="mean")

    def forward(self, data):  # :154-202
        x, ei = data.x_dict, data.edge_index_dict
        if self.pre_mp is not None:
            x.update(self.pre_mp(x, ei))
        for l, conv in enumerate(self.convs):
            x = conv(x, ei, data.edge_attr_dict) if self.conv_block == "GAT_edge" else conv(x, ei)
            if l != self.num_layers - 1:
                x = {k: self._act_drop(v) for k, v in x.items()}
        if self.classification_task != "room":
            x = {k: self._act_drop(v) for k, v in x.items()}
        rooms = self.post_mp(x["room"], ei["room", "r_to_rv", "room_virtual"])[0: data["room_virtual"].num_nodes]
        if self.classification_task == "room":
            return rooms
        objects = self.post_mp(x["object"], ei["object", "o_to_ov", "object_virtual"])[
            0: data["object_virtual"].num_nodes]
        return rooms, objects


# ----------------------------------------------------------------------------- homogeneous
class HomogeneousNetwork(_Common):
    """homogeneous_network.py:12-148."""

    def __init__(self, input_dim, output_dim=None, output_dim_dict=None, conv_block="GCN", hidden_dim=None,
                 num_layers=None, GAT_hidden_dims=None, GAT_heads=None, GAT_concats=None, dropout=0.25, **kwargs):
        super().__init__()
        self.conv_block = conv_block
        gat = conv_block[:3] == "GAT"
        if output_dim is not None:  # :44-59
            assert output_dim_dict is None
            self.classification_task = "room"
            mp_out, gat_dims = output_dim, (GAT_hidden_dims + [output_dim]) if gat else None
        else:
            assert output_dim_dict is not None
            self.classification_task = "all"
            mp_out, gat_dims = hidden_dim, GAT_hidden_dims
        self.num_layers = num_layers if not gat else len(GAT_heads)
        self.dropout = dropout
        if conv_block == "GAT":  # :64-71
            self.convs = make_gat_stack(input_dim, gat_dims, GAT_heads, GAT_concats, dropout=dropout)
        elif conv_block == "GAT_edge":  # :72-82
            self.convs = make_gat_stack(input_dim, gat_dims, GAT_heads, GAT_concats, dropout=dropout, edge_dim=3,
                                        add_self_loop=True, fill_value=torch.zeros(3, dtype=torch.float64))
        else:  # :83-91
            widths = [input_dim] + [hidden_dim] * (self.num_layers - 1) + [mp_out]
            self.convs = nn.ModuleList(make_conv(conv_block, widths[l], widths[l + 1]) for l in range(self.num_layers))
        if conv_block == "GIN":  # :94-97 (one per layer, last one unused)
            self.batch_norms = nn.ModuleList(BatchNorm(hidden_dim) for _ in range(self.num_layers))
        if self.classification_task == "all":  # :99-120
            if not gat:
                final = hidden_dim
            else:
                final = gat_dims[-1] * GAT_heads[-1] if GAT_concats[-1] else gat_dims[-1]
            n_room = output_dim_dict["rooms"] if "rooms" in output_dim_dict else output_dim_dict["room"]
            n_obj = output_dim_dict["objects"] if "objects" in output_dim_dict else output_dim_dict["object"]
            self.post_mp_room = nn.Linear(final, n_room)
            self.post_mp_object = nn.Linear(final, n_obj)

    def _mp(self, x, data, use_bn):
        for l, conv in enumerate(self.convs):
            if self.conv_block == "GAT_edge":
                x = conv(x, edge_index=data.edge_index, edge_attr=data.edge_attr)
            else:
                x = conv(x, edge_index=data.edge_index)
            if l != self.num_layers - 1:
                if use_bn and self.conv_block == "GIN":
                    x = self.batch_norms[l](x)
                x = self._act_drop(x)
        return x

    def forward(self, data):  # :122-145
        x = self._mp(data.x, data, use_bn=True)
        if self.classification_task == "room":
            return x[data.room_mask, :]
        x = self._act_drop(x)
        return self.post_mp_room(x[data.room_mask, :]), self.post_mp_object(x[~data.room_mask, :])


class HomogeneousNeuralTreeNetwork(HomogeneousNetwork):
    """homogeneous_neural_tree_network.py:7-109.  NB the reference's forward applies NO
    batch norm for GIN (:88-97) although the parent registers the modules."""

    def __init__(self, input_dim, output_dim=None, output_dim_dict=None, conv_block="GCN",
                 disable_initialization=False, hidden_dim=None, num_layers=None, GAT_hidden_dims=None,
                 GAT_heads=None, GAT_concats=None, dropout=0.25, **kwargs):
        super().__init__(input_dim, output_dim, output_dim_dict, conv_block, hidden_dim, num_layers,
                         GAT_hidden_dims, GAT_heads, GAT_concats, dropout, **kwargs)
        assert conv_block in ["GCN", "GraphSAGE", "GAT", "GAT_edge", "GIN"]  # :56
        if disable_initialization:
            self.pre_mp = None
        else:  # :63-70; int in_channels => shared lin_src/lin_dst; overwrites ALL rows (:86)
            self.pre_mp = pyg_nn.GATConv(input_dim, input_dim, heads=1, concat=False, dropout=0.0,
                                         add_self_loops=False)
        self.post_mp_pool = LeafPool(aggr="mean")

    def forward(self, data):  # :75-109
        x = data.x
        if self.pre_mp is not None:
            x = self.pre_mp(x, edge_index=data.init_edge_index)
        x = self._mp(x, data, use_bn=False)
        x = self.post_mp_pool(x, data.pool_edge_index)
        if self.classification_task == "room":
            return x[data.room_mask, :]
        x = self._act_drop(x)
        return self.post_mp_room(x[data.room_mask, :]), self.post_mp_object(x[data.object_mask, :])
