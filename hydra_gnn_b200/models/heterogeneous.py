"""``HeterogeneousNetwork`` and ``HeterogeneousNeuralTreeNetwork`` (drop-in).

Constructor kwargs, attributes, ``forward(data)`` contract and parameter names follow
heterogeneous_network.py:12-136 and heterogeneous_neural_tree_network.py:34-205.  Each
``HeteroConv`` layer runs as: one fused projection per node type, one fused GAT / SAGE
aggregation per destination type with the cross-relation sum and the activation in the
epilogue; the CSR plan is built once per batch and shared by all layers.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import nn as hnn
from ..nn import PlanRequest
from ..schema import EDGE_TYPES, HTREE_EDGE_TYPES, HTREE_INIT_EDGE_TYPES, HTREE_NODE_TYPES
from .common import _Base, build_GAT_hetero_conv, build_hetero_conv


def _build_layers(model, edge_types, node_types, input_dim_dict, output_dim_dict, hidden_dim, GAT_hidden_dims,
                  GAT_heads, GAT_concats, dropout):
    """heterogeneous_network.py:55-97 / heterogeneous_neural_tree_network.py:106-149."""
    if model.conv_block == "GAT":
        return build_GAT_hetero_conv(edge_types, input_dim_dict, output_dim_dict, GAT_hidden_dims, GAT_heads,
                                     GAT_concats, dropout)
    if model.conv_block == "GAT_edge":
        return build_GAT_hetero_conv(edge_types, input_dim_dict, output_dim_dict, GAT_hidden_dims, GAT_heads,
                                     GAT_concats, dropout, edge_dim=3,
                                     fill_value=torch.zeros(3, dtype=torch.float64))
    hidden = {t: hidden_dim for t in node_types}
    dims = [input_dim_dict] + [hidden] * (model.num_layers - 1) + [output_dim_dict]
    return nn.ModuleList(build_hetero_conv(model.conv_block, edge_types, dims[l], dims[l + 1])
                         for l in range(model.num_layers))


class HeterogeneousNetwork(_Base):
    """Message passing on heterogeneous room-object graphs (4 relations ``EDGE_TYPES``)."""

    def __init__(self, input_dim_dict, output_dim=None, output_dim_dict=None, conv_block="GraphSAGE",
                 hidden_dim=None, num_layers=None, GAT_hidden_dims=None, GAT_heads=None, GAT_concats=None,
                 dropout=0.25, **kwargs):
        super().__init__()
        assert conv_block in ["GraphSAGE", "GAT", "GAT_edge"]
        self.conv_block = conv_block
        if output_dim is not None:
            assert output_dim_dict is None
            self.classification_task = "room"
            output_dim_dict = {"rooms": output_dim, "objects": output_dim}  # object logits are never returned
        else:
            assert output_dim_dict is not None
            self.classification_task = "all"
        self.num_layers = num_layers if conv_block[:3] != "GAT" else len(GAT_heads)
        self.dropout = dropout
        self.convs = _build_layers(self, EDGE_TYPES, ["rooms", "objects"], input_dim_dict, output_dim_dict,
                                   hidden_dim, GAT_hidden_dims, GAT_heads, GAT_concats, dropout)
        self._init_runtime()

    def _forward(self, data):
        x_dict, ei_dict = dict(data.x_dict), data.edge_index_dict
        ea_dict = data.edge_attr_dict if self.conv_block == "GAT_edge" else None
        plan = self._plans.get(self.convs[0].plan_requests(x_dict, ei_dict))
        room_only = self.classification_task == "room"
        for l, conv in enumerate(self.convs):
            last = l == self.num_layers - 1
            # task "all" applies one more activation + dropout after the last layer (:123-132)
            act = None if (last and room_only) else self._act_name
            x_dict = conv.run(x_dict, ei_dict, plan, act=act, edge_attr_dict=ea_dict,
                              only_dst=["rooms"] if (last and room_only) else None)
            if act is not None:
                x_dict = self._drop(x_dict)
        if room_only:
            return x_dict["rooms"]
        return x_dict["rooms"], x_dict["objects"]


class HeterogeneousNeuralTreeNetwork(_Base):
    """Message passing on augmented H-trees: optional clique initialisation from the virtual
    nodes (``pre_mp``), L layers over the 10 ``HTREE_EDGE_TYPES``, leaf -> virtual mean readout."""

    def __init__(self, input_dim_dict, output_dim=None, output_dim_dict=None, conv_block="GraphSAGE",
                 disable_initialization=False, hidden_dim=None, num_layers=None, GAT_hidden_dims=None,
                 GAT_heads=None, GAT_concats=None, dropout=0.25, **kwargs):
        super().__init__()
        assert conv_block in ["GraphSAGE", "GAT", "GAT_edge"]
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

        assert input_dim_dict["object"] == input_dim_dict["object_virtual"]
        assert input_dim_dict["room"] == input_dim_dict["room_virtual"]
        assert input_dim_dict["object-room"] == input_dim_dict["room-room"]

        if disable_initialization:
            self.pre_mp = None
            print("diable initialization")  # the reference's message (heterogeneous_neural_tree_network.py:91)
        else:
            self.pre_mp = hnn.HeteroConv({
                (s, r, t): hnn.GATConv((input_dim_dict[s], input_dim_dict[t]), input_dim_dict[t], heads=1,
                                       concat=False, dropout=0.0, add_self_loops=False)
                for s, r, t in HTREE_INIT_EDGE_TYPES}, aggr="mean")

        mp_in = {t: input_dim_dict[t] for t in HTREE_NODE_TYPES}
        self.convs = _build_layers(self, HTREE_EDGE_TYPES, HTREE_NODE_TYPES, mp_in, output_dim_dict, hidden_dim,
                                   GAT_hidden_dims, GAT_heads, GAT_concats, dropout)
        self.post_mp = hnn.LeafPool(aggr="mean")
        self._init_runtime()

    _ROOM_POOL = ("room", "r_to_rv", "room_virtual")
    _OBJECT_POOL = ("object", "o_to_ov", "object_virtual")

    def _forward(self, data):
        x_dict, ei_dict = dict(data.x_dict), data.edge_index_dict
        ea_dict = data.edge_attr_dict if self.conv_block == "GAT_edge" else None
        room_only = self.classification_task == "room"
        n_rv = int(data["room_virtual"].num_nodes)
        pools = [(self._ROOM_POOL, n_rv)]
        if not room_only:
            pools.append((self._OBJECT_POOL, int(data["object_virtual"].num_nodes)))

        requests = list(self.convs[0].plan_requests(x_dict, ei_dict))
        if self.pre_mp is not None:
            requests += self.pre_mp.plan_requests(x_dict, ei_dict)
        for et, n_out in pools:
            requests.append(PlanRequest(et, ei_dict[et], x_dict[et[0]].size(0), n_out))
        plan = self._plans.get(requests)

        if self.pre_mp is not None:  # only the clique types are overwritten (x_dict.update, :159)
            x_dict.update(self.pre_mp.run(x_dict, ei_dict, plan))

        keep = ["room"] if room_only else ["room", "object"]
        for l, conv in enumerate(self.convs):
            last = l == self.num_layers - 1
            act = None if (last and room_only) else self._act_name
            x_dict = conv.run(x_dict, ei_dict, plan, act=act, edge_attr_dict=ea_dict,
                              only_dst=keep if last else None)
            if act is not None:
                x_dict = self._drop(x_dict)

        from .. import ops
        outs = [hnn.seg_reduce(x_dict[et[0]], plan[(et, False)], "mean") for et, _ in pools]
        return outs[0] if room_only else (outs[0], outs[1])
