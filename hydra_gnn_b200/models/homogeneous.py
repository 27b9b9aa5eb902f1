"""``HomogeneousNetwork`` and ``HomogeneousNeuralTreeNetwork`` (drop-in).

Follows homogeneous_network.py:12-148 and homogeneous_neural_tree_network.py:7-109
(constructor kwargs, attributes, ``forward(data)`` contract, parameter names).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import nn as hnn
from .. import ops
from ..nn import PlanRequest
from .common import _Base, build_conv_layer, build_GAT_conv_layers

_REL = ("_", "edge_index", "_")


class HomogeneousNetwork(_Base):
    """L conv layers on a flat ``Data`` (``x``, ``edge_index``, ``room_mask``)."""

    def __init__(self, input_dim, output_dim=None, output_dim_dict=None, conv_block="GCN", hidden_dim=None,
                 num_layers=None, GAT_hidden_dims=None, GAT_heads=None, GAT_concats=None, dropout=0.25, **kwargs):
        super().__init__()
        self.conv_block = conv_block
        is_gat = conv_block[:3] == "GAT"
        if output_dim is not None:
            assert output_dim_dict is None
            self.classification_task = "room"
            mp_output_dim = output_dim
            gat_dims = (list(GAT_hidden_dims) + [output_dim]) if is_gat else None
        else:
            assert output_dim_dict is not None
            self.classification_task = "all"
            mp_output_dim = hidden_dim
            gat_dims = GAT_hidden_dims
        self.num_layers = num_layers if not is_gat else len(GAT_heads)
        self.dropout = dropout

        if conv_block == "GAT":
            self.convs = build_GAT_conv_layers(input_dim, gat_dims, GAT_heads, GAT_concats, dropout=dropout)
        elif conv_block == "GAT_edge":
            self.convs = build_GAT_conv_layers(input_dim, gat_dims, GAT_heads, GAT_concats, dropout=dropout,
                                               edge_dim=3, add_self_loop=True,
                                               fill_value=torch.zeros(3, dtype=torch.float64))
        else:
            widths = [input_dim] + [hidden_dim] * (self.num_layers - 1) + [mp_output_dim]
            self.convs = nn.ModuleList(build_conv_layer(conv_block, widths[l], widths[l + 1])
                                       for l in range(self.num_layers))

        if conv_block == "GIN":  # one BatchNorm per layer is registered; the last is never applied (:94-97,133)
            self.batch_norms = nn.ModuleList(hnn.BatchNorm(hidden_dim) for _ in range(self.num_layers))

        if self.classification_task == "all":
            if not is_gat:
                final = hidden_dim
            else:
                final = gat_dims[-1] * GAT_heads[-1] if GAT_concats[-1] else gat_dims[-1]
            n_room = output_dim_dict["rooms"] if "rooms" in output_dim_dict else output_dim_dict["room"]
            n_obj = output_dim_dict["objects"] if "objects" in output_dim_dict else output_dim_dict["object"]
            self.post_mp_room = nn.Linear(final, n_room)
            self.post_mp_object = nn.Linear(final, n_obj)
        self._init_runtime()

    # ------------------------------------------------------------------ plan
    def _main_request(self, data) -> PlanRequest:
        n = data.x.size(0)
        if self.conv_block[:3] == "GAT":
            return PlanRequest(_REL, data.edge_index, n, n, True, want_perm=self.conv_block == "GAT_edge",
                               want_window=any(c.wants_window() for c in self.convs))
        if self.conv_block == "GCN":
            return PlanRequest(_REL, data.edge_index, n, n, True, want_norm=True)
        return PlanRequest(_REL, data.edge_index, n, n, False)

    @staticmethod
    def _mask_rows(plan, name: str, mask: torch.Tensor, negate: bool = False) -> torch.Tensor:
        """``nonzero(mask)`` as int32, once per mask tensor (``PlanCache.mask_rows``)."""
        return plan.mask_rows(name, mask, negate)

    # ------------------------------------------------------------------ layers
    def _message_passing(self, x, data, plan, use_batch_norm: bool, final_act: bool):
        csr = plan[(_REL, self._main_request(data).self_loops)]
        for l, conv in enumerate(self.convs):
            last = l == self.num_layers - 1
            act = self._act_name if (not last or final_act) else None
            if isinstance(conv, hnn.GATConv):
                et = ("n", "e", "n")
                ea = {et: data.edge_attr} if self.conv_block == "GAT_edge" else None
                x = hnn.run_gat_layer({et: conv}, [et], {"n": x}, {et: csr}, act=act, edge_attr=ea, owner=conv)["n"]
            elif isinstance(conv, hnn.SAGEConv):
                x = hnn.run_sage(conv, x, x, csr, act=act)
            elif isinstance(conv, hnn.GCNConv):
                x = hnn.run_gcn(conv, x, csr, act=act)
            elif isinstance(conv, hnn.GINConv):
                bn = self.batch_norms[l] if (use_batch_norm and not last) else None
                x = hnn.run_gin(conv, x, csr, act=act, bn=bn)
            else:
                raise NotImplementedError(type(conv).__name__)
            if act is not None:
                x = self._drop(x)
        return x

    def _heads(self, x, room_rows, object_rows):
        xr, xo = hnn.gather_rows(x, room_rows), hnn.gather_rows(x, object_rows)
        return (hnn.linear(xr, self.post_mp_room.weight, self.post_mp_room.bias),
                hnn.linear(xo, self.post_mp_object.weight, self.post_mp_object.bias))

    def _forward(self, data):
        plan = self._plans.get([self._main_request(data)])
        room_only = self.classification_task == "room"
        x = self._message_passing(data.x, data, plan, use_batch_norm=True, final_act=not room_only)
        room_rows = self._mask_rows(self._plans, "room_rows", data.room_mask)
        if room_only:
            return hnn.gather_rows(x, room_rows)
        return self._heads(x, room_rows, self._mask_rows(self._plans, "not_room_rows", data.room_mask, negate=True))


class HomogeneousNeuralTreeNetwork(HomogeneousNetwork):
    """Flat H-tree variant: ``pre_mp`` GATConv over ``init_edge_index`` (overwrites every row,
    :86), L layers over the tree edges, ``LeafPool`` over ``pool_edge_index``, then the
    ``room_mask`` (= room_virtual) rows.  The reference applies no BatchNorm here (:88-97)."""

    def __init__(self, input_dim, output_dim=None, output_dim_dict=None, conv_block="GCN",
                 disable_initialization=False, hidden_dim=None, num_layers=None, GAT_hidden_dims=None,
                 GAT_heads=None, GAT_concats=None, dropout=0.25, **kwargs):
        super().__init__(input_dim, output_dim, output_dim_dict, conv_block, hidden_dim, num_layers,
                         GAT_hidden_dims, GAT_heads, GAT_concats, dropout, **kwargs)
        assert conv_block in ["GCN", "GraphSAGE", "GAT", "GAT_edge", "GIN"]
        if disable_initialization:
            self.pre_mp = None
            print("diable initialization")  # the reference's message (homogeneous_neural_tree_network.py:61)
        else:
            self.pre_mp = hnn.GATConv(input_dim, input_dim, heads=1, concat=False, dropout=0.0, add_self_loops=False)
        self.post_mp_pool = hnn.LeafPool(aggr="mean")

    _INIT = ("_", "init_edge_index", "_")
    _POOL = ("_", "pool_edge_index", "_")

    def _forward(self, data):
        n = data.x.size(0)
        requests = [self._main_request(data), PlanRequest(self._POOL, data.pool_edge_index, n, n)]
        if self.pre_mp is not None:
            requests.append(PlanRequest(self._INIT, data.init_edge_index, n, n))
        plan = self._plans.get(requests)
        room_only = self.classification_task == "room"

        x = data.x
        if self.pre_mp is not None:
            et = ("n", "init", "n")
            x = hnn.run_gat_layer({et: self.pre_mp}, [et], {"n": x}, {et: plan[(self._INIT, False)]},
                                  owner=self.pre_mp)["n"]
        x = self._message_passing(x, data, plan, use_batch_norm=False, final_act=False)

        pool = plan[(self._POOL, False)]
        room_rows = self._mask_rows(self._plans, "room_rows", data.room_mask)
        if room_only:  # LeafPool evaluated only on the rows the caller keeps
            return hnn.seg_reduce(x, pool, "mean", row_map=room_rows)
        # task "all": activation (+ dropout) after the pooling (:105-109)
        object_rows = self._mask_rows(self._plans, "object_rows", data.object_mask)
        xr = self._drop(hnn.seg_reduce(x, pool, "mean", row_map=room_rows, act=self._act_name))
        xo = self._drop(hnn.seg_reduce(x, pool, "mean", row_map=object_rows, act=self._act_name))
        return (hnn.linear(xr, self.post_mp_room.weight, self.post_mp_room.bias),
                hnn.linear(xo, self.post_mp_object.weight, self.post_mp_object.bias))
