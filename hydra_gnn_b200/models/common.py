"""Layer factories and shared behaviour of the four drop-in model classes.

Mirror of the reference's ``models/utils.py`` (layer wiring :9-140, loss :143-161) on top
of ``hydra_gnn_b200.nn``; the module tree -- hence every ``state_dict`` key -- is the
reference's.
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import nn as hnn


def build_conv_layer(conv_block: str, input_dim, output_dim: int, **kwargs):
    """models/utils.py:9-28 (SAGEConv normalize=False bias=True; GCNConv add_self_loops=True;
    GINConv(MLP(Linear, ReLU, Linear), eps=0, train_eps=True))."""
    if conv_block == "GraphSAGE":
        return hnn.SAGEConv(input_dim, output_dim, normalize=False, bias=True)
    if conv_block == "GCN":
        return hnn.GCNConv(input_dim, output_dim, add_self_loops=True)
    if conv_block == "GIN":
        mlp = nn.Sequential(nn.Linear(input_dim, output_dim), nn.ReLU(), nn.Linear(output_dim, output_dim))
        return hnn.GINConv(mlp, eps=0.0, train_eps=True)
    return NotImplemented


def build_GAT_conv_layers(input_dim, hidden_dims: Sequence[int], heads: Sequence[int], concats: Sequence[bool],
                          dropout: float = 0.0, add_self_loop: bool = True, edge_dim=None, fill_value="mean"):
    """models/utils.py:31-87: a stack of GATConv; the input width of layer l is
    ``hidden_dims[l-1] * heads[l-1]`` if ``concats[l-1]`` else ``hidden_dims[l-1]``."""
    assert len(hidden_dims) == len(heads)
    assert len(hidden_dims) == len(concats)
    layers = nn.ModuleList()
    width = input_dim
    for c, h, cat in zip(hidden_dims, heads, concats):
        layers.append(hnn.GATConv(width, c, heads=h, concat=cat, dropout=dropout, add_self_loops=add_self_loop,
                                  edge_dim=edge_dim, fill_value=fill_value))
        width = c * h if cat else c
    return layers


def build_hetero_conv(conv_block: str, edge_types, input_dim_dict, output_dim_dict, aggr: str = "sum"):
    """models/utils.py:90-100: one bipartite conv per relation inside a HeteroConv."""
    return hnn.HeteroConv(
        {(s, r, t): build_conv_layer(conv_block, (input_dim_dict[s], input_dim_dict[t]), output_dim_dict[t])
         for s, r, t in edge_types}, aggr=aggr)


def build_GAT_hetero_conv(edge_types, input_dim_dict, output_dim_dict, GAT_hidden_dims, GAT_heads, GAT_concats,
                          dropout, aggr: str = "sum", edge_dim=None, fill_value="mean"):
    """models/utils.py:103-140: a GAT stack per relation (self loops iff source == target, :124),
    layer l of all relations assembled into one HeteroConv."""
    stacks = {}
    for s, r, t in edge_types:
        stacks[s, r, t] = build_GAT_conv_layers((input_dim_dict[s], input_dim_dict[t]),
                                                list(GAT_hidden_dims) + [output_dim_dict[t]], GAT_heads, GAT_concats,
                                                dropout, add_self_loop=(s == t), edge_dim=edge_dim,
                                                fill_value=fill_value)
    return nn.ModuleList(hnn.HeteroConv({et: st[l] for et, st in stacks.items()}, aggr=aggr)
                         for l in range(len(GAT_heads)))


def cross_entropy_loss(pred, label, mask=None):
    """models/utils.py:143-161."""
    if isinstance(pred, torch.Tensor):
        if mask is None:
            return F.cross_entropy(pred, label)
        return F.cross_entropy(pred[mask, :], label[mask])
    if mask is None:
        total = sum(F.cross_entropy(p, l, reduction="sum") for p, l in zip(pred, label))
        count = sum(torch.numel(l) for l in label)
    else:
        total = sum(F.cross_entropy(p[m, :], l[m], reduction="sum") for p, l, m in zip(pred, label, mask))
        count = sum(torch.numel(l[m]) for l, m in zip(label, mask))
    return total / count


def remap_checkpoint_keys(state_dict, own_keys):
    """PyG >= 2.4 parameter spelling -> the PyG 2.3 spelling this package keeps (pure renaming)."""
    own = set(own_keys)
    out = {}
    for k, v in state_dict.items():
        if "<" in k and "___" in k:  # ModuleDict key of HeteroConv: '<a___b___c>' -> 'a__b__c'
            head, rest = k.split("<", 1)
            name, tail = rest.split(">", 1)
            k = head + "__".join(name.split("___")) + tail
        if k not in own and k.endswith(".lin.weight"):  # GATConv(in_channels=int): lin -> lin_src (= lin_dst)
            base = k[: -len("lin.weight")]
            if base + "lin_src.weight" in own:
                out[base + "lin_src.weight"] = v
                if base + "lin_dst.weight" in own:
                    out[base + "lin_dst.weight"] = v
                continue
        out[k] = v
    return out


class _Base(nn.Module):
    """Shared pieces: activation choice (ELU for GAT*, ReLU otherwise -- section 9.8), dropout,
    loss, the per-batch plan cache, and the inference-mode guard."""

    conv_block: str
    dropout: float

    def _init_runtime(self):
        self._plans = hnn.PlanCache()

    def load_state_dict(self, state_dict, strict: bool = True, **kwargs):
        """Accepts checkpoints written by the reference under PyG 2.3 (this module's own spelling) and
        under PyG >= 2.4: hetero keys ``convs.<src___rel___dst>.`` and a GATConv
        with one shared projection stored as ``lin.weight`` instead of ``lin_src`` / ``lin_dst``."""
        return super().load_state_dict(remap_checkpoint_keys(state_dict, self.state_dict().keys()), strict=strict,
                                       **kwargs)

    @property
    def plan_cache(self) -> hnn.PlanCache:
        return self._plans

    def train(self, mode: bool = True):
        """``net.train()`` / ``net.eval()`` (base_training_job.py:203,248): besides the mode flag, drops the packed weight
        images, so the validation pass after an epoch of optimizer steps -- which update the weights in place, possibly
        inside a replayed CUDA graph -- folds the CURRENT weights."""
        if mode != self.training:
            hnn.invalidate_weight_caches(self)
        return super().train(mode)

    @property
    def _act_name(self) -> str:
        return "elu" if self.conv_block[:3] == "GAT" else "relu"

    def _drop(self, x):
        if self.training and self.dropout > 0:
            if isinstance(x, dict):
                return {k: F.dropout(v, p=self.dropout, training=True) for k, v in x.items()}
            return F.dropout(x, p=self.dropout, training=True)
        return x

    def loss(self, pred, label, mask=None):
        return cross_entropy_loss(pred, label, mask)

    def double(self):
        raise NotImplementedError("fp64 mode (base_training_job.py:50-51) is not supported by the sm_100a kernels")

    def forward(self, data):
        try:
            if self.training:
                return self._forward(data)
            with torch.no_grad():
                return self._forward(data)
        finally:
            self._plans.join()  # a side-stream plan that no kernel consumed
