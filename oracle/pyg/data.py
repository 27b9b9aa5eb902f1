"""Restatement of the ``torch_geometric.data`` containers on the reference's path.

ORACLE -- test infrastructure only (see ``oracle/__init__.py``).

``Data`` / ``HeteroData`` / ``Batch`` as used by the reference: models read
``data.x_dict / edge_index_dict / edge_attr_dict / data[type].num_nodes`` and
``data.x / edge_index / room_mask / ...``; the dataset code calls
``HeteroData.to_homogeneous`` (mp3d_dataset.py:54,60), ``Data.to_heterogeneous``
(neural_tree/construct.py:475-480) and the ``DataLoader`` collates with
``Batch.from_data_list`` (base_training_job.py:164-178).  Semantics follow the
published PyG 2.3.1 behaviour:

* collate: every tensor attribute is concatenated on dim 0, except attributes whose
  name contains ``"index"`` which are concatenated on dim -1 after adding the running
  node count (per src / dst node type for hetero);
* ``to_homogeneous``: nodes concatenated in ``node_types`` order, edges in
  ``edge_types`` order with per-type node offsets, ``node_type`` / ``edge_type`` added;
* ``to_heterogeneous``: order-preserving masking + local re-indexing per type.
"""
from __future__ import annotations

import copy
from typing import Any, Dict, List, Tuple

import torch
from torch import Tensor

_N_KEYS = ("x", "feat", "pos", "batch", "node_type", "n_id")


def _cat_dim(key: str) -> int:
    return -1 if ("index" in key or key == "face") else 0


class _Storage(dict):
    """Attribute-style dict (PyG ``BaseStorage``)."""

    def __getattr__(self, key):
        if key.startswith("__"):
            raise AttributeError(key)
        try:
            return self[key]
        except KeyError:
            raise AttributeError(key) from None

    def __setattr__(self, key, value):
        self[key] = value

    def __delattr__(self, key):
        del self[key]

    def tensor_items(self):
        return [(k, v) for k, v in self.items() if isinstance(v, Tensor)]


class NodeStorage(_Storage):
    @property
    def num_nodes(self):
        if "num_nodes" in self:
            return self["num_nodes"]
        for k in _N_KEYS:
            if k in self and isinstance(self[k], Tensor):
                return self[k].size(0)
        for k, v in self.items():
            if isinstance(v, Tensor) and v.dim() >= 1 and "index" not in k and "edge" not in k:
                return v.size(0)
        return None

    @property
    def num_node_features(self):
        x = self.get("x", None)
        if x is None:
            return 0
        return 1 if x.dim() == 1 else x.size(-1)

    num_features = num_node_features


class EdgeStorage(_Storage):
    @property
    def num_edges(self):
        ei = self.get("edge_index", None)
        return 0 if ei is None else ei.size(1)


class Data:
    """Homogeneous graph: a flat attribute store."""

    def __init__(self, x=None, edge_index=None, edge_attr=None, y=None, pos=None, **kwargs):
        object.__setattr__(self, "_store", NodeStorage())
        for k, v in dict(x=x, edge_index=edge_index, edge_attr=edge_attr, y=y, pos=pos, **kwargs).items():
            if v is not None:
                self._store[k] = v

    def __getattr__(self, key):
        if key.startswith("__") or key == "_store":
            raise AttributeError(key)
        try:
            return self._store[key]
        except KeyError:
            raise AttributeError(key) from None

    def __setattr__(self, key, value):
        self._store[key] = value

    def __delattr__(self, key):
        del self._store[key]

    def __getitem__(self, key):
        return self._store[key]

    def __setitem__(self, key, value):
        self._store[key] = value

    def __contains__(self, key):
        return key in self._store

    def keys(self):
        return list(self._store.keys())

    def items(self):
        return list(self._store.items())

    @property
    def num_nodes(self):
        return NodeStorage.num_nodes.fget(self._store)

    @property
    def num_edges(self):
        ei = self._store.get("edge_index", None)
        return 0 if ei is None else ei.size(1)

    @property
    def num_node_features(self):
        return NodeStorage.num_node_features.fget(self._store)

    num_features = num_node_features

    def is_node_attr(self, key):
        v = self._store[key]
        return (isinstance(v, Tensor) and v.dim() >= 1 and "edge" not in key and "index" not in key
                and v.size(_cat_dim(key)) == self.num_nodes)

    def is_edge_attr(self, key):
        v = self._store[key]
        if key == "edge_index" or not isinstance(v, Tensor) or v.dim() < 1:
            return False
        if self.is_node_attr(key):
            return False
        return "index" not in key and v.size(0) == self.num_edges

    def to(self, device):
        for k, v in self._store.items():
            if isinstance(v, Tensor):
                self._store[k] = v.to(device)
        return self

    def clone(self):
        out = Data()
        for k, v in self._store.items():
            out._store[k] = v.clone() if isinstance(v, Tensor) else copy.deepcopy(v)
        return out

    def to_heterogeneous(self, node_type=None, edge_type=None, node_type_names=None, edge_type_names=None):
        node_type = self._store["node_type"] if node_type is None else node_type
        edge_type = self._store["edge_type"] if edge_type is None else edge_type
        node_ids, index_map = {}, torch.empty_like(node_type, dtype=torch.long)
        for i, _ in enumerate(node_type_names):
            node_ids[i] = (node_type == i).nonzero(as_tuple=False).view(-1)
            index_map[node_ids[i]] = torch.arange(len(node_ids[i]))
        edge_ids = {i: (edge_type == i).nonzero(as_tuple=False).view(-1) for i, _ in enumerate(edge_type_names)}
        data = HeteroData()
        for i, key in enumerate(node_type_names):
            for attr, value in self.items():
                if attr in ("node_type", "edge_type", "ptr"):
                    continue
                if isinstance(value, Tensor) and self.is_node_attr(attr):
                    data[key][attr] = value.index_select(0, node_ids[i])
            if len(data[key]) == 0:
                data[key].num_nodes = node_ids[i].size(0)
        for i, key in enumerate(edge_type_names):
            for attr, value in self.items():
                if attr in ("node_type", "edge_type", "ptr"):
                    continue
                if attr == "edge_index":
                    ei = value[:, edge_ids[i]].clone()
                    ei[0] = index_map[ei[0]]
                    ei[1] = index_map[ei[1]]
                    data[key].edge_index = ei
                elif isinstance(value, Tensor) and self.is_edge_attr(attr):
                    data[key][attr] = value.index_select(0, edge_ids[i])
        return data


class HeteroData:
    """Heterogeneous graph: one ``NodeStorage`` per node type, one ``EdgeStorage`` per
    ``(src, rel, dst)``; stores keep insertion order (this order is what
    ``to_homogeneous`` and ``HeteroConv`` iterate in)."""

    def __init__(self):
        object.__setattr__(self, "_node_stores", {})
        object.__setattr__(self, "_edge_stores", {})
        object.__setattr__(self, "_global", _Storage())

    def __getitem__(self, key):
        if isinstance(key, tuple):
            assert len(key) == 3, "edge types are (src, rel, dst) triplets"
            if key not in self._edge_stores:
                self._edge_stores[key] = EdgeStorage()
            return self._edge_stores[key]
        if key not in self._node_stores:
            self._node_stores[key] = NodeStorage()
        return self._node_stores[key]

    def __getattr__(self, key):
        if key.startswith("__") or key in ("_node_stores", "_edge_stores", "_global"):
            raise AttributeError(key)
        if key.endswith("_dict"):
            return self.collect(key[:-5])
        try:
            return self._global[key]
        except KeyError:
            raise AttributeError(key) from None

    def __setattr__(self, key, value):
        self._global[key] = value

    def collect(self, key) -> Dict[Any, Any]:
        out = {}
        for t, s in list(self._node_stores.items()) + list(self._edge_stores.items()):
            if key in s:
                out[t] = s[key]
        return out

    @property
    def node_types(self) -> List[str]:
        return list(self._node_stores.keys())

    @property
    def edge_types(self) -> List[Tuple[str, str, str]]:
        return list(self._edge_stores.keys())

    @property
    def node_stores(self):
        return list(self._node_stores.values())

    @property
    def edge_stores(self):
        return list(self._edge_stores.values())

    @property
    def num_nodes(self):
        return sum(s.num_nodes or 0 for s in self._node_stores.values())

    def to(self, device):
        for s in self.node_stores + self.edge_stores:
            for k, v in s.items():
                if isinstance(v, Tensor):
                    s[k] = v.to(device)
        return self

    def clone(self):
        out = HeteroData()
        for t, s in self._node_stores.items():
            for k, v in s.items():
                out[t][k] = v.clone() if isinstance(v, Tensor) else copy.deepcopy(v)
        for t, s in self._edge_stores.items():
            for k, v in s.items():
                out[t][k] = v.clone() if isinstance(v, Tensor) else copy.deepcopy(v)
        return out

    def to_homogeneous(self, node_attrs=None, edge_attrs=None, add_node_type=True, add_edge_type=True):
        def consistent(stores):
            sizes: Dict[str, list] = {}
            for s in stores:
                for k, v in s.items():
                    if k == "edge_index" or not isinstance(v, Tensor) or v.dim() < 1:
                        continue
                    sizes.setdefault(k, []).append(tuple(v.shape[1:]))
            return [k for k, sz in sizes.items() if len(sz) == len(stores) and len(set(sz)) == 1]

        offsets, cum = {}, 0
        for t, s in self._node_stores.items():
            offsets[t] = cum
            cum += s.num_nodes
        eis = []
        for (src, _, dst), s in self._edge_stores.items():
            ei = s["edge_index"].clone().long()
            ei[0] += offsets[src]
            ei[1] += offsets[dst]
            eis.append(ei)
        data = Data(**dict(self._global))
        if eis:
            data.edge_index = torch.cat(eis, dim=1)
        for k in (node_attrs if node_attrs is not None else consistent(self.node_stores)):
            vals = [s[k] for s in self.node_stores]
            data[k] = torch.cat(vals, 0) if len(vals) > 1 else vals[0]
        if data.num_nodes is None:
            data.num_nodes = cum
        for k in (edge_attrs if edge_attrs is not None else consistent(self.edge_stores)):
            vals = [s[k] for s in self.edge_stores]
            data[k] = torch.cat(vals, 0) if len(vals) > 1 else vals[0]
        if add_node_type:
            sizes = torch.tensor([s.num_nodes for s in self.node_stores], dtype=torch.long)
            data.node_type = torch.arange(len(sizes)).repeat_interleave(sizes)
        if add_edge_type and eis:
            sizes = torch.tensor([e.size(1) for e in eis], dtype=torch.long)
            data.edge_type = torch.arange(len(sizes)).repeat_interleave(sizes)
        return data


class Batch:
    """``Batch.from_data_list``: disjoint union."""

    @staticmethod
    def from_data_list(data_list):
        first = data_list[0]
        if isinstance(first, HeteroData):
            out = HeteroData()
            cum = {t: 0 for t in first.node_types}
            node_vals = {t: {} for t in first.node_types}
            edge_vals = {t: {} for t in first.edge_types}
            batch_vec = {t: [] for t in first.node_types}
            ptr = {t: [0] for t in first.node_types}
            for g, d in enumerate(data_list):
                for t in first.edge_types:
                    src, _, dst = t
                    for k, v in d[t].items():
                        if not isinstance(v, Tensor):
                            continue
                        if "index" in k:
                            inc = torch.tensor([[cum[src]], [cum[dst]]], dtype=v.dtype)
                            v = v + inc
                        edge_vals[t].setdefault(k, []).append(v)
                for t in first.node_types:
                    s = d[t]
                    n = s.num_nodes
                    for k, v in s.items():
                        if isinstance(v, Tensor):
                            node_vals[t].setdefault(k, []).append(v)
                    batch_vec[t].append(torch.full((n,), g, dtype=torch.long))
                    cum[t] += n
                    ptr[t].append(cum[t])
            for t in first.node_types:
                for k, vs in node_vals[t].items():
                    out[t][k] = torch.cat(vs, _cat_dim(k))
                if "x" not in out[t]:
                    out[t].num_nodes = cum[t]
                out[t].batch = torch.cat(batch_vec[t])
                out[t].ptr = torch.tensor(ptr[t], dtype=torch.long)
            for t in first.edge_types:
                for k, vs in edge_vals[t].items():
                    out[t][k] = torch.cat(vs, _cat_dim(k))
            out.num_graphs = len(data_list)
            return out
        out = Data()
        cum, vals, batch_vec, ptr = 0, {}, [], [0]
        for g, d in enumerate(data_list):
            n = d.num_nodes
            for k, v in d.items():
                if not isinstance(v, Tensor):
                    continue
                if "index" in k:
                    v = v + cum
                vals.setdefault(k, []).append(v)
            batch_vec.append(torch.full((n,), g, dtype=torch.long))
            cum += n
            ptr.append(cum)
        for k, vs in vals.items():
            out[k] = torch.cat(vs, _cat_dim(k))
        out.batch = torch.cat(batch_vec)
        out.ptr = torch.tensor(ptr, dtype=torch.long)
        out.num_graphs = len(data_list)
        return out
