"""Graph containers accepted by the models (PyG-shaped, no PyG dependency).

The reference feeds its models PyG ``Data`` / ``HeteroData`` / ``Batch`` objects
(base_training_job.py:164-178,205; bin/room_classification_server:286).  The models in
this package read them only through the attribute surface listed in SURVEY.md section 8b
(``x_dict``, ``edge_index_dict``, ``edge_attr_dict``, ``data[type].num_nodes``, ``x``,
``edge_index``, ``room_mask`` ...), so a real PyG object works unchanged.  PyG is not
installable on the target image, hence these minimal stand-ins with the same surface and
the same collate rule (``Batch.from_data_list``: concat on dim 0; ``*index*`` attributes
concatenated on the last dim after adding the running node count -- int64, bit-exact).
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Sequence, Tuple

import torch
from torch import Tensor

EdgeType = Tuple[str, str, str]


class Store(dict):
    """Attribute-style dict holding the tensors of one node type / relation."""

    __slots__ = ()

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        try:
            return self[name]
        except KeyError:
            raise AttributeError(name) from None

    def __setattr__(self, name, value):
        self[name] = value

    def __delattr__(self, name):
        del self[name]

    def apply_(self, fn):
        for k, v in self.items():
            if isinstance(v, Tensor):
                self[k] = fn(v)
        return self


class NodeStore(Store):
    __slots__ = ()

    @property
    def num_nodes(self):
        if "num_nodes" in self:
            return self["num_nodes"]
        for k in ("x", "pos", "batch", "y"):
            v = self.get(k)
            if isinstance(v, Tensor):
                return v.size(0)
        return 0

    @property
    def num_node_features(self):
        x = self.get("x")
        return 0 if x is None else (1 if x.dim() == 1 else x.size(-1))

    num_features = num_node_features


class EdgeStore(Store):
    __slots__ = ()

    @property
    def num_edges(self):
        ei = self.get("edge_index")
        return 0 if ei is None else ei.size(1)


def _is_index_key(key: str) -> bool:
    return "index" in key


class Data:
    """Homogeneous graph (``data.x``, ``data.edge_index``, ``data.room_mask`` ...)."""

    def __init__(self, **tensors):
        object.__setattr__(self, "_s", NodeStore())
        for k, v in tensors.items():
            if v is not None:
                self._s[k] = v

    def __getattr__(self, name):
        if name.startswith("__") or name == "_s":
            raise AttributeError(name)
        try:
            return self._s[name]
        except KeyError:
            raise AttributeError(name) from None

    def __setattr__(self, name, value):
        self._s[name] = value

    def __delattr__(self, name):
        del self._s[name]

    def __getitem__(self, name):
        return self._s[name]

    def __setitem__(self, name, value):
        self._s[name] = value

    def __contains__(self, name):
        return name in self._s

    def keys(self):
        return list(self._s.keys())

    def items(self):
        return list(self._s.items())

    @property
    def num_nodes(self):
        return self._s.num_nodes

    @property
    def num_edges(self):
        ei = self._s.get("edge_index")
        return 0 if ei is None else ei.size(1)

    @property
    def num_node_features(self):
        return self._s.num_node_features

    num_features = num_node_features

    def to(self, device, non_blocking: bool = False):
        self._s.apply_(lambda t: t.to(device, non_blocking=non_blocking))
        return self

    def pin_memory(self):
        self._s.apply_(lambda t: t.pin_memory())
        return self


class HeteroData:
    """Heterogeneous graph: ``data[node_type]`` / ``data[src, rel, dst]`` stores in
    insertion order, ``*_dict`` views collected on access."""

    def __init__(self):
        object.__setattr__(self, "_nodes", {})
        object.__setattr__(self, "_edges", {})
        object.__setattr__(self, "_g", Store())

    def __getitem__(self, key):
        if isinstance(key, tuple):
            if len(key) != 3:
                raise KeyError(f"edge types are (src, rel, dst) triplets, got {key!r}")
            return self._edges.setdefault(key, EdgeStore())
        return self._nodes.setdefault(key, NodeStore())

    def __getattr__(self, name):
        if name.startswith("__") or name in ("_nodes", "_edges", "_g"):
            raise AttributeError(name)
        if name.endswith("_dict"):
            return self.collect(name[:-5])
        try:
            return self._g[name]
        except KeyError:
            raise AttributeError(name) from None

    def __setattr__(self, name, value):
        self._g[name] = value

    def collect(self, key: str) -> Dict:
        out = {}
        for t, s in self._nodes.items():
            if key in s:
                out[t] = s[key]
        for t, s in self._edges.items():
            if key in s:
                out[t] = s[key]
        return out

    @property
    def node_types(self) -> List[str]:
        return list(self._nodes)

    @property
    def edge_types(self) -> List[EdgeType]:
        return list(self._edges)

    def _stores(self) -> Iterable[Store]:
        yield from self._nodes.values()
        yield from self._edges.values()

    def to(self, device, non_blocking: bool = False):
        for s in self._stores():
            s.apply_(lambda t: t.to(device, non_blocking=non_blocking))
        return self

    def pin_memory(self):
        for s in self._stores():
            s.apply_(lambda t: t.pin_memory())
        return self

    def nbytes(self) -> int:
        return sum(v.numel() * v.element_size() for s in self._stores() for v in s.values() if isinstance(v, Tensor))


class Batch:
    """Disjoint-union collate (PyG ``Batch.from_data_list`` semantics)."""

    @staticmethod
    def from_data_list(graphs: Sequence, device=None):
        """``device=None``: the CPU collate (torch.cat + per-graph offset adds, as PyG does it).
        ``device="cuda"``: the GPU collate -- every attribute is concatenated UN-offset on the host
        (one memcpy per graph into a pinned buffer), copied once, and the node offsets are added to
        the index attributes by ``hg_collate_offsets`` on the device; ``batch`` comes from
        ``hg_batch_vector``.  Bit-identical to ``from_data_list(graphs).to(device)``."""
        if len(graphs) == 0:
            raise ValueError("empty batch")
        if device is not None and torch.device(device).type == "cuda":
            return Batch._device(graphs, torch.device(device))
        if isinstance(graphs[0], HeteroData):
            return Batch._hetero(graphs)
        return Batch._homo(graphs)

    @staticmethod
    def _device(graphs: Sequence, device) -> "Data | HeteroData":
        from . import ops

        G = len(graphs)

        def up(parts, dim):  # host concatenation into pinned memory, one asynchronous H2D copy
            shape = list(parts[0].shape)
            shape[dim] = sum(p.size(dim) for p in parts)
            host = torch.empty(shape, dtype=parts[0].dtype, pin_memory=True)
            torch.cat(parts, dim=dim, out=host)
            return host.to(device, non_blocking=True)

        def prefix(counts):  # [G + 1] exclusive prefix sums, int64, on host and device
            p = torch.zeros(G + 1, dtype=torch.long)
            p[1:] = torch.tensor(counts, dtype=torch.long).cumsum(0)
            return p, p.to(device, non_blocking=True)

        hetero = isinstance(graphs[0], HeteroData)
        first = graphs[0]
        out = HeteroData() if hetero else Data()
        if hetero:
            ptrs = {}
            for nt in first.node_types:
                ptr, ptr_d = prefix([g[nt].num_nodes for g in graphs])
                ptrs[nt] = ptr_d
                for key, v in first[nt].items():
                    if isinstance(v, Tensor):
                        out[nt][key] = up([g[nt][key] for g in graphs], -1 if _is_index_key(key) else 0)
                if "x" not in out[nt]:
                    out[nt].num_nodes = int(ptr[-1])
                out[nt].batch = ops.batch_vector(ptr_d, int(ptr[-1]))
                out[nt].ptr = ptr_d
            for et in first.edge_types:
                src, _, dst = et
                for key, v in first[et].items():
                    if not isinstance(v, Tensor):
                        continue
                    parts = [g[et][key] for g in graphs]
                    if _is_index_key(key):
                        _, eptr_d = prefix([p.size(-1) for p in parts])
                        out[et][key] = ops.collate_offsets_(up(parts, -1), eptr_d, ptrs[src], ptrs[dst])
                    else:
                        out[et][key] = up(parts, 0)
        else:
            ptr, ptr_d = prefix([g.num_nodes for g in graphs])
            for key in first.keys():
                parts = [g[key] for g in graphs]
                if not isinstance(parts[0], Tensor):
                    continue
                if _is_index_key(key):
                    _, eptr_d = prefix([p.size(-1) for p in parts])
                    out[key] = ops.collate_offsets_(up(parts, -1), eptr_d, ptr_d, ptr_d)
                else:
                    out[key] = up(parts, 0)
            out.batch = ops.batch_vector(ptr_d, int(ptr[-1]))
            out.ptr = ptr_d
        out.num_graphs = G
        return out

    @staticmethod
    def _homo(graphs: Sequence[Data]) -> Data:
        counts = torch.tensor([g.num_nodes for g in graphs], dtype=torch.long)
        ptr = torch.zeros(len(graphs) + 1, dtype=torch.long)
        ptr[1:] = counts.cumsum(0)
        out = Data()
        for key in graphs[0].keys():
            parts = [g[key] for g in graphs]
            if not isinstance(parts[0], Tensor):
                continue
            if _is_index_key(key):
                parts = [p + int(off) for p, off in zip(parts, ptr[:-1])]
                out[key] = torch.cat(parts, dim=-1)
            else:
                out[key] = torch.cat(parts, dim=0)
        out.batch = torch.repeat_interleave(torch.arange(len(graphs)), counts)
        out.ptr = ptr
        out.num_graphs = len(graphs)
        return out

    @staticmethod
    def _hetero(graphs: Sequence[HeteroData]) -> HeteroData:
        first = graphs[0]
        out = HeteroData()
        ptrs = {}
        for nt in first.node_types:
            counts = torch.tensor([g[nt].num_nodes for g in graphs], dtype=torch.long)
            ptr = torch.zeros(len(graphs) + 1, dtype=torch.long)
            ptr[1:] = counts.cumsum(0)
            ptrs[nt] = ptr
            for key, v in first[nt].items():
                if isinstance(v, Tensor):
                    out[nt][key] = torch.cat([g[nt][key] for g in graphs], dim=-1 if _is_index_key(key) else 0)
            if "x" not in out[nt]:
                out[nt].num_nodes = int(ptr[-1])
            out[nt].batch = torch.repeat_interleave(torch.arange(len(graphs)), counts)
            out[nt].ptr = ptr
        for et in first.edge_types:
            src, _, dst = et
            for key, v in first[et].items():
                if not isinstance(v, Tensor):
                    continue
                parts = [g[et][key] for g in graphs]
                if _is_index_key(key):
                    parts = [p + torch.tensor([[int(a)], [int(b)]], dtype=p.dtype)
                             for p, a, b in zip(parts, ptrs[src][:-1], ptrs[dst][:-1])]
                    out[et][key] = torch.cat(parts, dim=-1)
                else:
                    out[et][key] = torch.cat(parts, dim=0)
        out.num_graphs = len(graphs)
        return out
