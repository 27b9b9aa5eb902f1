"""Shared definitions of the golden cases: model kwargs, seeded inputs, fixture (de)serialisation.

Imported by ``make_golden.py`` (build container: runs the REFERENCE's own files) and by the
tests (anywhere: compare oracle / product against the committed outputs).  Nothing here reads
/root/reference.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from hydra_gnn_b200 import synth
from hydra_gnn_b200.data import Batch, HeteroData
from hydra_gnn_b200.schema import (EDGE_TYPES, HTREE_EDGE_TYPES, HTREE_INIT_EDGE_TYPES, HTREE_NODE_TYPES,
                                   HTREE_POOL_EDGE_TYPES, HTREE_VIRTUAL_NODE_TYPES)

HERE = os.path.dirname(os.path.abspath(__file__))
HTREE_ALL_NODE_TYPES = HTREE_NODE_TYPES + HTREE_VIRTUAL_NODE_TYPES
HTREE_ALL_EDGE_TYPES = HTREE_EDGE_TYPES + HTREE_POOL_EDGE_TYPES + HTREE_INIT_EDGE_TYPES

# synthetic graph ids whose H-trees are committed (small ones: <= 6 rooms)
HTREE_SYNTH_IDS = [0, 1, 2, 4, 5, 6, 8, 10]
HTREE_STANFORD_IDS = list(range(0, 482, 20))

GAT3 = dict(GAT_hidden_dims=[64, 64], GAT_heads=[3, 3, 3], GAT_concats=[False, False, False])
HT_DIMS = {"object": 306, "room": 6, "object-room": 6, "room-room": 6, "object_virtual": 306, "room_virtual": 6}
HT_DIMS_ST = {k: 6 for k in HT_DIMS}

# name -> (class name, kwargs, input name)
CASES = {
    "mgat": ("HeterogeneousNetwork", dict(input_dim_dict={"objects": 306, "rooms": 6}, output_dim=26, conv_block="GAT",
                                          dropout=0.4, **GAT3), "synth16"),
    "mgat_edge": ("HeterogeneousNetwork", dict(input_dim_dict={"objects": 303, "rooms": 3}, output_dim=26,
                                               conv_block="GAT_edge", dropout=0.4, **GAT3), "synth16_relpos"),
    "msage_all": ("HeterogeneousNetwork", dict(input_dim_dict={"objects": 6, "rooms": 6},
                                               output_dim_dict={"rooms": 15, "objects": 35}, conv_block="GraphSAGE",
                                               hidden_dim=128, num_layers=3, dropout=0.25), "stanford"),
    "nt_het_gat": ("HeterogeneousNeuralTreeNetwork", dict(input_dim_dict=HT_DIMS, output_dim=26, conv_block="GAT",
                                                          disable_initialization=True, **GAT3), "htree_synth"),
    "nt_het_gat_init": ("HeterogeneousNeuralTreeNetwork", dict(input_dim_dict=HT_DIMS, output_dim=26, conv_block="GAT",
                                                               disable_initialization=False, **GAT3), "htree_synth"),
    "nt_het_sage_all": ("HeterogeneousNeuralTreeNetwork",
                        dict(input_dim_dict=HT_DIMS_ST, output_dim_dict={"room": 15, "object": 35, "object-room": 64,
                                                                         "room-room": 64},
                             conv_block="GraphSAGE", hidden_dim=64, num_layers=3, disable_initialization=False),
                        "htree_stanford"),
    "nt_hom_gcn": ("HomogeneousNeuralTreeNetwork", dict(input_dim=306, output_dim=26, conv_block="GCN", hidden_dim=128,
                                                        num_layers=4, disable_initialization=True), "htree_synth_hom"),
    "nt_hom_sage": ("HomogeneousNeuralTreeNetwork", dict(input_dim=306, output_dim=26, conv_block="GraphSAGE",
                                                         hidden_dim=128, num_layers=4, disable_initialization=True),
                    "htree_synth_hom"),
    "nt_hom_gcn_init": ("HomogeneousNeuralTreeNetwork", dict(input_dim=306, output_dim=26, conv_block="GCN",
                                                             hidden_dim=64, num_layers=3, disable_initialization=False),
                        "htree_synth_hom"),
    "nt_hom_gat_all": ("HomogeneousNeuralTreeNetwork",
                       dict(input_dim=6, output_dim_dict={"room": 15, "object": 35}, conv_block="GAT",
                            GAT_hidden_dims=[32, 32], GAT_heads=[2, 2], GAT_concats=[True, True],
                            disable_initialization=False), "htree_stanford_hom"),
    "hom_gcn": ("HomogeneousNetwork", dict(input_dim=306, output_dim=26, conv_block="GCN", hidden_dim=64, num_layers=3),
                "synth16_hom"),
    "hom_gin": ("HomogeneousNetwork", dict(input_dim=306, output_dim=26, conv_block="GIN", hidden_dim=64, num_layers=3),
                "synth16_hom"),
    "hom_sage_all": ("HomogeneousNetwork", dict(input_dim=306, output_dim_dict={"rooms": 26, "objects": 28},
                                                conv_block="GraphSAGE", hidden_dim=64, num_layers=3), "synth16_hom"),
    "hom_gat": ("HomogeneousNetwork", dict(input_dim=306, output_dim=26, conv_block="GAT", GAT_hidden_dims=[64, 64],
                                           GAT_heads=[6, 6, 6], GAT_concats=[True, True, False]), "synth16_hom"),
    "hom_gat_edge": ("HomogeneousNetwork", dict(input_dim=303, output_dim=26, conv_block="GAT_edge",
                                                GAT_hidden_dims=[64], GAT_heads=[3, 3], GAT_concats=[True, False]),
                     "synth16_relpos_hom"),
}


# ------------------------------------------------------------------------------ fixtures on disk
def load_stanford():
    """The 482 graphs of the reference's bundled data/Stanford3DSG.pkl, converted by make_golden.py to
    flat arrays (the pkl itself is not on the GPU box).  Returns a list of per-graph dicts."""
    z = np.load(os.path.join(HERE, "stanford3dsg.npz"))
    nptr, eptr = z["node_ptr"], z["edge_ptr"]
    out = []
    for g in range(len(nptr) - 1):
        n0, n1, e0, e1 = nptr[g], nptr[g + 1], eptr[g], eptr[g + 1]
        out.append(dict(x=z["x"][n0:n1], y=z["y"][n0:n1], room_mask=z["room_mask"][n0:n1],
                        edge_index=z["edge_index"][:, e0:e1]))
    return out


def stanford_hetero(g) -> HeteroData:
    """Room-object HeteroData of one preprocessed Stanford3DSG graph: restates
    Stanford3DSG_dataset.py:236-287 (room = node 0, objects = the rest, object-object edges kept,
    object-room edges = one per object) + spark_dsg ``to_torch`` (objects_to_objects
    holds both directions, rooms_to_objects = room -> each object) + ``fill_missing_edge_index``
    (mp3d_dataset.py:221-255: objects_to_rooms = flip, rooms_to_rooms empty [2, 0])."""
    x, y, ei = torch.from_numpy(g["x"]), torch.from_numpy(g["y"]), g["edge_index"]
    assert bool(g["room_mask"][0]) and int(g["room_mask"].sum()) == 1
    n_obj = x.size(0) - 1
    pairs = set()
    for i, j in ei.T.tolist():
        if i == 0 or j == 0 or i == j:
            continue
        pairs.add((min(i, j) - 1, max(i, j) - 1))
    pairs = sorted(pairs)
    oo = torch.tensor([[a for a, b in pairs] + [b for a, b in pairs], [b for a, b in pairs] + [a for a, b in pairs]],
                      dtype=torch.long).reshape(2, -1)
    r2o = torch.stack([torch.zeros(n_obj, dtype=torch.long), torch.arange(n_obj)])
    d = HeteroData()
    d["objects"].x, d["objects"].pos, d["objects"].y = x[1:].float(), x[1:, :3].float(), y[1:].long()
    d["rooms"].x, d["rooms"].pos, d["rooms"].y = x[:1].float(), x[:1, :3].float(), y[:1].long()
    d["objects", "objects_to_objects", "objects"].edge_index = oo
    d["rooms", "rooms_to_rooms", "rooms"].edge_index = torch.zeros(2, 0, dtype=torch.long)
    d["objects", "objects_to_rooms", "rooms"].edge_index = r2o.flip(0)
    d["rooms", "rooms_to_objects", "objects"].edge_index = r2o
    return d


def save_htrees(path, trees, wide_object: bool):
    """trees: list of HeteroData-like (reference output).  Object features beyond the first 6 columns
    are NOT stored: they are word2vec_table()[label] by construction (checked by the caller)."""
    arrs = {"n_graphs": np.int64(len(trees)), "wide_object": np.int64(wide_object)}
    for k, t in enumerate(trees):
        for nt in HTREE_ALL_NODE_TYPES:
            arrs[f"g{k}|n|{nt}|x6"] = t[nt].x[:, :6].numpy().astype(np.float32)
            arrs[f"g{k}|n|{nt}|label"] = t[nt].label.numpy().astype(np.int32)
            arrs[f"g{k}|n|{nt}|clique_has"] = t[nt].clique_has.numpy().astype(np.int32)
        for et in HTREE_ALL_EDGE_TYPES:
            arrs[f"g{k}|e|{et[1]}"] = t[et].edge_index.numpy().astype(np.int32)
    np.savez_compressed(path, **arrs)


def load_htrees(name: str):
    """Per-graph augmented H-trees (HeteroData) as produced by the reference's
    generate_htree -> add_virtual_nodes_to_htree -> nx_htree_to_torch -> feature truncation
    (mp3d_dataset.py:439-458), committed by make_golden.py."""
    z = np.load(os.path.join(HERE, name + ".npz"))
    wide = bool(z["wide_object"])
    table = torch.from_numpy(synth.word2vec_table()) if wide else None
    out = []
    for k in range(int(z["n_graphs"])):
        d = HeteroData()
        for nt in HTREE_ALL_NODE_TYPES:
            x6 = torch.from_numpy(z[f"g{k}|n|{nt}|x6"])
            label = torch.from_numpy(z[f"g{k}|n|{nt}|label"])
            if wide and nt in ("object", "object_virtual"):
                x6 = torch.cat([x6, table[label.long()]], dim=1)
            d[nt].x, d[nt].label = x6, label
            d[nt].pos = x6[:, :3].clone()
            d[nt].clique_has = torch.from_numpy(z[f"g{k}|n|{nt}|clique_has"])
        for et in HTREE_ALL_EDGE_TYPES:
            d[et].edge_index = torch.from_numpy(z[f"g{k}|e|{et[1]}"]).long()
        out.append(d)
    return out


# ------------------------------------------------------------------------------ seeded inputs
def _per_graph_hom(graphs, fn):
    return Batch.from_data_list([fn(g) for g in graphs])


def build_input(name: str):
    """The input batch of a golden case (product-side containers; also consumed by the oracle and,
    in make_golden.py, by the reference's own model classes)."""
    from hydra_gnn_b200 import convert

    if name == "synth16":
        return synth.synth_mp3d_batch(range(16))
    if name == "synth16_relpos":
        return synth.synth_mp3d_batch(range(16), relative_pos=True)
    if name == "synth16_hom":
        return _per_graph_hom([synth.synth_mp3d_batch([g]) for g in range(16)], convert.dsg_to_homogeneous)
    if name == "synth16_relpos_hom":
        return _per_graph_hom([synth.synth_mp3d_batch([g], relative_pos=True) for g in range(16)],
                              convert.dsg_to_homogeneous)
    if name == "stanford":
        return Batch.from_data_list([stanford_hetero(g) for g in load_stanford()])
    if name == "htree_synth":
        return Batch.from_data_list(load_htrees("htree_synth"))
    if name == "htree_stanford":
        return Batch.from_data_list(load_htrees("htree_stanford"))
    if name == "htree_synth_hom":
        return _per_graph_hom(load_htrees("htree_synth"), convert.heterogeneous_htree_to_homogeneous)
    if name == "htree_stanford_hom":
        return _per_graph_hom(load_htrees("htree_stanford"), convert.heterogeneous_htree_to_homogeneous)
    raise KeyError(name)


def build_model(namespace, case: str, seed: int = 0):
    """``namespace`` = a module exposing the four model classes (reference / oracle / product)."""
    cls, kw, _ = CASES[case]
    model = getattr(namespace, cls)(**kw)
    synth.random_state_dict_(model, seed=seed)
    return model.eval()


def load_golden_logits():
    z = np.load(os.path.join(HERE, "logits.npz"))
    out = {}
    for case in CASES:
        keys = sorted(k for k in z.files if k.split("|")[0] == case)
        out[case] = [torch.from_numpy(z[k]) for k in keys]
    return out


# ------------------------------------------------------------------------------ gradient fingerprints (training parity)
def grad_fingerprint(model, inp):
    """Three float64 numbers per parameter -- sum, l2 norm and the dot product with a fixed pseudo-random vector -- of the
    gradient of a fixed linear functional of the model's outputs (eval mode: dropout off, so the forward is
    deterministic; autograd does not care).  Small enough to commit, sharp enough to pin a backward."""
    import zlib

    model.zero_grad(set_to_none=True)
    out = model(inp)
    outs = list(out) if isinstance(out, (tuple, list)) else [out]
    g = torch.Generator().manual_seed(1234)
    sum((o * torch.randn(o.shape, generator=g).to(o.device)).sum() for o in outs).backward()
    fp = {}
    for name, p in sorted(model.named_parameters()):
        if p.grad is None:
            fp[name] = np.full(3, np.nan)
            continue
        gr = p.grad.detach().double().cpu().reshape(-1)
        probe = torch.randn(gr.numel(), generator=torch.Generator().manual_seed(zlib.crc32(name.encode())),
                            dtype=torch.float64)
        fp[name] = np.array([float(gr.sum()), float(gr.norm()), float(gr.dot(probe))])
    return fp


def load_golden_grads():
    path = os.path.join(HERE, "grads.npz")
    if not os.path.exists(path):
        return {}
    z = np.load(path)
    out = {}
    for k in z.files:
        case, name = k.split("|", 1)
        out.setdefault(case, {})[name] = z[k]
    return out
