#!/usr/bin/env python
"""Generates the committed fixtures of tests/golden/ by running the REFERENCE's own code.

Run in the build container only (needs /root/reference; the GPU box never has it):

    python tests/golden/make_golden.py

What runs from the reference, unmodified, imported from where it lies:
  * hydra_gnn/neural_tree/{construct,generate_junction_tree_hierarchies}.py -- the networkx H-tree
    construction (generate_htree -> add_virtual_nodes_to_htree -> nx_htree_to_torch), on synthetic
    MP3D-shaped graphs and on graphs of the bundled data/Stanford3DSG.pkl;
  * hydra_gnn/mp3d_dataset.py -- heterogeneous_data_to_homogeneous / heterogeneous_htree_to_homogeneous;
  * hydra_gnn/models/*.py -- the four model classes (control flow, layer wiring, slicing).
torch_geometric is not installable here, so those files run on ``oracle.pyg`` (the restatement of
the PyG 2.3.1 operators, registered under the name ``torch_geometric``); if the real package ever
is importable it is used instead and ``meta.json`` records ``pyg-<version>``.

Outputs: stanford3dsg.npz, htree_synth.npz, htree_stanford.npz, convert_golden.npz, logits.npz, grads.npz,
htree_sha256.json, meta.json.
"""
from __future__ import annotations

import hashlib
import json
import os
import pickle
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)

from oracle import ref_import  # noqa: E402

ref_models, ref_construct, ref_mp3d, ORACLE_KIND = ref_import.load()

import cases  # noqa: E402
from hydra_gnn_b200 import synth  # noqa: E402
from oracle.pyg import data as od  # noqa: E402  (the container type the reference code expects)


def to_ref_hetero(g, pad_to=None, with_label=True):
    """product HeteroData -> the (oracle.pyg / PyG) HeteroData the reference's H-tree code consumes."""
    h = od.HeteroData()
    for nt in g.node_types:
        x = g[nt].x
        if pad_to is not None and x.size(1) < pad_to:
            x = torch.cat([x, torch.zeros(x.size(0), pad_to - x.size(1))], dim=1)
        h[nt].x = x
        for k in ("pos", "label", "clique_has", "y"):
            if k in g[nt]:
                h[nt][k] = g[nt][k]
        if with_label and "label" not in g[nt] and "y" in g[nt]:
            h[nt].label = g[nt].y.int()
    for et in g.edge_types:
        h[et].edge_index = g[et].edge_index
        if "edge_attr" in g[et]:
            h[et].edge_attr = g[et].edge_attr
    return h


def reference_htree(g, pad_to, trunc):
    """mp3d_dataset.py:437-458 (Hydra_mp3d_htree_data.compute_torch_data) minus the spark_dsg part."""
    h = to_ref_hetero(g, pad_to=pad_to)
    htree_nx = ref_construct.generate_htree(h, verbose=False)
    aug = ref_construct.add_virtual_nodes_to_htree(htree_nx)
    t = ref_construct.nx_htree_to_torch(aug)
    if trunc:
        for nt in ("room", "room_virtual", "object-room", "room-room"):
            t[nt].x = t[nt].x[:, :-trunc]
    sig = repr([(n, d["node_type"], d["clique_has"]) for n, d in aug.nodes.items()]) + repr(list(aug.edges))
    return t, sig


def stanford_nx_signature(graphs):
    """SURVEY.md section 0 / 8c golden hash: H-trees of the 482 Stanford3DSG graphs built from an
    nx.Graph with objects first and the room last (the to_homogeneous() order)."""
    import networkx as nx

    sha = hashlib.sha256()
    for g in graphs:
        n = g["x"].shape[0]
        nxg = nx.Graph()
        for k in range(1, n):
            nxg.add_node(k - 1, x=g["x"][k].tolist(), pos=g["x"][k, :3].tolist(), label=int(g["y"][k]), node_type="object")
        nxg.add_node(n - 1, x=g["x"][0].tolist(), pos=g["x"][0, :3].tolist(), label=int(g["y"][0]), node_type="room")
        for i, j in g["edge_index"].T.tolist():
            a = n - 1 if i == 0 else i - 1
            b = n - 1 if j == 0 else j - 1
            if a != b:
                nxg.add_edge(a, b)
        for k in range(n - 1):
            nxg.add_edge(k, n - 1)
        sha.update(_htree_sig_from_nx(nxg).encode())
    return sha.hexdigest()


def _htree_sig_from_nx(dsg_nx):
    """Body of construct.generate_htree (construct.py:249-310) for an already-built nx graph."""
    import networkx as nx

    c = ref_construct
    htree_nx = nx.Graph()
    for comp in nx.connected_components(dsg_nx):
        comp_g = dsg_nx.subgraph(comp).copy()
        room_g, room_idx = c.get_room_graph(comp_g)
        room_jth, room_roots = c.generate_component_jth(dsg_nx_component=room_g, component_type="rooms",
                                                        num_zero_padding=6, verbose=False)
        ht = c.HTree(room_jth=room_jth, room_root_nodes=room_roots)
        for r in room_idx:
            obj_g, _ = c.get_object_graph(comp_g, r)
            for oc in nx.connected_components(obj_g):
                sub = dsg_nx.subgraph(oc).copy()
                jth, roots = c.generate_component_jth(dsg_nx_component=sub, component_type="objects",
                                                      room_node_data=(r, comp_g.nodes[r]), num_zero_padding=6,
                                                      verbose=False)
                ht.add_object_jth(object_jth=jth, object_root_nodes=roots, room_idx=r)
        if len(room_idx) == 1 and ht.jth.number_of_nodes() > 1:
            ht.jth.nodes[0]["clique_has"] = ht.jth.nodes[0]["clique_has"][0]
        htree_nx = nx.disjoint_union(htree_nx, ht.jth)
    aug = c.add_virtual_nodes_to_htree(htree_nx)
    return repr([(n, d["node_type"], d["clique_has"]) for n, d in aug.nodes.items()]) + repr(list(aug.edges))


def main():
    meta = {"oracle_kind": ORACLE_KIND, "torch": torch.__version__}
    import networkx

    meta["networkx"] = networkx.__version__

    # ---- 1. Stanford3DSG pkl -> flat arrays (data, not source; 2,820 nodes)
    with open("/root/reference/data/Stanford3DSG.pkl", "rb") as f:
        lists, _sem, num_labels = pickle.load(f)
    xs, ys, eis, rms = lists["x_list"], lists["y_list"], lists["edge_index_list"], lists["room_mask_list"]
    node_ptr = np.concatenate([[0], np.cumsum([np.asarray(x).shape[0] for x in xs])]).astype(np.int64)
    edge_ptr = np.concatenate([[0], np.cumsum([np.asarray(e).shape[1] for e in eis])]).astype(np.int64)
    np.savez_compressed(os.path.join(HERE, "stanford3dsg.npz"),
                        x=np.concatenate([np.asarray(x, dtype=np.float32) for x in xs]),
                        y=np.concatenate([np.asarray(y, dtype=np.int64) for y in ys]),
                        room_mask=np.concatenate([np.asarray(m, dtype=bool) for m in rms]),
                        edge_index=np.concatenate([np.asarray(e, dtype=np.int64) for e in eis], axis=1),
                        node_ptr=node_ptr, edge_ptr=edge_ptr, num_labels=np.asarray(num_labels))
    stanford = cases.load_stanford()
    meta["stanford_graphs"] = len(stanford)

    # ---- 2. golden hash of the 482 Stanford H-trees
    sha = stanford_nx_signature(stanford)
    meta["stanford_htree_sha256"] = sha
    print("stanford H-tree sha256:", sha)

    # ---- 3. H-tree fixtures through the reference's networkx code
    table = torch.from_numpy(synth.word2vec_table())
    trees, sigs = [], []
    for gid in cases.HTREE_SYNTH_IDS:
        g = synth.synth_mp3d_batch([gid])
        t, sig = reference_htree(g, pad_to=306, trunc=300)
        for nt in ("object", "object_virtual"):  # the 300 columns not stored are reconstructible
            assert torch.equal(t[nt].x[:, 6:], table[t[nt].label.long()])
        trees.append(t)
        sigs.append(hashlib.sha256(sig.encode()).hexdigest())
    cases.save_htrees(os.path.join(HERE, "htree_synth.npz"), trees, wide_object=True)
    meta["htree_synth_sha256"] = dict(zip(map(str, cases.HTREE_SYNTH_IDS), sigs))
    trees_st = []
    for gid in cases.HTREE_STANFORD_IDS:
        t, _ = reference_htree(cases.stanford_hetero(stanford[gid]), pad_to=None, trunc=0)
        trees_st.append(t)
    cases.save_htrees(os.path.join(HERE, "htree_stanford.npz"), trees_st, wide_object=False)
    for a, b in zip(trees, cases.load_htrees("htree_synth")):  # round trip of the fixture format
        for nt in cases.HTREE_ALL_NODE_TYPES:
            assert torch.equal(a[nt].x, b[nt].x), nt
        for et in cases.HTREE_ALL_EDGE_TYPES:
            assert torch.equal(a[et].edge_index.long(), b[et].edge_index)

    # ---- 4. reference hetero -> homogeneous conversions (integer golden)
    conv = {}
    for k, t in enumerate(trees[:3]):
        hom = ref_mp3d.heterogeneous_htree_to_homogeneous(t.clone())
        for key in ("edge_index", "init_edge_index", "pool_edge_index", "room_mask", "object_mask", "node_type"):
            conv[f"htree{k}|{key}"] = hom[key].numpy()
        conv[f"htree{k}|x_sum"] = hom.x.double().sum(0).numpy()
    for k, gid in enumerate([0, 3]):
        g = to_ref_hetero(synth.synth_mp3d_batch([gid]))
        hom, node_types = ref_mp3d.heterogeneous_data_to_homogeneous(g)
        assert node_types == ["objects", "rooms"]
        conv[f"dsg{k}|edge_index"] = hom.edge_index.numpy()
        conv[f"dsg{k}|node_type"] = hom.node_type.numpy()
        conv[f"dsg{k}|x_sum"] = hom.x.double().sum(0).numpy()
    np.savez_compressed(os.path.join(HERE, "convert_golden.npz"), **conv)

    # ---- 5. logits of the reference's own model classes
    logits = {}
    for case, (cls, kw, inp) in cases.CASES.items():
        model = cases.build_model(ref_models, case)
        with torch.no_grad():
            out = model(cases.build_input(inp))
        outs = list(out) if isinstance(out, (tuple, list)) else [out]
        for k, o in enumerate(outs):
            assert torch.isfinite(o).all(), case
            logits[f"{case}|{k}"] = o.numpy().astype(np.float32)
        print(f"{case:18s} {cls:32s} -> {[tuple(o.shape) for o in outs]}")
    np.savez_compressed(os.path.join(HERE, "logits.npz"), **logits)

    # ---- 6. gradient fingerprints of the reference's own model classes (training parity)
    grads = {}
    for case, (cls, kw, inp) in cases.CASES.items():
        model = cases.build_model(ref_models, case)
        for name, v in cases.grad_fingerprint(model, cases.build_input(inp)).items():
            grads[f"{case}|{name}"] = v
        print(f"{case:18s} gradient fingerprint of {sum(1 for k in grads if k.startswith(case + '|'))} parameters")
    np.savez_compressed(os.path.join(HERE, "grads.npz"), **grads)

    with open(os.path.join(HERE, "meta.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    print(json.dumps(meta, indent=1, sort_keys=True))


if __name__ == "__main__":
    main()
