#!/usr/bin/env python
"""Differential check of the native H-tree builder against the REFERENCE's networkx code (build container
only: needs /root/reference).  usage: htree_diff.py [first_id] [count] [--stanford]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests", "golden")]
import make_golden as mg  # noqa: E402  (imports the reference)
import cases  # noqa: E402
from hydra_gnn_b200 import htree, synth  # noqa: E402


def hom_inputs(g):
    """(is_room, edge_index) of a product HeteroData in to_homogeneous() order."""
    off, cum = {}, 0
    for nt in g.node_types:
        off[nt] = cum
        cum += g[nt].x.size(0)
    is_room = np.zeros(cum, np.uint8)
    is_room[off["rooms"]:off["rooms"] + g["rooms"].x.size(0)] = 1
    eis = []
    for et in g.edge_types:
        ei = g[et].edge_index.clone().long()
        ei[0] += off[et[0]]
        ei[1] += off[et[2]]
        eis.append(ei)
    return is_room, torch.cat(eis, 1).numpy()


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    first, count = (int(args[0]) if args else 0), (int(args[1]) if len(args) > 1 else 64)
    if "--stanford" in sys.argv:
        graphs = [cases.stanford_hetero(g) for g in cases.load_stanford()[first:first + count]]
        pad = None
    else:
        graphs = [synth.synth_mp3d_batch([i]) for i in range(first, first + count)]
        pad = 306
    t0 = time.time()
    raw = htree.build_raw([hom_inputs(g) for g in graphs], n_threads=1)
    t_native = time.time() - t0
    bad, t_ref = [], 0.0
    for k, g in enumerate(graphs):
        t0 = time.time()
        h = mg.to_ref_hetero(g, pad_to=pad)
        aug = mg.ref_construct.add_virtual_nodes_to_htree(mg.ref_construct.generate_htree(h, verbose=False))
        t_ref += time.time() - t0
        want = repr([(n, d["node_type"], d["clique_has"]) for n, d in aug.nodes.items()]) + repr(list(aug.edges))
        got = htree.signature(raw, k)
        if got != want:
            bad.append(first + k)
            if len(bad) <= 3:
                i = next(j for j in range(min(len(got), len(want))) if got[j] != want[j]) if got[:len(want)] != want[:len(got)] else min(len(got), len(want))
                print(f"graph {first + k}: mismatch at char {i}\n  want ...{want[max(0, i - 120):i + 120]}\n  got  ...{got[max(0, i - 120):i + 120]}")
    print(f"{len(graphs) - len(bad)}/{len(graphs)} identical; native {t_native * 1e3 / len(graphs):.3f} ms/graph (1 thread), "
          f"reference {t_ref * 1e3 / len(graphs):.1f} ms/graph; mismatches: {bad[:20]}")


if __name__ == "__main__":
    main()
