"""BASELINE.json configs[2]: Neural Tree networks over the H-tree decompositions of synthetic MP3D-shaped graphs
(M-NT (i) homogeneous GCN, (ii) homogeneous GraphSAGE, (iii) heterogeneous GAT), 1 GPU.

H-trees come from the native builder (libhydra_htree.so, untimed prep as in the reference, where the dataset
build does it; its wall time is reported).  Per model: eager ms/forward, CUDA-graph replay ms/forward, graphs/s,
the per-kernel table with algorithmic GB/s, and (small sample) the max deviation from the CPU oracle.

    python tools/bench_nt.py [--graphs 4096] [--iters 10] > profiles/r01_nt.json
"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import hydra_gnn_b200 as hg
from hydra_gnn_b200 import convert, htree, ops, synth
from hydra_gnn_b200.data import Batch, Data, HeteroData
from bench import roofline_from, _peaks

HT_DIMS = {"object": 306, "room": 6, "object-room": 6, "room-room": 6, "object_virtual": 306, "room_virtual": 6}
MODELS = {
    "nt_hom_gcn": ("HomogeneousNeuralTreeNetwork", dict(input_dim=306, output_dim=26, conv_block="GCN", hidden_dim=128,
                                                        num_layers=4, disable_initialization=True), True),
    "nt_hom_sage": ("HomogeneousNeuralTreeNetwork", dict(input_dim=306, output_dim=26, conv_block="GraphSAGE",
                                                         hidden_dim=128, num_layers=4, disable_initialization=True), True),
    "nt_het_gat": ("HeterogeneousNeuralTreeNetwork", dict(input_dim_dict=HT_DIMS, output_dim=26, conv_block="GAT",
                                                          GAT_hidden_dims=[64, 64], GAT_heads=[3, 3, 3],
                                                          GAT_concats=[False] * 3, disable_initialization=True), False),
    "nt_het_gat_premp": ("HeterogeneousNeuralTreeNetwork", dict(input_dim_dict=HT_DIMS, output_dim=26, conv_block="GAT",
                                                                GAT_hidden_dims=[64, 64], GAT_heads=[3, 3, 3],
                                                                GAT_concats=[False] * 3, disable_initialization=False), False),
}


def _to(data, dev):
    if isinstance(data, HeteroData):
        d = HeteroData()
        for t in data.node_types:
            d[t].x = data[t].x.to(dev)
            d[t].num_nodes = data[t].x.size(0)
        for et in data.edge_types:
            d[et].edge_index = data[et].edge_index.to(dev)
        return d
    d = Data()  # Data.to() moves in place (as PyG's does): build a device copy instead
    for k in ("x", "edge_index", "init_edge_index", "pool_edge_index", "room_mask", "object_mask"):
        setattr(d, k, getattr(data, k).to(dev))
    return d


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--graphs", type=int, default=4096)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--check", type=int, default=16, help="graphs of the oracle comparison (0 = skip)")
    ap.add_argument("--only", default=None)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    peaks = _peaks()
    out = {"graphs": args.graphs, "peak_hbm_gbs": peaks["hbm_gbs"], "models": {}}

    # one batched call of the native builder, timed on its own (all host threads)
    t0 = time.perf_counter()
    src = synth.synth_mp3d_batch(range(args.graphs))
    t1 = time.perf_counter()
    het = htree.generate_htree_batch(src)
    t2 = time.perf_counter()
    out["htree_builder"] = {"seconds_batched_call": round(t2 - t1, 3), "graphs_per_s": round(args.graphs / (t2 - t1)),
                            "threads": os.cpu_count(), "synth_seconds": round(t1 - t0, 3)}
    # the homogeneous form (reference: per-graph conversion + collate) from the batched H-tree in one pass;
    # tests/test_htree.py pins it bit-exactly to the per-graph route, and --check below uses that route
    t3 = time.perf_counter()
    hom = convert.batched_htree_to_homogeneous(het)
    out["htree_builder"]["homogeneous_conversion_seconds"] = round(time.perf_counter() - t3, 3)
    out["sizes"] = {"hetero_nodes": {t: int(het[t].x.size(0)) for t in het.node_types},
                    "hetero_edges": {et[1]: int(het[et].edge_index.size(1)) for et in het.edge_types},
                    "homogeneous_nodes": int(hom.x.size(0)), "tree_edges": int(hom.edge_index.size(1)),
                    "init_edges": int(hom.init_edge_index.size(1)), "pool_edges": int(hom.pool_edge_index.size(1)),
                    "room_virtual": int(hom.room_mask.sum())}
    print(json.dumps(out["sizes"]), file=sys.stderr)
    het_d, hom_d = _to(het, dev), _to(hom, dev)
    del hom

    small = None
    if args.check:
        from oracle import models as om  # checker only
        s_trees = [htree.generate_htree(synth.synth_mp3d_batch([g])) for g in range(args.check)]
        s_het = Batch.from_data_list(s_trees)
        s_het_x = {t: s_het[t].x.clone() for t in s_het.node_types}
        s_hom = Batch.from_data_list([convert.heterogeneous_htree_to_homogeneous(t) for t in s_trees])
        s_het2 = HeteroData()
        for t in s_het.node_types:
            s_het2[t].x = s_het_x[t]
            s_het2[t].num_nodes = s_het_x[t].size(0)
        for et in s_het.edge_types:
            s_het2[et].edge_index = s_het[et].edge_index
        small = (s_het2, s_hom, om)

    for name, (cls, kw, use_hom) in MODELS.items():
        if args.only and args.only != name:
            continue
        net = synth.random_state_dict_(getattr(hg, cls)(**kw), seed=0).eval().to(dev)
        net.plan_cache.enabled = False
        data = hom_d if use_hom else het_d
        rec = {"class": cls, "kwargs": kw}
        if small is not None:
            s_het2, s_hom, om = small
            ref = getattr(om, cls)(**kw)
            ref.load_state_dict(net.state_dict())
            ref = ref.eval()
            with torch.no_grad():
                want = ref(s_hom if use_hom else s_het2)
            got = net(_to(s_hom if use_hom else s_het2, dev)).cpu()
            rec["oracle_check"] = {"graphs": args.check, "max_abs_err": float((got - want).abs().max()),
                                   "max_abs_ref": float(want.abs().max()),
                                   "argmax_equal": bool(torch.equal(got.argmax(1), want.argmax(1)))}
        for _ in range(3):
            net(data)
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(args.iters):
            net(data)
        e.record()
        torch.cuda.synchronize()
        ms = s.elapsed_time(e) / args.iters
        # CUDA-graph replay of the same forward (plan included)
        cap = torch.cuda.Stream(dev)
        with torch.cuda.stream(cap):
            net(data)
            cap.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=cap):
                res = net(data)
        torch.cuda.synchronize()
        for _ in range(3):
            g.replay()
        s.record()
        for _ in range(args.iters):
            g.replay()
        e.record()
        torch.cuda.synchronize()
        ms_g = s.elapsed_time(e) / args.iters
        del g, res
        sink = []
        ops.set_profile_sink(sink)
        saved = ops.PARALLEL_BRANCHES, net.plan_cache.overlap
        ops.PARALLEL_BRANCHES, net.plan_cache.overlap = False, False
        for _ in range(args.iters):
            net(data)
        torch.cuda.synchronize()
        ops.PARALLEL_BRANCHES, net.plan_cache.overlap = saved
        ops.set_profile_sink(None)
        _, table = roofline_from(sink, args.iters, peaks)
        for k in table:
            k["hbm_frac"] = round(k["gbs"] / peaks["hbm_gbs"], 4) if k["gbs"] else None
        alg = sum(k["alg_bytes"] * k["launches_per_step"] for k in table)
        rec.update({"ms_per_forward_eager": round(ms, 4), "ms_per_forward_graph": round(ms_g, 4),
                    "graphs_per_s": round(args.graphs / ms_g * 1e3), "launches_per_forward": sum(k["launches_per_step"] for k in table),
                    "alg_bytes_per_forward": int(alg), "alg_gbs_whole_forward": round(alg / ms_g / 1e6, 1),
                    "kernels": table})
        out["models"][name] = rec
        print(f"{name}: eager {ms:.3f} ms, graph {ms_g:.3f} ms, {args.graphs / ms_g * 1e3:.0f} graphs/s, "
              f"check {rec.get('oracle_check')}", file=sys.stderr)
        for k in table[:8]:
            print(f"   {k['kernel']:64s} x{k['launches_per_step']:.0f} {k['mean_us']:8.1f} us {k['gbs']} GB/s", file=sys.stderr)
        del net
        torch.cuda.empty_cache()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
