"""Per-kernel tables of the OTHER model families at the bench scale (4096 MP3D-shaped graphs, 1 GPU):
the segmented-aggregation kernel (hg_seg_reduce) behind GraphSAGE / GCN / GIN, and the homogeneous GAT.
Not a bench line (bench.py's workload is M-GAT); evidence for the aggregation kernels' HBM fraction.

    python tools/bench_models.py [--graphs 4096] [--iters 10] > profiles/r01_models.json
"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import hydra_gnn_b200 as hg
from hydra_gnn_b200 import ops, synth, convert
from bench import roofline_from, _peaks

MODELS = {
    "hetero_sage": ("HeterogeneousNetwork", dict(input_dim_dict={"objects": 306, "rooms": 6}, output_dim=26,
                                                 conv_block="GraphSAGE", hidden_dim=128, num_layers=3, dropout=0.25), False),
    "hom_gcn": ("HomogeneousNetwork", dict(input_dim=306, output_dim=26, conv_block="GCN", hidden_dim=128, num_layers=4), True),
    "hom_sage": ("HomogeneousNetwork", dict(input_dim=306, output_dim=26, conv_block="GraphSAGE", hidden_dim=128, num_layers=4), True),
    "hom_gin": ("HomogeneousNetwork", dict(input_dim=306, output_dim=26, conv_block="GIN", hidden_dim=128, num_layers=3), True),
    "hom_gat": ("HomogeneousNetwork", dict(input_dim=306, output_dim=26, conv_block="GAT", GAT_hidden_dims=[64, 64],
                                           GAT_heads=[3, 3, 3], GAT_concats=[False, False, False]), True),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--graphs", type=int, default=4096)
    ap.add_argument("--iters", type=int, default=10)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    het = synth.synth_mp3d_batch(range(args.graphs), device=dev)
    # the conversion zero-pads the features of its argument in place (as the reference's does): give it its own copy
    hom = convert.dsg_to_homogeneous(synth.synth_mp3d_batch(range(args.graphs), device=dev))
    peaks = _peaks()
    out = {"graphs": args.graphs, "peak_hbm_gbs": peaks["hbm_gbs"], "models": {}}
    for name, (cls, kw, use_hom) in MODELS.items():
        net = synth.random_state_dict_(getattr(hg, cls)(**kw), seed=0).eval().to(dev)
        net.plan_cache.enabled = False
        data = hom if use_hom else het
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
        sink = []
        ops.set_profile_sink(sink)
        for _ in range(args.iters):
            net(data)
        torch.cuda.synchronize()
        ops.set_profile_sink(None)
        _, table = roofline_from(sink, args.iters, peaks)
        for k in table:
            k["hbm_frac"] = round(k["gbs"] / peaks["hbm_gbs"], 4) if k["gbs"] else None
        out["models"][name] = {"class": cls, "kwargs": {k: v for k, v in kw.items()}, "ms_per_forward_eager": round(ms, 4),
                               "graphs_per_s_eager": round(args.graphs / ms * 1e3), "kernels": table}
        print(f"{name}: {ms:.3f} ms/forward eager", file=sys.stderr)
        for k in table[:6]:
            print(f"   {k['kernel']:60s} x{k['launches_per_step']:.0f} {k['mean_us']:8.1f} us {k['gbs']} GB/s", file=sys.stderr)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
