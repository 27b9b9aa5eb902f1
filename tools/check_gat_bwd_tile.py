"""First-run check of the opt-in tile kernel for pass 1 of the GAT backward (csrc/gat_train_tile.cu,
HG_GAT_BWD_TILE=1) against the default, GPU-verified kernel.  Own process on purpose:

    timeout 300 python tools/check_gat_bwd_tile.py [--graphs 512]

For the heterogeneous GAT (M-GAT) and the heterogeneous Neural-Tree GAT, with dropout 0 and 0.4 (same host seed for
both runs, so both draw the same masks): one forward + backward per kernel, every parameter gradient compared
(max |a - b| <= 2e-5 * max |b| + 1e-7: the two kernels sum each row in a different order).  Then the M-GAT backward is
timed with each kernel.  Prints one JSON line per case; exits non-zero on the first mismatch.
"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--graphs", type=int, default=512)
    args = ap.parse_args()
    import hydra_gnn_b200 as hg
    from hydra_gnn_b200 import htree, synth
    from bench import M_GAT

    dev = torch.device("cuda", 0)

    def to_dev(host):
        d = hg.data.HeteroData()
        for t in host.node_types:
            d[t].x = host[t].x.to(dev)
        for t in host.edge_types:
            d[t].edge_index = host[t].edge_index.to(dev)
        return d

    src = synth.synth_mp3d_batch(range(args.graphs))
    flat = to_dev(src)
    y_flat = src["rooms"].y.long().to(dev)
    tree_host = htree.generate_htree_batch(src)
    tree = to_dev(tree_host)
    y_tree = tree_host["room_virtual"].label.long().to(dev)
    nt_kw = dict(input_dim_dict={t: int(tree_host[t].x.size(1)) for t in tree_host.node_types}, output_dim=26,
                 conv_block="GAT", GAT_hidden_dims=[64, 64], GAT_heads=[3, 3, 3], GAT_concats=[False] * 3,
                 disable_initialization=True)
    cases = [("mgat", hg.HeterogeneousNetwork, dict(M_GAT), flat, y_flat),
             ("nt_het_gat", hg.HeterogeneousNeuralTreeNetwork, nt_kw, tree, y_tree)]

    def grads(cls, kw, data, y, dropout, tile):
        if tile:
            os.environ["HG_GAT_BWD_TILE"] = "1"
        else:
            os.environ.pop("HG_GAT_BWD_TILE", None)
        net = synth.random_state_dict_(cls(**dict(kw, dropout=dropout)), seed=0).train().to(dev)
        torch.manual_seed(1234)  # host-drawn dropout seeds and torch's feature dropout: identical in both runs
        torch.cuda.manual_seed(1234)
        keep = (y != 25).nonzero().view(-1)
        loss = F.cross_entropy(net(data).index_select(0, keep), y.index_select(0, keep))
        loss.backward()
        torch.cuda.synchronize()
        return {n: p.grad.detach().clone() for n, p in net.named_parameters() if p.grad is not None}, net

    for name, cls, kw, data, y in cases:
        for dropout in (0.0, 0.4):
            base, _ = grads(cls, kw, data, y, dropout, tile=False)
            got, _ = grads(cls, kw, data, y, dropout, tile=True)
            worst, worst_name = 0.0, ""
            for n, b in base.items():
                err = (got[n] - b).abs().max().item() / (2e-5 * b.abs().max().item() + 1e-7)
                if not (err <= worst):
                    worst, worst_name = err, n
            ok = worst <= 1.0 and all(torch.isfinite(g).all() for g in got.values())
            print(json.dumps(dict(case=name, dropout=dropout, params=len(base), worst_err_over_bound=round(worst, 4),
                                  worst_param=worst_name, ok=bool(ok))), flush=True)
            if not ok:
                sys.exit(1)

    # timing: M-GAT forward + backward, each kernel
    name, cls, kw, data, y = cases[0]
    res = {}
    for tile in (False, True):
        _, net = grads(cls, kw, data, y, 0.4, tile)
        keep = (y != 25).nonzero().view(-1)

        def step():
            net.zero_grad(set_to_none=True)
            F.cross_entropy(net(data).index_select(0, keep), y.index_select(0, keep)).backward()

        for _ in range(3):
            step()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(10):
            step()
        e.record()
        torch.cuda.synchronize()
        res["tile" if tile else "default"] = round(s.elapsed_time(e) / 10, 3)
    print(json.dumps(dict(case="mgat fwd+bwd ms", graphs=args.graphs, **res)), flush=True)


if __name__ == "__main__":
    main()
