"""BASELINE.json configs[4]: heterogeneous GAT (M-GAT) TRAINING step -- forward (dropout 0.4: feature dropout +
in-kernel attention dropout) + masked cross-entropy + backward + gradient all-reduce (NCCL, N > 1) + Adam --
on synthetic MP3D-shaped graphs, 8192 graphs per GPU per step (reference: base_training_job.py:203-219).

    python tools/bench_train.py [--graphs 8192] [--steps 10] [--warmup 3]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_train.py ...

One JSON line on rank 0: graphs/s of the whole job (max over ranks, CUDA events), the per-kernel table of one
eager pass (algorithmic bytes / flops per launch), and the CPU oracle's train step on a bounded sample.
"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
import torch.nn.functional as F

from bench import M_GAT, ClockSampler, _peaks, algorithmic, kernel_key


def alg(name, m):
    if name == "hg_linear_wgrad":
        return 4 * (m["M"] * m["N"] + m["M"] * m["K"] + m["N"] * m["K"]), 2 * m["M"] * m["N"] * m["K"]
    if name in ("hg_gat_train_fwd", "hg_gat_train_bwd"):
        HC, H = m["H"] * m["C"], m["H"]
        b = 4 * m["n_dst"] * m["width"] * (1 if name.endswith("fwd") else 3)  # out | out, dout, g
        for r in m["rels"]:
            b += 4 * r["n_src"] * (HC + H) + 4 * m["n_dst"] * H + 4 * (m["n_dst"] + 1) + 4 * r["cap"] + 4 * r["cap"] * H
            if name.endswith("bwd"):  # + dlogit, reversed CSR, gradient rows
                b += 8 * r["cap"] * H + 8 * r["cap"] + 4 * (r["n_src"] + 1) + 4 * r["n_src"] * (HC + H) + 4 * m["n_dst"] * H
        return b, 0
    if name == "hg_adam_step":
        return 28 * m["n"], 0
    return algorithmic(name, m)


def key(name, m):
    if name == "hg_linear_wgrad":
        return f"hg_linear_wgrad[M={m['M']},N={m['N']},K={m['K']}]"
    if name in ("hg_gat_train_fwd", "hg_gat_train_bwd"):
        return f"{name}[n_dst={m['n_dst']},H={m['H']},C={m['C']},rels={len(m['rels'])}]"
    return kernel_key(name, m)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--graphs", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--dropout", type=float, default=0.4)
    ap.add_argument("--cpu-graphs", type=int, default=512)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--edge-attr", action="store_true",
                    help="M-GATe: conv_block GAT_edge on 303 / 3-d inputs with relative-position edge attributes (what the "
                         "reference's MP3D configs train); first GPU run pending, see DESIGN.md section 4.4")
    ap.add_argument("--no-graph", action="store_true", help="launch every step eagerly instead of replaying a CUDA graph")
    ap.add_argument("--graph", action="store_true",
                    help="(default since round 2) capture the step, NCCL all-reduce included, as a CUDA graph")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    import hydra_gnn_b200 as hg
    from hydra_gnn_b200 import _lib, ops, synth
    from hydra_gnn_b200.dist import shard_range
    from hydra_gnn_b200 import train as hgt
    from hydra_gnn_b200.train import FlatAdam

    G = args.graphs
    lo, hi = shard_range(G * world, rank, world)
    host = synth.synth_mp3d_batch(range(lo, hi), relative_pos=args.edge_attr)
    data = hg.data.HeteroData()
    for t in host.node_types:
        data[t].x = host[t].x.to(dev)
    for t in host.edge_types:
        data[t].edge_index = host[t].edge_index.to(dev)
        if args.edge_attr:
            data[t].edge_attr = host[t].edge_attr.to(dev)
    labels = host["rooms"].y.long().to(dev)
    mask = labels != 25  # base_training_job.py:63 ignored_label
    keep_rows = mask.nonzero().view(-1)
    n_local = torch.tensor([float(keep_rows.numel())], device=dev)
    if world > 1:
        dist.all_reduce(n_local)
    n_global = float(n_local.item())

    kw = dict(M_GAT, dropout=args.dropout)
    if args.edge_attr:
        kw.update(conv_block="GAT_edge", input_dim_dict={"objects": 303, "rooms": 3})
    net = synth.random_state_dict_(hg.HeterogeneousNetwork(**kw), seed=0).train().to(dev)
    net.plan_cache.enabled = False  # plan (forward + reversed CSRs) rebuilt every step
    opt = FlatAdam(net.parameters(), lr=1e-3, weight_decay=1e-3)

    # The step is replayed as ONE CUDA graph at every N; at N > 1 the graph contains the NCCL gradient all-reduce
    # (verified on 2 x B200 in round 2: 8.50 ms against 9.09 ms eager and 8.34 ms on one GPU).
    use_graph = not args.no_graph  # N > 1: the captured step contains the NCCL all-reduce (thread_local capture mode)
    if use_graph:  # dropout masks must change on every replay: the seed lives on the device
        hgt.enable_device_seed(dev, 20240229 + rank)

    def step():
        hgt.advance_seed()
        opt.zero_grad()
        logits = net(data)
        # global masked mean: local sum / global count, gradients summed over ranks
        loss = F.cross_entropy(logits.index_select(0, keep_rows), labels.index_select(0, keep_rows), reduction="sum")
        (loss / n_global).backward()
        opt.flatten()  # first call only: re-homes the live parameters / gradients into the flat buffers
        if world > 1:
            dist.all_reduce(opt.flat_grad)  # the one collective of the step (sum; the loss is already / global count)
        opt.step()
        return loss

    def fence():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    eager_step = step
    side = torch.cuda.Stream(dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):  # eager warm-up: flattens the parameters, sizes workspaces, primes NCCL
        for _ in range(3):
            step()
    torch.cuda.current_stream(dev).wait_stream(side)
    fence()
    if use_graph:
        # the whole step (plan, forward, loss, backward, gradient all-reduce, Adam) as ONE CUDA graph
        held = {"graph": torch.cuda.CUDAGraph()}
        mode = dict(capture_error_mode="thread_local") if world > 1 else {}  # the NCCL watchdog thread polls events
        with torch.cuda.graph(held["graph"], stream=side, **mode):
            static_loss = eager_step()
        fence()

        def step():
            held["graph"].replay()
            return static_loss

    for _ in range(max(args.warmup, 3)):
        step()
    fence()
    with ClockSampler(local) as clocks:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            loss = step()
        e1.record()
        fence()
    ms = e0.elapsed_time(e1)
    final_loss = float(loss.detach()) / max(float(keep_rows.numel()), 1.0)

    sink = []
    ops.set_profile_sink(sink)
    l0 = _lib.launch_count
    saved = ops.PARALLEL_BRANCHES, net.plan_cache.overlap
    ops.PARALLEL_BRANCHES, net.plan_cache.overlap = False, False
    t_eager0 = torch.cuda.Event(enable_timing=True); t_eager1 = torch.cuda.Event(enable_timing=True)
    t_eager0.record()
    for _ in range(args.steps):
        eager_step()
    t_eager1.record()
    fence()
    ops.PARALLEL_BRANCHES, net.plan_cache.overlap = saved
    ops.set_profile_sink(None)
    launches = _lib.launch_count - l0
    ms_prof = t_eager0.elapsed_time(t_eager1)

    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
    if rank == 0:
        peaks = _peaks()
        agg = {}
        for name, meta, s, e in sink:
            k = key(name, meta)
            a = agg.setdefault(k, dict(ms=0.0, n=0, name=name, meta=meta))
            a["ms"] += s.elapsed_time(e)
            a["n"] += 1
        ours_ms = sum(a["ms"] for a in agg.values())
        table = []
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["ms"]):
            b, fl = alg(a["name"], a["meta"])
            mean = a["ms"] / a["n"]
            table.append(dict(kernel=k, launches_per_step=a["n"] / args.steps, mean_us=round(mean * 1e3, 1),
                              share_of_step=round(a["ms"] / ms_prof, 4), alg_bytes=b, alg_flops=fl,
                              gbs=round(b / mean / 1e6, 1) if mean > 0 and b else None,
                              hbm_frac=round(b / mean / 1e6 / peaks["hbm_gbs"], 4) if mean > 0 and b else None,
                              tflops=round(fl / mean / 1e9, 2) if mean > 0 and fl else None))
        line = {
            "metric": "scene_graphs_per_s_train_step", "value": G * world * args.steps / (ms / 1e3), "unit": "graphs/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "dtype": "f32", "data": "synthetic",
            "config": {"workload": ("configs[4] variant M-GATe (GAT_edge): " if args.edge_attr else "configs[4]: ") +
                                   "heterogeneous GAT (M-GAT) training step: forward (dropout "
                                   f"{args.dropout}) + masked CE + backward + gradient all-reduce + Adam, {G} graphs per GPU",
                       "graphs_per_gpu": G, "objects": int(host["objects"].x.size(0)), "rooms": int(host["rooms"].x.size(0)),
                       "collective": "all_reduce(flat gradient, %d floats)" % opt.flat_grad.numel() if world > 1 else "none",
                       "cuda_graph": use_graph},
            "gpu_launches": launches, "clocks": clocks.summary(),
            "our_kernels_share_of_step": round(ours_ms / ms_prof, 4),
            "loss_last_step": final_loss, "kernels": table,
        }
        if world == 1 and not args.no_cpu:
            from oracle import models as om  # CPU baseline: the oracle's autograd + torch.optim.Adam
            torch.set_num_threads(os.cpu_count() or 1)
            n = min(args.cpu_graphs, G)
            ref = synth.random_state_dict_(om.HeterogeneousNetwork(**kw), seed=0).train()
            o = torch.optim.Adam(ref.parameters(), lr=1e-3, weight_decay=1e-3)
            d = synth.synth_mp3d_batch(range(n), relative_pos=args.edge_attr)
            y = d["rooms"].y.long()
            mk = y != 25
            def cpu_step():
                o.zero_grad()
                ref.loss(ref(d), y, mk).backward()
                o.step()
            cpu_step()
            it, t0 = 0, time.perf_counter()
            while it < 20 and (it == 0 or time.perf_counter() - t0 < 10.0):
                cpu_step()
                it += 1
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": n * it / dt, "unit": "graphs/s", "cores": torch.get_num_threads(), "kind": "port",
                                    "sample": f"{it} train steps over a {n}-graph batch ({dt:.1f} s; oracle port, torch autograd + Adam)"}
        print(json.dumps(line), flush=True)
    if world > 1:
        # a captured graph that contains NCCL kernels must be gone before the communicator is torn down (a first 2-GPU run
        # printed its line and then hung in destroy_process_group)
        if use_graph:
            held.clear()
        torch.cuda.synchronize(dev)
        dist.barrier()
        dist.destroy_process_group()
    sys.stdout.flush()
    if world > 1:
        os._exit(0)  # skip interpreter teardown of NCCL state captured in graphs


if __name__ == "__main__":
    main()
