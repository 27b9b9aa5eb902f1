"""BASELINE.json configs[3]: 1M synthetic scene graphs sharded by graph across the GPUs of one box, heterogeneous GAT
(M-GAT) forward in micro-batches of 16,384 graphs, ONE NCCL gather of the room logits at the end.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_1m.py \
        [--total 1000000] [--micro 16384] [--repeat 3]

Rank r owns graph ids [r T/W, (r+1) T/W) (dist.shard_range).  Two forms:
  default    the micro-batches are generated once (host, untimed) and kept resident in HBM; the local pass is one CUDA graph
  --stream   every micro-batch is GENERATED ON THE DEVICE inside the timed region (hg_synth_mp3d_*: same graphs, bit for
             bit, as the host generator), consumed by the forward and dropped: nothing but one micro-batch is resident,
             which is the only form a 1 M-graph job (~75 GB of inputs) has on a host that cannot hold it.  The counting
             pass of all micro-batches (output sizes) runs once up front with one host sync.
Timed region (CUDA events between two barriers, max over ranks): [generation +] every micro-batch forward (plan rebuilt
per micro-batch) + the ragged logit gather.  One JSON line on rank 0.
"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

from bench import M_GAT, ClockSampler


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--total", type=int, default=1_000_000)
    ap.add_argument("--micro", type=int, default=16384)
    ap.add_argument("--repeat", type=int, default=10)
    ap.add_argument("--no-graph", action="store_true", help="launch every micro-batch eagerly")
    ap.add_argument("--stream", action="store_true", help="generate every micro-batch on the device inside the timed region")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    dist.init_process_group("nccl" if world > 1 else "gloo", rank=rank, world_size=world,
                            **({"device_id": dev} if world > 1 else {}))

    import hydra_gnn_b200 as hg
    from hydra_gnn_b200 import _lib, synth
    from hydra_gnn_b200.dist import LogitGather, shard_range

    lo, hi = shard_range(args.total, rank, world)
    t0 = time.perf_counter()
    micro, rooms = [], 0
    gen = synth.DeviceSynth(dev) if args.stream else None
    if gen is not None:
        ranges = [(a, min(a + args.micro, hi) - a) for a in range(lo, hi, args.micro)]
        for a, n in ranges:
            gen.counts_async(a, n)  # all counting passes queued, then ONE sync reads the totals
        for a, n in ranges:
            sz = gen.sizes(a, n)
            micro.append(((a, n, sz), rooms, sz[0]))
            rooms += sz[0]
    for a in ([] if gen is not None else range(lo, hi, args.micro)):
        host = synth.synth_mp3d_batch(range(a, min(a + args.micro, hi)))
        d = hg.data.HeteroData()
        for t in host.node_types:
            d[t].x = host[t].x.to(dev)
        for t in host.edge_types:
            d[t].edge_index = host[t].edge_index.to(dev)
        micro.append((d, rooms, host["rooms"].x.size(0)))
        rooms += host["rooms"].x.size(0)
    t_gen = time.perf_counter() - t0
    net = synth.random_state_dict_(hg.HeterogeneousNetwork(**M_GAT), seed=0).eval().to(dev)
    net.plan_cache.enabled = False
    logits = torch.empty(rooms, 26, dtype=torch.float32, device=dev)
    gather = LogitGather(26).prepare(rooms, dev) if world > 1 else None

    def forward_all():
        for d, off, n in micro:
            if gen is not None:  # (first id, graphs, sizes): generate, consume, drop
                gen.counts_async(d[0], d[1])
                d = gen.batch(d[0], d[1], d[2], labels=False)
            logits[off:off + n].copy_(net(d))

    graph = None
    if not args.no_graph:  # --stream too: generation is sync-free once the sizes are known, and every replay regenerates
        # the micro-batches have fixed shapes: the whole local pass (plans included) is captured once and replayed, so the
        # job is bound by the device, not by ~340 eager launches per pass on a host shared by eight ranks (round 1: 26-50 ms)
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            forward_all()
            side.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                forward_all()
        torch.cuda.current_stream(dev).wait_stream(side)

    def job():
        if graph is not None:
            graph.replay()
        else:
            forward_all()
        if gather is not None:
            gather(logits)

    def fence():
        dist.barrier()
        torch.cuda.synchronize(dev)

    job()  # warm-up: weight images, workspaces, NCCL channels
    fence()
    l0 = _lib.launch_count
    times = []
    with ClockSampler(local) as clocks:
        for _ in range(args.repeat):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            fence()
            e0.record()
            job()
            e1.record()
            fence()
            times.append(e0.elapsed_time(e1))
    launches = _lib.launch_count - l0
    t = torch.tensor(times, device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    times = sorted(t.tolist())
    ms = times[len(times) // 2]
    n_rooms_all = sum(gather.counts) if gather is not None else rooms
    if rank == 0:
        got = gather.unpad() if gather is not None else logits
        assert got.size(0) == n_rooms_all and bool(torch.isfinite(got).all())
        # graph order check: this rank's rows come first and equal its local logits
        assert torch.equal(got[:rooms], logits)
        print(json.dumps({
            "metric": "scene_graphs_per_s_forward", "value": args.total / (ms / 1e3), "unit": "graphs/s", "n_gpus": world,
            "ms_per_job": ms, "ms_all_repeats": times, "higher_is_better": True, "scaling": "strong", "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": "configs[3]: 1M synthetic MP3D-shaped scene graphs sharded by graph, M-GAT forward in "
                                   f"micro-batches of {args.micro}, one NCCL gather of the room logits",
                       "total_graphs": args.total, "micro_batch": args.micro, "micro_batches_per_rank": len(micro),
                       "rooms_total": n_rooms_all, "gathered_bytes": n_rooms_all * 26 * 4,
                       "collective": "all_gather_into_tensor(max-padded room logits)" if world > 1 else "none",
                       "cuda_graph": graph is not None,
                       "inputs": ("generated on the device inside the timed region, one micro-batch resident"
                                  if gen is not None else "generated on the host (untimed), resident in HBM"),
                       "setup_s_per_rank": round(t_gen, 1)},
            "gpu_launches": launches, "clocks": clocks.summary()}), flush=True)
    graph = None
    torch.cuda.synchronize(dev)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
