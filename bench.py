#!/usr/bin/env python
"""Benchmark of the hot path: heterogeneous GAT forward over batched MP3D-shaped scene graphs.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload = BASELINE.json configs[1]: synthetic MP3D-shaped room/object HeteroData (306-d object
features), M-GAT, 4096 graphs per GPU per step.  A step = graph plan (K1) +
the full 3-layer forward (+ the NCCL logit gather when N > 1).  One JSON line on rank 0.
See DESIGN.md "Measurement" for the definition of every field.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "scene_graphs_per_s_forward"
UNIT = "graphs/s"
GRAPHS_PER_GPU = 4096
M_GAT = dict(input_dim_dict={"objects": 306, "rooms": 6}, output_dim=26, conv_block="GAT", GAT_hidden_dims=[64, 64],
             GAT_heads=[3, 3, 3], GAT_concats=[False, False, False], dropout=0.4)


def _peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(hbm_gbs=p["hbm_gbs"], tflops=p.get("bf16_tflops_sustained", p["bf16_tflops"]), source="measured")
    return dict(hbm_gbs=6650.0, tflops=1400.0, source="fallback")  # B200_PROFILING.md fallback


# ------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i",
                 str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def __exit__(self, *exc):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self):
        sm, smax, reasons = [], [], set()
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                smax.append(float(parts[1]))
            except ValueError:
                continue
            for name, val in zip(self.NAMES, parts[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=[], samples=0)
        return dict(sm_mhz=statistics.median(sm), sm_max_mhz=max(smax), reasons=sorted(reasons), samples=len(sm))


# ------------------------------------------------------------------------------ algorithmic bytes
def algorithmic(name: str, m: dict):
    """(bytes, flops) of one ABI call: every distinct input element read once, every output
    written once, fp32 = 4 B, CSR int32; no credit for re-gathered rows."""
    if name == "hg_plan_build":
        b = 0
        for E, n_dst, cap in zip(m["edges"], m["n_dst"], m["caps"]):
            b += 16 * E + 4 * (n_dst + 1) + 4 * cap
        return b, 0
    if name == "hg_seg_reduce":
        b = 4 * m["n_src"] * m["F"] + 4 * (m["n_out"] + 1) + 4 * m["cap"] + 4 * m["n_out"] * m["F"]
        if m["self"]:
            b += 4 * m["n_out"] * m["F"]
        return b, 0
    if name == "hg_gat_fused":
        b = 4 * m["n_dst"] * m["width"] * (2 if m.get("addend") else 1)
        for r in m["rels"]:
            b += 4 * r["n_src"] * m.get("src_width", m["H"] * m["C"]) + 4 * r["n_src"] * m["H"] + 4 * m["n_dst"] * m["H"]
            b += 4 * (m["n_dst"] + 1) + 4 * r["cap"] + (4 * r["cap"] * m["H"] if r["edge"] else 0)
        return b, 0
    if name == "hg_linear":
        K = sum(m["K"])
        return 4 * (m["M"] * K + m["N"] * K + m["M"] * m["N"]), 2 * m["M"] * m["N"] * K
    return 0, 0


def kernel_key(name, m):
    if name == "hg_linear":
        return f"hg_linear[M={m['M']},N={m['N']},K={'+'.join(map(str, m['K']))},{m['precision']}]"
    if name == "hg_gat_fused":
        return f"hg_gat_fused[n_dst={m['n_dst']},H={m['H']},C={m['C']},rels={len(m['rels'])}]"
    if name == "hg_seg_reduce":
        return f"hg_seg_reduce[n_out={m['n_out']},F={m['F']}]"
    return name


def roofline_from(sink, steps, peaks):
    """Dominant kernel of the timed region: per-launch algorithmic bytes / mean CUDA-event time."""
    agg = {}
    for name, meta, s, e in sink:
        key = kernel_key(name, meta)
        a = agg.setdefault(key, dict(ms=0.0, n=0, name=name, meta=meta))
        a["ms"] += s.elapsed_time(e)
        a["n"] += 1
    total_ms = sum(a["ms"] for a in agg.values())
    table = []
    for key, a in sorted(agg.items(), key=lambda kv: -kv[1]["ms"]):
        b, fl = algorithmic(a["name"], a["meta"])
        mean_ms = a["ms"] / a["n"]
        table.append(dict(kernel=key, launches_per_step=a["n"] / steps, mean_us=round(mean_ms * 1e3, 2),
                          share=round(a["ms"] / total_ms, 4), alg_bytes=b, alg_flops=fl,
                          gbs=round(b / mean_ms / 1e6, 1) if mean_ms > 0 else None,
                          tflops=round(fl / mean_ms / 1e9, 2) if mean_ms > 0 and fl else None))
    top = table[0]
    intensity = top["alg_flops"] / max(top["alg_bytes"], 1)
    ridge = peaks["tflops"] * 1e12 / (peaks["hbm_gbs"] * 1e9)
    if top["alg_flops"] and intensity > ridge:
        roof = dict(bound="tensor", achieved=top["tflops"], peak=peaks["tflops"], unit="TFLOP/s",
                    frac=round(top["tflops"] / peaks["tflops"], 4))
    else:
        roof = dict(bound="hbm", achieved=top["gbs"], peak=peaks["hbm_gbs"], unit="GB/s",
                    frac=round(top["gbs"] / peaks["hbm_gbs"], 4))
    roof.update(kernel=top["kernel"], traffic=None, peak_source=peaks["source"], share_of_step=top["share"],
                alg_bytes_per_launch=top["alg_bytes"])
    return roof, table


# ------------------------------------------------------------------------------ CPU arm (oracle port)
def cpu_forward_rate(n_graphs: int, iters: int, warmup: int = 1):
    """The reference's CPU path for this workload: PyG is not installable here, so this is the
    oracle port (plain-torch restatement of PyG's gather/scatter operators hosting the
    reference's model control flow), all host threads.  Returns (graphs/s, threads, seconds)."""
    from hydra_gnn_b200 import synth
    from oracle import models as om  # the one place bench.py executes oracle/: the CPU baseline

    torch.set_num_threads(os.cpu_count() or 1)
    ref = synth.random_state_dict_(om.HeterogeneousNetwork(**M_GAT), seed=0).eval()
    data = synth.synth_mp3d_batch(range(n_graphs))
    with torch.no_grad():
        for _ in range(warmup):
            ref(data)
        t0 = time.perf_counter()
        for _ in range(iters):
            ref(data)
        dt = time.perf_counter() - t0
    return n_graphs * iters / dt, torch.get_num_threads(), dt


def run_reference(args, rank, world):
    if rank != 0:
        return
    sample = 512
    rates = []
    for _ in range(args.warmup):
        cpu_forward_rate(sample, 1, warmup=0)
    t0 = time.perf_counter()
    rate, threads, _ = cpu_forward_rate(sample, args.steps, warmup=0)
    dt = time.perf_counter() - t0
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sample / rate, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[1]: synthetic MP3D-shaped hetero GAT (M-GAT) forward", "graphs_per_step": sample},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{sample} graphs per step x {args.steps} steps of the same synthetic workload "
                                   "(oracle port of the PyG CPU path; PyG itself is not installable on this image)"},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------ GPU arm
def run_ours(args, rank, world, local_rank):
    import hydra_gnn_b200 as hg
    from hydra_gnn_b200 import _lib, ops, synth
    import torch.distributed as dist

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    G = args.graphs
    ids = range(rank * G, (rank + 1) * G)  # shard by contiguous graph-id range; no data-path exchange
    host = synth.synth_mp3d_batch(ids, pin=True)
    ref_sd = None
    net = hg.HeterogeneousNetwork(**M_GAT)
    synth.random_state_dict_(net, seed=0)
    net = net.eval().to(dev)
    net.plan_cache.enabled = False  # K1 (graph plan) is rebuilt every step: it is part of the path

    resident = synth.synth_mp3d_batch(ids, device=dev)
    n_rooms = resident["rooms"].x.size(0)
    h2d = host.nbytes()
    max_rooms = None
    if world > 1:
        cnt = torch.tensor([n_rooms], device=dev, dtype=torch.int64)
        dist.all_reduce(cnt, op=dist.ReduceOp.MAX)
        max_rooms = int(cnt.item())
        gather_buf = torch.empty(world * max_rooms, 26, device=dev)
        pad = torch.zeros(max_rooms, 26, device=dev)

    def step(data):
        logits = net(data)
        if world > 1:  # the only collective on the path: gather of the room logits (config 4)
            pad[:n_rooms].copy_(logits)
            dist.all_gather_into_tensor(gather_buf, pad)
        return logits

    def fence():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- device-resident arm (value): inputs already in HBM
    for _ in range(args.warmup):
        step(resident)
    fence()
    sink = []
    ops.set_profile_sink(sink)
    launches0 = _lib.launch_count
    with ClockSampler(local_rank) as clocks:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            step(resident)
        e1.record()
        fence()
    ops.set_profile_sink(None)
    launches = _lib.launch_count - launches0
    ms = e0.elapsed_time(e1)

    # ---- end-to-end arm (e2e): pinned host buffers -> H2D -> forward -> D2H of the logits
    out_host = torch.empty(n_rooms, 26, pin_memory=True)

    def e2e_step():
        d = hg.data.HeteroData()
        for t in host.node_types:
            d[t].x = host[t].x.to(dev, non_blocking=True)
        for t in host.edge_types:
            d[t].edge_index = host[t].edge_index.to(dev, non_blocking=True)
        out_host.copy_(step(d), non_blocking=True)

    h2d = sum(host[t].x.numel() * 4 for t in host.node_types) + sum(host[t].edge_index.numel() * 8 for t in host.edge_types)
    d2h = out_host.numel() * 4
    for _ in range(args.warmup):
        e2e_step()
    fence()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for _ in range(args.steps):
        e2e_step()
    f1.record()
    fence()
    ms_e2e = f0.elapsed_time(f1)

    if world > 1:  # max over ranks, measured on the device
        t = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e = float(t[0]), float(t[1])
    if rank != 0:
        return
    peaks = _peaks()
    roof, table = roofline_from(sink, args.steps, peaks)
    total_graphs = G * world
    line = {
        "metric": METRIC, "value": total_graphs * args.steps / (ms / 1e3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[1]: synthetic MP3D-shaped hetero GAT (M-GAT) forward, plan (K1) rebuilt every step",
                   "graphs_per_gpu": G, "objects": int(resident["objects"].x.size(0)), "rooms": int(n_rooms),
                   "edges": {et[1]: int(resident[et].edge_index.size(1)) for et in resident.edge_types},
                   "l2": "inputs larger than L2 (object features 297 MB + 390 MB projections per layer vs 126 MB L2)",
                   "projection_precision": ops.DEFAULT_PRECISION,
                   "collective": "all_gather_into_tensor(room logits)" if world > 1 else "none"},
        "e2e": {"value": total_graphs * args.steps / (ms_e2e / 1e3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": launches,
        "clocks": clocks.summary(),
        "roofline": roof,
        "kernels": table,
    }
    if world == 1 and not args.no_cpu:
        rate, threads, secs = cpu_forward_rate(args.cpu_graphs, args.cpu_iters)
        line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"{args.cpu_graphs} graphs x {args.cpu_iters} forwards of the same workload "
                                          f"({secs:.1f} s; oracle port of the PyG CPU path, all host threads)"}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--graphs", type=int, default=GRAPHS_PER_GPU, help="graphs per GPU per step")
    ap.add_argument("--precision", default=None, choices=[None, "fp32", "tf32x3"])
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-graphs", type=int, default=1024)
    ap.add_argument("--cpu-iters", type=int, default=3)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.precision:
        from hydra_gnn_b200 import ops
        ops.DEFAULT_PRECISION = args.precision
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        run_ours(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
