"""Batch-1 latency of the M-GAT forward (the room_classification_server operating point):
wall clock per forward incl. synchronisation -- eager, CUDA-graph replay, and the CPU oracle port.

    python tools/latency.py > profiles/r01_latency.json
"""
import json, os, statistics, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import hydra_gnn_b200 as hg
from hydra_gnn_b200 import synth
from bench import M_GAT


def wall(fn, n=200, warm=20):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ts.append((time.perf_counter() - t0) * 1e6)
    return dict(median_us=round(statistics.median(ts), 1), p10_us=round(sorted(ts)[n // 10], 1),
                p90_us=round(sorted(ts)[9 * n // 10], 1))


def main():
    out = {}
    net = synth.random_state_dict_(hg.HeterogeneousNetwork(**M_GAT), seed=0).eval().cuda()
    for n_graphs in (1, 16):
        data = synth.synth_mp3d_batch(range(n_graphs), device="cuda")
        net.plan_cache.enabled = False
        r = {"eager": wall(lambda: net(data))}
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            net(data)
            s.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                logits = net(data)
        torch.cuda.synchronize()
        r["cuda_graph_replay"] = wall(g.replay)
        try:
            from oracle import models as om  # CPU baseline only (test infrastructure)
            ref = synth.random_state_dict_(om.HeterogeneousNetwork(**M_GAT), seed=0).eval()
            cpu = synth.synth_mp3d_batch(range(n_graphs))
            torch.set_num_threads(os.cpu_count() or 1)
            with torch.no_grad():
                for _ in range(5):
                    ref(cpu)
                ts = []
                for _ in range(30):
                    t0 = time.perf_counter()
                    ref(cpu)
                    ts.append((time.perf_counter() - t0) * 1e6)
            r["cpu_oracle_port"] = dict(median_us=round(statistics.median(ts), 1), threads=torch.get_num_threads())
        except ImportError:
            pass
        out[f"graphs={n_graphs}"] = r
    print(json.dumps(out))


if __name__ == "__main__":
    main()
