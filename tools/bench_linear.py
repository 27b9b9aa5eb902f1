"""Times hg_linear alone on given shapes: python tools/bench_linear.py M,N,K [M,N,K ...]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hydra_gnn_b200 import ops

shapes = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]] or [(242968, 396, 306), (242968, 204, 64), (20277, 396, 64)]
for M, N, K in shapes:
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn(M, K, device="cuda", generator=g)
    w = torch.randn(N, K, device="cuda", generator=g) / K ** 0.5
    cache = {}
    y = ops.linear([(x, w)], N, pack_cache=cache)
    ref = (x.double() @ w.double().t()).float()
    err = (y - ref).abs().max().item()
    flush = torch.empty(64 << 20, dtype=torch.float32, device="cuda")
    ts = []
    for _ in range(12):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        ops.linear([(x, w)], N, pack_cache=cache)
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    ts = sorted(ts)[1:-1]
    us = sum(ts) / len(ts)
    gb = 4 * (M * K + N * K + M * N) / 1e9
    print(f"M={M} N={N} K={K}: {us:8.1f} us  {gb / us * 1e6:7.0f} GB/s  {2 * M * N * K / us / 1e6:6.1f} TFLOP/s  max|err|={err:.2e}")
