"""Device generator timing (hg_synth_mp3d_counts / _fill) for one 16,384-graph micro-batch: CUDA events per ABI call."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hydra_gnn_b200 import ops, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
gen = synth.DeviceSynth("cuda")
sizes = gen.sizes(0, n)
for _ in range(3):
    gen.counts_async(0, n)
    d = gen.batch(0, n, sizes, labels=False)
torch.cuda.synchronize()
sink = []
ops.set_profile_sink(sink)
for _ in range(5):
    gen.counts_async(0, n)
    d = gen.batch(0, n, sizes, labels=False)
torch.cuda.synchronize()
ops.set_profile_sink(None)
by = {}
for name, meta, s, e in sink:
    by.setdefault(name, []).append(s.elapsed_time(e) * 1e3)
nbytes = d["objects"].x.numel() * 4 + sum(d[et].edge_index.numel() * 8 for et in d.edge_types)
for k, v in by.items():
    v.sort()
    print(f"{k:24s} median {v[len(v) // 2]:8.1f} us")
fill = sorted(by["hg_synth_mp3d_fill"])[2]
print(f"fill writes {nbytes / 1e6:.1f} MB -> {nbytes / fill / 1e3:.0f} GB/s; sizes {sizes}")
