"""Timeline of one eager M-GAT forward with the side streams ON (plan and rooms-side chains forked as in the value arm):
start / end of every ABI call relative to the first launch, from the CUDA-event pairs of ops.set_profile_sink.
Eager launch gaps are larger than under CUDA-graph replay; what it shows is which chain ends last."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hydra_gnn_b200 as hg
from hydra_gnn_b200 import ops, synth
from bench import M_GAT, kernel_key

net = synth.random_state_dict_(hg.HeterogeneousNetwork(**M_GAT), seed=0).eval().cuda()
net.plan_cache.enabled = False
data = synth.synth_mp3d_batch(range(int(sys.argv[1]) if len(sys.argv) > 1 else 4096), device="cuda")
for _ in range(3):
    net(data)
torch.cuda.synchronize()
sink = []
base = torch.cuda.Event(enable_timing=True)
base.record()
ops.set_profile_sink(sink)
net(data)
torch.cuda.synchronize()
ops.set_profile_sink(None)
rows = sorted((base.elapsed_time(s) * 1e3, base.elapsed_time(e) * 1e3, kernel_key(n, m)) for n, m, s, e in sink)
t0 = rows[0][0]
for s, e, k in rows:
    print(f"{s - t0:8.1f} {e - t0:8.1f}  {e - s:7.1f} us  {k}")
