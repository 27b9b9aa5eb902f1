"""Times hg_gat_fused at the bench's objects-layer shape in isolation (no stream concurrency)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hydra_gnn_b200 as hg
from hydra_gnn_b200 import ops, synth
from bench import M_GAT, roofline_from, _peaks
ops.PARALLEL_BRANCHES = False
net = synth.random_state_dict_(hg.HeterogeneousNetwork(**M_GAT), seed=0).eval().cuda()
net.plan_cache.enabled = False
net.plan_cache.overlap = False
data = synth.synth_mp3d_batch(range(int(sys.argv[1]) if len(sys.argv) > 1 else 4096), device="cuda")
for _ in range(3):
    net(data)
sink = []
ops.set_profile_sink(sink)
for _ in range(10):
    net(data)
torch.cuda.synchronize()
ops.set_profile_sink(None)
_, table = roofline_from(sink, 10, _peaks())
for k in table:
    print(f"{k['kernel']:60s} x{k['launches_per_step']:.0f} {k['mean_us']:8.1f} us {k['gbs']} GB/s")
