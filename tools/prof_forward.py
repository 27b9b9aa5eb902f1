"""Short profiling target: a few M-GAT forwards at the bench shape (for ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hydra_gnn_b200 as hg
from hydra_gnn_b200 import synth

G = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
kw = dict(input_dim_dict={"objects": 306, "rooms": 6}, output_dim=26, conv_block="GAT", GAT_hidden_dims=[64, 64],
          GAT_heads=[3, 3, 3], GAT_concats=[False] * 3, dropout=0.4)
net = synth.random_state_dict_(hg.HeterogeneousNetwork(**kw), seed=0).eval().cuda()
net.plan_cache.enabled = False
data = synth.synth_mp3d_batch(range(G), device="cuda")
for _ in range(iters):
    out = net(data)
torch.cuda.synchronize()
print("ok", tuple(out.shape))
