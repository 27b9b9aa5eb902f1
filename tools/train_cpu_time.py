"""CPU issue time vs GPU time of the training step (is the step launch-bound?)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
import hydra_gnn_b200 as hg
from hydra_gnn_b200 import synth
from hydra_gnn_b200.train import FlatAdam
from bench import M_GAT
G = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
dev = torch.device("cuda", 0)
host = synth.synth_mp3d_batch(range(G))
data = hg.data.HeteroData()
for t in host.node_types: data[t].x = host[t].x.to(dev)
for t in host.edge_types: data[t].edge_index = host[t].edge_index.to(dev)
labels = host["rooms"].y.long().to(dev)
keep = (labels != 25).nonzero().view(-1)
net = synth.random_state_dict_(hg.HeterogeneousNetwork(**dict(M_GAT, dropout=0.4)), seed=0).train().to(dev)
net.plan_cache.enabled = False
opt = FlatAdam(net.parameters(), lr=1e-3, weight_decay=1e-3)
def step():
    opt.zero_grad()
    loss = F.cross_entropy(net(data).index_select(0, keep), labels.index_select(0, keep), reduction="sum") / keep.numel()
    loss.backward(); opt.step()
for _ in range(5): step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20): step()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"graphs {G}: CPU issue {1e3*(t1-t0)/20:.2f} ms/step, wall with sync {1e3*(t2-t0)/20:.2f} ms/step")
