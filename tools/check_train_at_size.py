"""Config-5-size gradient parity: M-GAT loss + every live parameter gradient on 8192 graphs, product (CUDA) vs the CPU
oracle's autograd; prints the worst error relative to the gradient scale (tests/test_gpu_train.py uses the same rule)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import hydra_gnn_b200 as hg
from hydra_gnn_b200 import synth
from oracle import models as om
from util import clone_data

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
kw = dict(input_dim_dict={"objects": 306, "rooms": 6}, output_dim=26, conv_block="GAT", GAT_hidden_dims=[64, 64],
          GAT_heads=[3, 3, 3], GAT_concats=[False, False, False], dropout=0.0)
host = synth.synth_mp3d_batch(range(n))
labels = host["rooms"].y.long(); mask = labels != 25
ref = synth.random_state_dict_(om.HeterogeneousNetwork(**kw), seed=0).train()
net = hg.HeterogeneousNetwork(**kw); net.load_state_dict(ref.state_dict()); net = net.train().cuda()
t0 = time.perf_counter()
lo = ref(clone_data(host)); loss_o = ref.loss(lo, labels, mask); loss_o.backward()
t1 = time.perf_counter()
lg = net(clone_data(host, "cuda")); loss_g = net.loss(lg, labels.cuda(), mask.cuda()); loss_g.backward()
torch.cuda.synchronize()
print(f"oracle {t1 - t0:.1f} s; loss {float(loss_o):.6f} vs {float(loss_g):.6f}; logits max abs err "
      f"{float((lg.detach().cpu() - lo.detach()).abs().max()):.3e} / scale {float(lo.detach().abs().max()):.3e}")
# the arbiter: the same oracle in float64 -- the errors of BOTH fp32 computations are measured against it
import copy
ref64 = copy.deepcopy(ref).double()
ref64.zero_grad(set_to_none=True)
d64 = clone_data(host)
for t in d64.node_types:
    for k, v in list(d64[t].items()):
        if torch.is_tensor(v) and v.dtype == torch.float32:
            d64[t][k] = v.double()
l64 = ref64(d64); loss64 = ref64.loss(l64, labels, mask); loss64.backward()
p64 = dict(ref64.named_parameters())
print(f"logits vs float64: product {float((lg.detach().cpu().double() - l64.detach()).abs().max()):.3e}, "
      f"fp32 oracle {float((lo.detach().double() - l64.detach()).abs().max()):.3e} (scale {float(l64.detach().abs().max()):.3e})")
ref_p, got_p = dict(ref.named_parameters()), dict(net.named_parameters())
floor = 1e-3 * max(float(p.grad.abs().max()) for p in ref_p.values() if p.grad is not None)
worst = worst64 = worst_o = 0.0
for name, p in ref_p.items():
    if p.grad is None:
        continue
    g = got_p[name].grad.detach().float().cpu()
    scale = max(float(p.grad.abs().max()), floor)
    rel = float((g - p.grad).abs().max()) / scale
    worst = max(worst, rel)
    e_prod = float((g.double() - p64[name].grad).abs().max()) / scale
    e_orac = float((p.grad.double() - p64[name].grad).abs().max()) / scale
    worst64 = max(worst64, e_prod); worst_o = max(worst_o, e_orac)
    if rel > 2e-4:
        print(f"  {name}: product vs fp32 oracle {rel:.2e}; vs float64: product {e_prod:.2e}, fp32 oracle {e_orac:.2e}")
print(f"worst gradient error relative to scale: product vs fp32 oracle {worst:.3e}; vs float64: product {worst64:.3e}, "
      f"fp32 oracle {worst_o:.3e}")
