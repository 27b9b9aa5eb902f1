"""One hg_linear_wgrad call at the config-5 layer-1 shape (for ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hydra_gnn_b200 import ops
M, N, K = 488696, 396, 306
dy = torch.randn(M, N, device="cuda"); x = torch.randn(M, K, device="cuda")
for _ in range(3):
    ops.linear_wgrad(dy, x)
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(5):
    ops.linear_wgrad(dy, x)
e.record(); torch.cuda.synchronize()
print("wgrad ms", s.elapsed_time(e) / 5)
