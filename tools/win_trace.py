"""Per-tile timeline of the window kernel's three roles in CTA 0 (debug build: tools/ab_build.sh trace gat_fused.cu -DHG_WIN_TRACE,
run with HYDRA_GNN_B200_SO=hydra_gnn_b200/ab/lib_trace.so).  Cycles relative to the first stamp.
  0/1 window warp: before / after the wait for a free window stage      2/3 edge producer: before / after the wait for its stage
  4 staging stores done   5 softmax done (stage published)
  6 consumer warp 0 starts waiting   7 edge stage full   8 window full   9 gather done   10 epilogue done"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from hydra_gnn_b200 import _lib, ops
from test_gpu_ops import _scene_batch_relations

g = torch.Generator().manual_seed(0)
rels, n_o = _scene_batch_relations(range(4096), 3, g)
for _ in range(3):
    out = ops.gat_fused(rels, n_o, 3, 64, False, act="elu")
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record(); out = ops.gat_fused(rels, n_o, 3, 64, False, act="elu"); e.record(); torch.cuda.synchronize()
print("kernel us", s.elapsed_time(e) * 1e3)
h = _lib.lib()
n_t, n_e = 64, 12
buf = (ctypes.c_longlong * (n_t * n_e))()
fn = h.hg_debug_win_trace
fn.argtypes, fn.restype = [ctypes.c_void_p, ctypes.c_int], ctypes.c_int
assert fn(buf, n_t * n_e) == 0
ev = [[buf[k * n_e + j] for j in range(n_e)] for k in range(n_t)]
t0 = min(v for row in ev for v in row if v > 0)
print("tile | win wait..issue | prod wait_start wait_end staged softmax | cons wait_start edge_full win_full gather_done epi_done")
for k in range(52):
    r = [(v - t0) if v > 0 else -1 for v in ev[k]]
    print(f"{k:3d} | {r[0]:7d} {r[1]:7d} | {r[2]:7d} {r[3]:7d} {r[4]:7d} {r[5]:7d} | {r[6]:7d} {r[7]:7d} {r[8]:7d} {r[9]:7d} {r[10]:7d}"
          f" || prod {r[5]-r[3]:6d} (stage {r[4]-r[3]:6d} sm {r[5]-r[4]:5d})  cons wait {r[8]-r[6]:6d} gather {r[9]-r[8]:6d} epi {r[10]-r[9]:5d}")
