"""torchrun --nproc-per-node N tools/check_peer_gather.py: the pipelined logit gather over NVLink peer memory (copy-engine
pulls) returns exactly what the NCCL all_gather_into_tensor path returns, ragged row counts, several steps in flight."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from hydra_gnn_b200 import dist as hdist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)
rows = 20000 + 37 * rank
g = torch.Generator(device=dev).manual_seed(rank)
hdist.USE_PEER_COPIES = True
peer = hdist.LogitGather(26).prepare(rows, dev, slots=2)
hdist.USE_PEER_COPIES = False
nccl = hdist.LogitGather(26).prepare(rows, dev, slots=2)
ok = peer._peer is not None
for step in range(6):
    k = step & 1
    x = torch.randn(rows, 26, device=dev, generator=g)
    for G in (peer, nccl):
        G.wait(k)
        G.stage(x, k)
        G.launch(k)
    peer.wait(k); nccl.wait(k)
    torch.cuda.synchronize()
    ok = ok and torch.equal(peer.unpad(k), nccl.unpad(k))
flag = torch.tensor([int(ok)], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("peer gather", "OK" if int(flag) else "MISMATCH / unavailable", "peer path:", peer._peer is not None, flush=True)
torch.cuda.synchronize()
dist.barrier()
dist.destroy_process_group()
os._exit(0 if int(flag) else 1)
