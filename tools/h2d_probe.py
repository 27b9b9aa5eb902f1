"""Aggregate pinned host -> device copy bandwidth of N ranks copying at the same time (the ceiling of bench.py's e2e arm:
each step moves 321.7 MB of fp32 inputs per GPU).  torchrun --nproc-per-node N tools/h2d_probe.py"""
import json, os
import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
src = torch.empty(320 << 20, dtype=torch.uint8, pin_memory=True)  # one step's worth
dst = torch.empty_like(src, device=dev)
for _ in range(3):
    dst.copy_(src, non_blocking=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(20):
    dst.copy_(src, non_blocking=True)
e.record()
torch.cuda.synchronize()
gbs = torch.tensor([20 * src.numel() / s.elapsed_time(e) / 1e6], device=dev, dtype=torch.float64)
every = [torch.zeros_like(gbs) for _ in range(world)]
if world > 1:
    dist.all_gather(every, gbs)
else:
    every = [gbs]
if rank == 0:
    per = [round(float(t), 1) for t in every]
    print(json.dumps({"n_gpus": world, "h2d_gbs_per_rank": per, "h2d_gbs_aggregate": round(sum(per), 1),
                      "cpu_affinity": sorted(os.sched_getaffinity(0))[:4] + ["..."], "cpus": os.cpu_count()}))
if world > 1:
    dist.destroy_process_group()
