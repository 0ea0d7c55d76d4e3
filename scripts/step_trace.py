#!/usr/bin/env python
"""Per-step device times of the bench loop (diagnosis of one-off costs at the start of the timed region).
usage: torchrun ... scripts/step_trace.py [steps]   (or plain python for one GPU)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
import bench
from trajectorycrafter_b200 import Warper

world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
frames, depths, M, src = bench.build_workload(1234 + rank)
n = M.shape[0]
w = Warper(device=f"cuda:{local}")
fd, dd = frames.to(dev), depths.to(dev)
M, src = M.to(dev), src.to(dev)
if world > 1:
    t = torch.ones(1, device=dev); dist.all_reduce(t); dist.barrier(); torch.cuda.synchronize()
out = (torch.empty((n, 3, bench.H, bench.W), device=dev), torch.empty((n, 1, bench.H, bench.W), device=dev),
       torch.empty((n, 1, bench.H, bench.W), device=dev), torch.empty((n, 2, bench.H, bench.W), device=dev))
flag = torch.zeros(1, dtype=torch.int32, device=dev)
def step():
    w.forward_warp_matrices(fd, dd, M, src_index=src, per_frame_max=True, check_range=flag, out=out)
for _ in range(3):
    step()
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
host = []
ev[0].record()
for i in range(steps):
    t0 = time.perf_counter(); step(); host.append((time.perf_counter() - t0) * 1e3)
    ev[i + 1].record()
torch.cuda.synchronize()
gpu = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
if rank in (0, world - 1):
    print(f"rank {rank} gpu ms/step:", " ".join(f"{g:.2f}" for g in gpu), flush=True)
    print(f"rank {rank} host ms/call:", " ".join(f"{h:.2f}" for h in host), flush=True)
if world > 1:
    dist.barrier(); dist.destroy_process_group()
