#!/usr/bin/env python
"""Small fixed program for ncu: config-2 units through Warper.forward_warp_matrices, a few calls, nothing else.
usage: python scripts/ncu_target.py [n_units=48] [calls=3] [accumulation=window] [first_unit=160]   (TCW_LANES / TCW_CHUNK_UNITS / TCW_NO_GRAPH from the env)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import build_workload, H, W
from trajectorycrafter_b200 import Warper

n = int(sys.argv[1]) if len(sys.argv) > 1 else 48
calls = int(sys.argv[2]) if len(sys.argv) > 2 else 3
acc = sys.argv[3] if len(sys.argv) > 3 else "window"
off = int(sys.argv[4]) if len(sys.argv) > 4 else 160          # first unit (default: frame 20, where the orbits have moved)
frames, depths, M, src = build_workload(1234)
dev = torch.device("cuda", 0)
fd, dd, Md, sd = frames.to(dev), depths.to(dev), M[off:off + n].contiguous().to(dev), src[off:off + n].contiguous().to(dev)
out = (torch.empty((n, 3, H, W), device=dev), torch.empty((n, 1, H, W), device=dev),
       torch.empty((n, 1, H, W), device=dev), torch.empty((n, 2, H, W), device=dev))
flag = torch.zeros(1, dtype=torch.int32, device=dev)
w = Warper(device="cuda:0")
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(calls):
    ev0.record()
    w.forward_warp_matrices(fd, dd, Md, src_index=sd, per_frame_max=True, check_range=flag, out=out, accumulation=acc)
    ev1.record()
    torch.cuda.synchronize()
    print(f"call {i}: {ev0.elapsed_time(ev1):.3f} ms for {n} units, flag {int(flag.item())}")
