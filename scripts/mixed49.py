"""Device time of one balanced 8-way shard of config 2 (49 units: a frame block of every orbit preset) on one GPU, for schedule
sweeps of small submissions (TCW_LANES / TCW_CHUNK_UNITS from the env).  usage: python scripts/mixed49.py [world=8] [rank=3]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from trajectorycrafter_b200 import Warper
from trajectorycrafter_b200.sharding import balanced_units, sources_of

world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
rank = int(sys.argv[2]) if len(sys.argv) > 2 else 3
frames, depths, M, src = bench.build_workload(1234)[:4]
dev = torch.device("cuda:0")
wp = Warper(device="cuda:0")
ids = balanced_units(M.shape[0], 49, world, rank)
needed, lsrc = sources_of(src, ids)
f, d = frames[needed].contiguous().to(dev), depths[needed].contiguous().to(dev)
Mp, sp = M[ids].contiguous().to(dev), lsrc.to(dev)
flag = torch.zeros(1, dtype=torch.int32, device=dev)
out = wp.forward_warp_matrices(f, d, Mp, src_index=sp, check_range=flag)
for _ in range(4):
    wp.forward_warp_matrices(f, d, Mp, src_index=sp, check_range=flag, out=out)
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(20):
    wp.forward_warp_matrices(f, d, Mp, src_index=sp, check_range=flag, out=out)
b.record(); torch.cuda.synchronize()
print(f"{ids.numel()} mixed units: {a.elapsed_time(b) / 20:.3f} ms")
