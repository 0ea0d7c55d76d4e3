"""Device time of each of config 2's eight orbit presets (49 units each) on one GPU: how uneven a contiguous 8-way shard is."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from trajectorycrafter_b200 import Warper

frames, depths, M, src = bench.build_workload(1234)[:4]
dev = torch.device("cuda:0")
wp = Warper(device="cuda:0")
f, d = frames.to(dev), depths.to(dev)
flag = torch.zeros(1, dtype=torch.int32, device=dev)
n = M.shape[0] // 8
for order, name in ((None, "contiguous (one preset per shard)"), ("interleaved", "interleaved (unit u -> shard u % 8)")):
    ts = []
    for p in range(8):
        idx = torch.arange(p * n, (p + 1) * n) if order is None else torch.arange(p, M.shape[0], 8)
        Mp, sp = M[idx].contiguous().to(dev), src[idx].to(dev)
        out = wp.forward_warp_matrices(f, d, Mp, src_index=sp, check_range=flag)
        for _ in range(3):
            wp.forward_warp_matrices(f, d, Mp, src_index=sp, check_range=flag, out=out)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10):
            wp.forward_warp_matrices(f, d, Mp, src_index=sp, check_range=flag, out=out)
        b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) / 10)
    print(name, " ".join(f"{t:.3f}" for t in ts), "ms; max", f"{max(ts):.3f}", "mean", f"{sum(ts)/8:.3f}")
