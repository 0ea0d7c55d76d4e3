#!/usr/bin/env python
"""How the two accumulation modes cope with a ROUGH depth map: the three full-size units of tests/golden/config1_p7.npz (frames of the
reference's fixture clip, luma-derived proxy depth), each repeated 16 times = 48 units per call, device-timed."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from _util import load_config1
from trajectorycrafter_b200 import Warper, prepare_matrices
frames, depths, z = load_config1()
M = prepare_matrices(torch.from_numpy(z["T1"]), torch.from_numpy(z["T2"]), torch.from_numpy(z["K"]), None)
w = Warper(device="cuda:0")
dev = w.device
for u in range(3):
    Mu = M[u:u + 1].repeat(16, 1).to(dev)
    src = torch.full((16,), u, dtype=torch.int32, device=dev)
    fd, dd = frames.to(dev), depths.to(dev)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    for mode in ("window", "atomic"):
        out = None
        for _ in range(3):
            out = w.forward_warp_matrices(fd, dd, Mu, src_index=src, check_range=flag, accumulation=mode)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            w.forward_warp_matrices(fd, dd, Mu, src_index=src, check_range=flag, accumulation=mode)
        b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 5
        print(f"frame {int(z['pick'][u]):2d} valid {float(out[1].mean()):.3f} {mode:7s} {ms:7.3f} ms per 16 units = {16 * 384 * 672 / ms / 1e6:6.2f} Gpix/s  status {int(flag.item())}", flush=True)
