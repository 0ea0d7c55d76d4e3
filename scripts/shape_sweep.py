#!/usr/bin/env python
"""Device-timed Gpix/s of the batched warp on the other shapes SURVEY.md §8d names (not the bench metric):
config 1-like gradual clip, the reference's own 576x1024 warp resolution, config 3 (1080p), config 4 (ring of 16
target cameras), each with resident inputs, 3 warm-up + 10 timed submissions, CUDA events.
usage: python scripts/shape_sweep.py   (one B200)"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from trajectorycrafter_b200 import Warper, prepare_matrices
from trajectorycrafter_b200.workloads import synthetic_clip, gradual_units, orbit_units, ORBIT_PRESETS

dev = torch.device("cuda:0")
w = Warper(device="cuda:0")
PEAK = 6531.3
if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")):
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])


def run(name, frames, depths, T1, T2, K, src):
    M = prepare_matrices(T1, T2, K).to(dev)
    n, (h, wd) = M.shape[0], frames.shape[-2:]
    fd, dd = frames.to(dev), depths.to(dev)
    sd = None if src is None else src.to(dev)
    out = (torch.empty((n, 3, h, wd), device=dev), torch.empty((n, 1, h, wd), device=dev),
           torch.empty((n, 1, h, wd), device=dev), torch.empty((n, 2, h, wd), device=dev))
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    step = lambda: w.forward_warp_matrices(fd, dd, M, src_index=sd, per_frame_max=True, check_range=flag, out=out)
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    px = n * h * wd
    print(json.dumps({"case": name, "units": n, "h": h, "w": wd, "ms": round(ms, 4), "gpix_s": round(px / ms / 1e6, 2),
                      "frac_of_hbm_roofline": round(44.0 * px / (ms * 1e-3) / 1e9 / PEAK, 4), "range_flag": int(flag.item())}), flush=True)
    del out, fd, dd


for (name, nf, h, wd, mode) in [("gradual 49x384x672 (one trajectory)", 49, 384, 672, "gradual"),
                                ("orbits 392 units 384x672 (bench metric)", 49, 384, 672, "orbit"),
                                ("gradual 49x576x1024 (reference warp resolution)", 49, 576, 1024, "gradual"),
                                ("orbits 392 units 576x1024", 49, 576, 1024, "orbit"),
                                ("gradual 49x1080x1920 (config 3)", 49, 1080, 1920, "gradual")]:
    frames, depths = synthetic_clip(nf, h, wd, seed=1234)
    if mode == "gradual":
        T1, T2, K = gradual_units(depths, (0.0, 30.0, 0.3, 0.0, 0.0), h, wd)
        run(name, frames, depths, T1, T2, K, None)
    else:
        T1, T2, K, src = orbit_units(depths, h, wd)
        run(name, frames, depths, T1, T2, K, src)
