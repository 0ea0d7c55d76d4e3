#!/usr/bin/env python
"""Small fixed program for compute-sanitizer: the default forward warp on the smoke shape and on an odd-size case with
mask1 holes, behind-camera pixels and mask=True; the generic splat; the point-cloud z-buffer.  Prints OK at the end."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from trajectorycrafter_b200 import GlobalPointCloudWarper, Warper, prepare_matrices
from trajectorycrafter_b200.workloads import orbit_units, synthetic_clip

w = Warper(device="cuda:0")
for (h, wd, nf, presets) in ((96, 160, 2, [(0, 30), (10, -45)]), (75, 131, 2, [(0, 60), (25, 170)])):
    frames, depths = synthetic_clip(nf, h, wd, seed=3)
    T1, T2, K, src = orbit_units(depths, h, wd, presets=presets)
    M = prepare_matrices(T1, T2, K)
    g = torch.Generator().manual_seed(1)
    mask1 = (torch.rand((nf, 1, h, wd), generator=g) > 0.2).float()
    for acc in ("window", "atomic"):
        out = w.forward_warp_matrices(frames, depths, M, src_index=src, mask1=mask1, accumulation=acc)
    out = w.forward_warp_matrices(frames, depths, M[:1], src_index=src[:1], mask=True)
    fl = out[3][:1].float()
    w.bilinear_splatting(frames[:1], None, depths[:1, 0], fl.cpu(), None, is_image=True)
    torch.cuda.synchronize()
    print("warp ok", h, wd)
pc = GlobalPointCloudWarper(device="cuda:0")
P = torch.rand((20000, 3)) * 4 - 2 + torch.tensor([0.0, 0.0, 5.0])
C = torch.rand((20000, 3))
f, m = pc.render_pointcloud_zbuffer_vectorized_point_size(P, C, torch.eye(4)[None], torch.tensor([[80.0, 0, 48], [0, 80.0, 30], [0, 0, 1]])[None], (60, 96), 3)
torch.cuda.synchronize()
print("OK")
