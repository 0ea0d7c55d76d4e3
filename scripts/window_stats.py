#!/usr/bin/env python
"""Per-strip target bounding boxes on config 2 (CPU model): how large a shared-memory window does a warp's
source strip need, and how many contributions fall outside a fixed window?  Used to size the windows of
csrc/warp_window.cuh.  usage: python scripts/window_stats.py [step]"""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

frames, depths, M, src = bench.build_workload(1234)
H, W = bench.H, bench.W
step = int(sys.argv[1]) if len(sys.argv) > 1 else 11
ys, xs = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
shapes = [(64, 8), (32, 8), (64, 4), (32, 16), (64, 16)]
wins = {(64, 8): [(80, 12), (96, 12), (80, 16), (96, 16)], (32, 8): [(48, 12), (40, 12), (48, 16)], (64, 4): [(80, 8), (96, 8)],
        (32, 16): [(48, 20), (48, 24)], (64, 16): [(80, 20), (96, 24)]}
res = {}
for u in range(0, M.shape[0], step):
    m = M[u]; K1i = m[:9].view(3, 3); T = m[9:21].view(3, 4); K2 = m[21:].view(3, 3)
    d = depths[src[u], 0]
    p = torch.stack([xs, ys, torch.ones_like(xs)], -1)
    X = (p @ K1i.T) * d[..., None]
    Y = X @ T[:, :3].T + T[:, 3]
    P = Y @ K2.T
    beh = Y[..., 2] <= 0.01
    ox = P[..., 0] / P[..., 2] + 1; oy = P[..., 1] / P[..., 2] + 1
    ix = torch.floor(ox).long(); iy = torch.floor(oy).long()
    live = (~beh) & (ox > 0) & (ox < W + 1) & (oy > 0) & (oy < H + 1)
    for (sw, sh) in shapes:
        nsx, nsy = (W + sw - 1) // sw, (H + sh - 1) // sh
        sid = ((torch.arange(H) // sh)[:, None] * nsx + (torch.arange(W) // sw)[None, :])
        big = 1 << 20
        l = live
        s = sid[l]; x = ix[l]; y = iy[l]
        ns = nsx * nsy
        xmin = torch.full((ns,), big).scatter_reduce(0, s, x, "amin"); xmax = torch.full((ns,), -big).scatter_reduce(0, s, x + 1, "amax")
        ymin = torch.full((ns,), big).scatter_reduce(0, s, y, "amin"); ymax = torch.full((ns,), -big).scatter_reduce(0, s, y + 1, "amax")
        cnt = torch.zeros(ns, dtype=torch.long).scatter_add(0, s, torch.ones_like(s))
        ok = cnt > 0
        bw = (xmax - xmin + 1)[ok]; bh = (ymax - ymin + 1)[ok]
        # median-anchored window: how many live px have all corners inside a window centred on the strip's median cell
        r = res.setdefault((sw, sh), dict(strips=0, px=0, live=0, bw=[], bh=[], fit={}, out={}))
        r["strips"] += int(ok.sum()); r["px"] += H * W; r["live"] += int(l.sum())
        r["bw"].append(bw.numpy()); r["bh"].append(bh.numpy())
        for (ww, wh) in wins[(sw, sh)]:
            fit = (bw <= ww) & (bh <= wh)
            r["fit"][(ww, wh)] = r["fit"].get((ww, wh), 0) + int(fit.sum())
            # anchor = bbox min of the FIRST live row's... approximate with centre of bbox-clamped median
            xmed = torch.zeros(ns); ymed = torch.zeros(ns)
            # use mean as a cheap stand-in for the sampled estimate
            sx_ = torch.zeros(ns).scatter_add(0, s, x.float()) / cnt.clamp(min=1); sy_ = torch.zeros(ns).scatter_add(0, s, y.float()) / cnt.clamp(min=1)
            x0 = (sx_ - ww / 2).floor().long(); y0 = (sy_ - wh / 2).floor().long()
            inside = (x >= x0[s]) & (x + 1 < x0[s] + ww) & (y >= y0[s]) & (y + 1 < y0[s] + wh)
            r["out"][(ww, wh)] = r["out"].get((ww, wh), 0) + int((~inside).sum())
for k, r in res.items():
    bw = np.concatenate(r["bw"]); bh = np.concatenate(r["bh"])
    print(f"strip {k}: strips {r['strips']}, live {r['live'] / r['px']:.3f}; bbox w p50/p90/p99/max {np.percentile(bw, [50, 90, 99]).astype(int)} {bw.max()}  h {np.percentile(bh, [50, 90, 99]).astype(int)} {bh.max()}  mean area/strip px {np.mean(bw * bh) / (k[0] * k[1]):.2f}")
    for wk in r["fit"]:
        print(f"    window {wk}: strips that fit {r['fit'][wk] / r['strips']:.4f}; live px outside a mean-centred window {r['out'][wk] / r['live']:.5f}")
