#!/usr/bin/env python
"""How many accumulator cells does the splat kernel emit per source pixel on config 2, against the ideal (distinct
cells per 64 x 8 strip / per unit)?  CPU-only model of the register-level merge of csrc/warp_pair.cuh, following the
kernel's routing row by row (ids only, no values): rules 1 = straight folds only (TCW_MERGE_RULES=1), rules 2 = straight
+ diagonal + same-column folds (the default).  Interior cells only (borders are approximate).
usage: python scripts/emit_stats.py   (~1-2 min; samples every 11th unit)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

frames, depths, M, src = bench.build_workload(1234)
H, W = bench.H, bench.W
SH, SW_ = 8, 64
n = M.shape[0]
ys, xs = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
tot = {}


def acc(k, v):
    tot[k] = tot.get(k, 0) + float(v)


def simulate(NW, NE, SW, SE, rules):
    """ids: (H, W) int64, dead pixels carry unique negative ids.  Returns the number of emitted cells."""
    A, B = slice(0, None, 2), slice(1, None, 2)
    nl = W // 2
    first = (torch.arange(nl) % 32) == 0                 # lane 0 of a warp receives lane 31's cells of ANOTHER strip row
    emits = 0
    cL = torch.full((nl,), -1, dtype=torch.long); cM = cL.clone()
    for y in range(H):
        if y % SH == 0:
            emits += int((cL >= 0).sum() + (cM >= 0).sum())
            cL = torch.full((nl,), -1, dtype=torch.long); cM = cL.clone()
        nwa, nea, swa, sea = NW[y, A], NE[y, A], SW[y, A], SE[y, A]
        nwb, neb, swb, seb = NW[y, B], NE[y, B], SW[y, B], SE[y, B]
        inNE = torch.roll(neb, 1); inSE = torch.roll(seb, 1)
        # lane 0 of each warp: the rotate wraps inside the warp, i.e. it gets lane 31's cells of the same strip
        idx0 = torch.nonzero(first).flatten()
        inNE[idx0] = neb[torch.clamp(idx0 + 31, max=nl - 1)]; inSE[idx0] = seb[torch.clamp(idx0 + 31, max=nl - 1)]
        if rules == 1:
            s1 = inNE != nwa; s2 = inSE != swa; s3 = nea != nwb; s4 = sea != swb
            fL = cL == nwa; fM = cM == nwb
            emits += int(((inNE >= 0) & s1).sum() + ((inSE >= 0) & s2).sum() + ((nea >= 0) & s3).sum() + ((sea >= 0) & s4).sum())
            emits += int(((cL >= 0) & ~fL).sum() + ((cM >= 0) & ~fM).sum())
            emits += int((nwa >= 0).sum() + (nwb >= 0).sum())
            cL, cM = swa.clone(), swb.clone()
        else:
            s1 = ~((inNE == nwa) | (inNE == swa) | (inNE == nea))
            s2 = ~((inSE == swa) | (inSE == nwa) | (inSE == sea))
            s3 = ~((nea == nwb) | (nea == swb))
            s4 = ~((sea == swb) | (sea == nwb))
            eL = ~((cL == nwa) | (cL == swa)); eM = ~((cM == nwb) | (cM == swb))
            sT = nwb == nwa; sB = swb == swa
            emits += int(((inNE >= 0) & s1).sum() + ((inSE >= 0) & s2).sum() + ((nea >= 0) & s3).sum() + ((sea >= 0) & s4).sum())
            emits += int(((cL >= 0) & eL).sum() + ((cM >= 0) & eM).sum())
            emits += int((nwa >= 0).sum() + ((nwb >= 0) & ~sT).sum())
            cL = swa.clone(); cM = torch.where(sB, torch.full_like(swb, -1), swb)
    emits += int((cL >= 0).sum() + (cM >= 0).sum())
    return emits


BIG = 1 << 40
for u in range(0, n, 11):
    m = M[u]; K1i = m[:9].view(3, 3); T = m[9:21].view(3, 4); K2 = m[21:].view(3, 3)
    d = depths[src[u], 0]
    p = torch.stack([xs, ys, torch.ones_like(xs)], -1)
    X = (p @ K1i.T) * d[..., None]
    Y = X @ T[:, :3].T + T[:, 3]
    P = Y @ K2.T
    beh = Y[..., 2] <= 0.01
    ox = P[..., 0] / P[..., 2] + 1; oy = P[..., 1] / P[..., 2] + 1
    fx = torch.floor(ox); fy = torch.floor(oy)
    ix = fx.long(); iy = fy.long()
    live = (~beh) & (ix >= 1) & (ix < W) & (iy >= 1) & (iy < H)
    dx = (ox != fx).long(); dy = (oy != fy).long()
    Wp = W + 2
    uid = torch.arange(H * W).view(H, W)

    def cell(a, b, k):
        c = (iy + a * dy) * Wp + ix + b * dx
        return torch.where(live, c, -(uid * 4 + k) - 2)

    NW, NE, SW, SE = cell(0, 0, 0), cell(0, 1, 1), cell(1, 0, 2), cell(1, 1, 3)
    acc("source px", H * W); acc("live px", live.sum())
    acc("emitted, rules 1 (straight)", simulate(NW, NE, SW, SE, 1))
    acc("emitted, rules 2 (straight + diagonal + same column)", simulate(NW, NE, SW, SE, 2))
    row = torch.arange(H)
    sid = ((row // SH)[:, None] * ((W + SW_ - 1) // SW_) + (torch.arange(W) // SW_)[None, :])
    acc("ideal: distinct cells per 64x8 strip", torch.unique(torch.cat([(sid[live] * BIG + c[live]) for c in (NW, NE, SW, SE)])).numel())
    acc("ideal: distinct cells per unit", torch.unique(torch.cat([c[live] for c in (NW, NE, SW, SE)])).numel())
for k, v in tot.items():
    print(f"{k:55s} {v:12.0f}  per source px {v / tot['source px']:.3f}  per live px {v / tot['live px']:.3f}")
