#!/usr/bin/env python
"""How many accumulator cells does the splat kernel emit per source pixel on config 2, by merge rule, against the
ideal (distinct cells per 64 x 8 strip / per unit)?  CPU-only model of the register-level merge of
csrc/warp_pair.cuh (in-lane NE==NW, cross-lane rotate, vertical carry); approximate at image borders.
usage: python scripts/emit_stats.py   (takes ~1 min; samples every 11th unit)"""
import sys, torch
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
frames, depths, M, src = bench.build_workload(1234)
H, W = bench.H, bench.W
n = M.shape[0]
ys, xs = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing='ij')
SH = 8
tot = {}
def acc(k, v): tot[k] = tot.get(k, 0) + float(v)
BIG = 1 << 40
for u in range(0, n, 11):
    m = M[u]; K1i = m[:9].view(3,3); T = m[9:21].view(3,4); K2 = m[21:].view(3,3)
    d = depths[src[u], 0]
    p = torch.stack([xs, ys, torch.ones_like(xs)], -1)
    X = (p @ K1i.T) * d[..., None]
    Y = X @ T[:, :3].T + T[:, 3]
    P = Y @ K2.T
    beh = Y[..., 2] <= 0.01
    ox = P[..., 0] / P[..., 2] + 1; oy = P[..., 1] / P[..., 2] + 1
    fx = torch.floor(ox); fy = torch.floor(oy)
    ix = fx.long(); iy = fy.long()
    live = (~beh) & (ix >= 1) & (ix < W) & (iy >= 1) & (iy < H)      # interior only (approx)
    dx = (ox != fx).long(); dy = (oy != fy).long()
    Wp = W + 2
    uid = torch.arange(H * W).view(H, W)
    def cell(a, b, k):   # dead -> unique negative ids (never equal to anything)
        c = (iy + a * dy) * Wp + ix + b * dx
        return torch.where(live, c, -(uid * 4 + k) - 1)
    NW, NE, SW, SE = cell(0,0,0), cell(0,1,1), cell(1,0,2), cell(1,1,3)
    lv = live
    A = slice(0, None, 2); B = slice(1, None, 2)
    acc('px', H * W); acc('live', lv.sum())
    # ---- current rules
    e = 0
    mT = NE[:, A] == NW[:, B]; mB = SE[:, A] == SW[:, B]
    t=(lv[:, A] & ~mT).sum() + (lv[:, A] & ~mB).sum(); e+=t; acc('inlane_singles', t)
    acc('inlane_diag', (lv[:, A] & ~mT & (NE[:, A] == SW[:, B])).sum() + (lv[:, A] & ~mB & (SE[:, A] == NW[:, B])).sum())
    acc('inlane_samecol', (lv[:, A] & ~mT & (NE[:, A] == NE[:, B])).sum() + (lv[:, A] & ~mB & (SE[:, A] == SE[:, B])).sum())
    acc('inlane_Bdead', (lv[:, A] & ~lv[:, B]).sum()*2)            # singles from A's right cells
    # cross-lane: B of lane l-1 -> A of lane l ; lane 0 of each 32-lane strip: never merges (different strip, receives lane 31)
    nl = W // 2
    lane = torch.arange(nl) % 32
    NEl = torch.roll(NE[:, B], 1, 1); SEl = torch.roll(SE[:, B], 1, 1); lvl = torch.roll(lv[:, B], 1, 1)
    first = (lane == 0)[None, :]
    lT = (NEl == NW[:, A]) & ~first; lB = (SEl == SW[:, A]) & ~first
    t=(lvl & ~lT).sum() + (lvl & ~lB).sum(); e+=t; acc('xlane_singles', t)
    acc('xlane_diag', (lvl & ~lT & ~first & (NEl == SW[:, A])).sum() + (lvl & ~lB & ~first & (SEl == NW[:, A])).sum())
    acc('xlane_first', (lvl & first).sum()*2)
    # vertical: carry SWA (prev row) vs NWA ; strip starts every SH rows
    row = torch.arange(H)
    has_prev = ((row % SH) != 0)[:, None]
    for S_, N_ in ((SW[:, A], NW[:, A]), (SW[:, B], NW[:, B])):
        prevS = torch.roll(S_, 1, 0); prevl = torch.roll(lv[:, A] if S_ is SW[:, A] else lv[:, B], 1, 0)
        fold = (prevS == N_) & has_prev
        # carried cell emitted when not folded (if it existed and was live-ish)
        t=(has_prev & prevl & ~fold).sum(); e+=t; acc('carry_unfolded', t)
    # last row of each strip: carries emitted
    last = ((row % SH) == SH - 1) | (row == H - 1)
    t=lv[last][:, A].sum() + lv[last][:, B].sum(); e+=t; acc('strip_end', t)
    # main emits NWA, NWB
    t=lv[:, A].sum() + lv[:, B].sum(); e+=t; acc('main', t)
    acc('emit_cur', e)
    # ---- ideal per strip: unique (strip, cell)
    sid = ((row // SH)[:, None] * ((W + 63) // 64) + (torch.arange(W) // 64)[None, :])
    allc = torch.cat([(sid[lv] * BIG + c[lv]) for c in (NW, NE, SW, SE)])
    acc('ideal_strip', torch.unique(allc).numel())
    allc = torch.cat([c[lv] for c in (NW, NE, SW, SE)])
    acc('ideal_unit', torch.unique(allc).numel())
    # row-level ideal: unique cells per (strip, source row) -- what a perfect horizontal merge could do w/o vertical
    rid = sid * H + row[:, None]
    allc = torch.cat([(rid[lv] * BIG + c[lv]) for c in (NW, NE, SW, SE)])
    acc('ideal_row', torch.unique(allc).numel())
for k, v in tot.items(): print(k, v, 'per px', v / tot['px'], 'per live', v / tot['live'])
