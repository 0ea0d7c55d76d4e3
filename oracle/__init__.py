"""CPU oracle for the forward re-rendering path — TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The product package
(``trajectorycrafter_b200``) never does; it fails loudly when its CUDA library is missing.

``liborc.so`` is the C restatement in ``warp_oracle.c`` (each function cites the reference
file:line it follows).  It is pinned against outputs of the reference itself
(``tests/golden/*.npz`` produced by ``oracle/gen_golden.py``), because the reference ships no
golden vectors of its own.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liborc.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile liborc.so with gcc (seconds)."""
    src = os.path.join(_HERE, "warp_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "liborc.so"])
    return _LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
    return _lib


def _fp(a: Optional[np.ndarray]):
    if a is None:
        return None
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _f32(a) -> Optional[np.ndarray]:
    if a is None:
        return None
    if hasattr(a, "detach"):
        a = a.detach().cpu().numpy()
    return np.ascontiguousarray(a, dtype=np.float32)


def project(depth, M, src_index=None):
    """compute_transformed_points (models/utils.py:350-420).  depth (n_src,1,h,w) -> (n,h,w,3)."""
    depth = _f32(depth)
    M = _f32(M).reshape(-1, 30)
    n_src, _, h, w = depth.shape
    n = M.shape[0]
    si = None if src_index is None else np.ascontiguousarray(src_index, dtype=np.int32)
    out = np.empty((n, h, w, 3), np.float32)
    rc = lib().orc_project(_fp(depth), None if si is None else si.ctypes.data_as(ctypes.POINTER(ctypes.c_int)),
                           _fp(M), n, n_src, h, w, _fp(out))
    assert rc == 0, rc
    return out


def forward_warp(frame, mask1, depth, M, src_index=None, per_unit_max=True):
    """forward_warp(twice=False) (models/utils.py:220-293) on prepared matrices M (n,30).

    Returns (warped_frame (n,3,h,w), mask (n,1,h,w), warped_depth (n,1,h,w), flow12 (n,2,h,w), range_flag).
    """
    frame, depth, mask1 = _f32(frame), _f32(depth), _f32(mask1)
    M = _f32(M).reshape(-1, 30)
    n_src, _, h, w = frame.shape
    n = M.shape[0]
    si = None if src_index is None else np.ascontiguousarray(src_index, dtype=np.int32)
    of = np.empty((n, 3, h, w), np.float32)
    om = np.empty((n, 1, h, w), np.float32)
    od = np.empty((n, 1, h, w), np.float32)
    ofl = np.empty((n, 2, h, w), np.float32)
    flag = ctypes.c_int(0)
    rc = lib().orc_forward_warp(_fp(frame), _fp(mask1), _fp(depth),
                                None if si is None else si.ctypes.data_as(ctypes.POINTER(ctypes.c_int)),
                                _fp(M), n, n_src, h, w, int(bool(per_unit_max)),
                                _fp(of), _fp(om), _fp(od), _fp(ofl), ctypes.byref(flag))
    assert rc == 0, rc
    return of, om, od, ofl, flag.value


def splat(frame, mask1, depth, flow, flow_mask=None, is_image=False):
    """bilinear_splatting (models/utils.py:422-583). depth is (b,h,w) as the reference passes it."""
    frame, depth, flow = _f32(frame), _f32(depth), _f32(flow)
    mask1, flow_mask = _f32(mask1), _f32(flow_mask)
    b, c, h, w = frame.shape
    out = np.empty((b, c, h, w), np.float32)
    om = np.empty((b, 1, h, w), np.float32)
    flag = ctypes.c_int(0)
    rc = lib().orc_splat(_fp(frame), _fp(mask1), _fp(depth), _fp(flow), _fp(flow_mask),
                         b, c, h, w, int(bool(is_image)), _fp(out), _fp(om), ctypes.byref(flag))
    assert rc == 0, rc
    return out, om, flag.value


def clean_points(frame, mask2):
    """clean_points (models/utils.py:585-626) for (1,c,h,w); float64 outputs like the reference."""
    frame, mask2 = _f32(frame), _f32(mask2)
    _, c, h, w = frame.shape
    of = np.empty((1, c, h, w), np.float64)
    om = np.empty((1, 1, h, w), np.float64)
    rc = lib().orc_clean_points(_fp(frame), _fp(mask2), c, h, w,
                                of.ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
                                om.ctypes.data_as(ctypes.POINTER(ctypes.c_double)))
    assert rc == 0, rc
    return of, om


def contrib_count(flow):
    """Contributions per output pixel (b,1,h,w) int32 for a flow field (b,2,h,w); see orc_contrib_count."""
    flow = _f32(flow)
    b, _, h, w = flow.shape
    out = np.zeros((b, 1, h, w), np.int32)
    rc = lib().orc_contrib_count(_fp(flow), b, h, w, out.ctypes.data_as(ctypes.POINTER(ctypes.c_int)))
    assert rc == 0, rc
    return out


def resize_pack(frame, mask, size):
    """(cat(warped)+1)/2 -> bilinear resize -> (1,3,F,h,w);  (1 - nearest(mask)) * 255 -> (1,1,F,h,w)  (demo.py:91-100,117-120)."""
    frame, mask = _f32(frame), _f32(mask)
    n, _, H, W = frame.shape
    h, w = int(size[0]), int(size[1])
    video = np.empty((1, 3, n, h, w), np.float32)
    masks = np.empty((1, 1, n, h, w), np.float32)
    rc = lib().orc_resize_pack(_fp(frame), _fp(mask), n, H, W, h, w, _fp(video), _fp(masks))
    assert rc == 0, rc
    return video, masks


def depth_post(res, near=0.0001, far=10000.0, radius_scale=1.0):
    """models/infer.py:79-92 restated in numpy float32, one correctly rounded operation at a time (no torch): channel
    mean ((a + b) + c) / 3, min / max normalisation over the whole clip, * 3900, floor at 1e-5, 10000 / x, clip; and the
    radius of demo.py:539-543.  res (n,h,w,c) or (n,h,w).  Returns (depths (n,1,h,w) float32, stats float32[4] = min, max,
    centre depth, min(centre * radius_scale, 5))."""
    res = np.asarray(res, dtype=np.float32)
    if res.ndim == 4:
        s = res[..., 0].copy()
        for k in range(1, res.shape[-1]):
            s = (s + res[..., k]).astype(np.float32)
        if res.shape[-1] > 1:
            s = (s / np.float32(res.shape[-1])).astype(np.float32)
        res = s
    lo, hi = np.float32(res.min()), np.float32(res.max())
    d = ((res - lo).astype(np.float32) / np.float32(hi - lo)).astype(np.float32)
    d = (d * np.float32(3900.0)).astype(np.float32)
    d = np.where(d < np.float32(1e-5), np.float32(1e-5), d).astype(np.float32)
    # torch evaluates `10000.0 / tensor` as tensor.reciprocal() * 10000.0: two roundings (probed on torch 2.11 CPU)
    d = ((np.float32(1.0) / d).astype(np.float32) * np.float32(10000.0)).astype(np.float32)
    d = np.clip(d, np.float32(near), np.float32(far)).astype(np.float32)[:, None]
    centre = d[0, 0, d.shape[-2] // 2, d.shape[-1] // 2]
    radius = np.float32(min(np.float32(centre * np.float32(radius_scale)), np.float32(5.0)))
    return d, np.array([lo, hi, centre, radius], dtype=np.float32)


# ---------------------------------------------------------------------------------------------
# the fork's global point-cloud renderer, restated in numpy
# (reference notebooks/06_10_25_vggt/warper_point_cloud.py; pinned by tests/golden/pointcloud_cases.npz)
# ---------------------------------------------------------------------------------------------
def _fma32(a, b, c):
    """float32 fused multiply-add through float64 (the product of two float32 is exact in float64)."""
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(np.float32)


def _gemm_rows(Mrows, cols):
    """Rows of `M @ X` the way torch's CPU GEMM evaluates them here: per element a chain of fmas from 0 in index order."""
    out = []
    for row in Mrows:
        acc = np.zeros(cols[0].shape, np.float32)
        for k, c in enumerate(cols):
            acc = _fma32(np.float32(row[k]), c, acc)
        out.append(acc)
    return out


def pc_project(points, T, K):
    """:534-543 / :199-226: camera = T @ [p, 1], projected = K @ camera, pixel = projected.xy / projected.z."""
    p = np.asarray(points, np.float32)
    one = np.ones(p.shape[0], np.float32)
    cam = _gemm_rows(np.asarray(T, np.float32)[:3], [p[:, 0], p[:, 1], p[:, 2], one])
    proj = _gemm_rows(np.asarray(K, np.float32), cam)
    with np.errstate(divide="ignore", invalid="ignore"):
        px = (proj[0] / proj[2]).astype(np.float32)
        py = (proj[1] / proj[2]).astype(np.float32)
    return cam, proj, px, py


def pc_lift(frame, mask1, depth, T1inv, K1inv, step=1):
    """lift_to_3d_pointcloud (:12-77): broadcast matmuls = sequential un-fused float32 like models/utils.py:388-399.
    frame (b,3,h,w), depth (b,1,h,w); T1inv (b,4,4), K1inv (b,3,3) already inverted (torch.linalg.inv on the host)."""
    frame, depth = _f32(frame), _f32(depth)
    b, _, h, w = frame.shape
    ys, xs = np.meshgrid(np.arange(h, dtype=np.float32) * np.float32(step), np.arange(w, dtype=np.float32) * np.float32(step), indexing="ij")
    m = np.ones((b, h, w), np.float32) if mask1 is None else _f32(mask1)[:, 0]
    pts = np.empty((b, h, w, 3), np.float32)

    def dot(row, cols):
        acc = np.zeros(cols[0].shape, np.float32)
        for k, c in enumerate(cols):
            acc = (acc + (np.float32(row[k]) * c).astype(np.float32)).astype(np.float32)
        return acc
    one = np.ones((h, w), np.float32)
    for i in range(b):
        Ki, Ti = np.asarray(K1inv[i], np.float32), np.asarray(T1inv[i], np.float32)
        u = [dot(Ki[r], [xs, ys, one]) for r in range(3)]
        X = [(depth[i, 0] * u[r]).astype(np.float32) for r in range(3)]
        for r in range(3):
            pts[i, ..., r] = (dot(Ti[r], [X[0], X[1], X[2], one]) * m[i]).astype(np.float32)
    cols = (np.transpose(frame, (0, 2, 3, 1)) * m[..., None]).astype(np.float32)
    return pts, cols


def pc_zbuffer(points, colors, T, K, h, w, point_size=1, clamp=False, first_wins=False):
    """render_pointcloud_zbuffer_vectorized_point_size (:522-646) / render_pointcloud_zbuffer_vectorized (:445-519, clamp=True).
    Nearest entry per pixel; among equal depths the LAST entry in (point, offset) order (the sequential CPU index_put).
    first_wins=True: render_pointcloud_zbuffer_vectorized_fixed (:650-765), whose loop `if depth < buffer` keeps the FIRST of
    equal depths (and which clamps for point_size <= 1: pass clamp=True).
    Returns frame (1,3,h,w) with -1 in empty pixels, mask (1,1,h,w), depth (1,1,h,w)."""
    colors = np.asarray(colors, np.float32)
    cam, _, px, py = pc_project(points, T, K)
    z = cam[2]
    valid = (z > np.float32(0.01)) & (px >= 0) & (px < w) & (py >= 0) & (py < h)
    idx = np.nonzero(valid)[0]
    r = 0 if clamp else point_size // 2
    offs = [(dx, dy) for dy in range(-r, r + 1) for dx in range(-r, r + 1)]
    frame = np.full((3, h * w), -1.0, np.float32)
    dbuf = np.zeros(h * w, np.float32)
    if idx.size:
        ex = np.stack([np.rint((px[idx] + np.float32(dx)).astype(np.float32)) for dx, _ in offs], 1).astype(np.int64)
        ey = np.stack([np.rint((py[idx] + np.float32(dy)).astype(np.float32)) for _, dy in offs], 1).astype(np.int64)
        if clamp:
            ex, ey = np.clip(ex, 0, w - 1), np.clip(ey, 0, h - 1)
        ez = np.repeat(z[idx][:, None], len(offs), 1)
        ep = np.repeat(idx[:, None], len(offs), 1)
        inb = (ex >= 0) & (ex < w) & (ey >= 0) & (ey < h)
        lin, ez, ep = (ey * w + ex)[inb], ez[inb], ep[inb]
        # sort by (pixel, depth, entry order): the first of each pixel's minimum-depth run ends at its last entry
        order = np.lexsort((np.arange(lin.size), ez, lin))
        lin, ez, ep = lin[order], ez[order], ep[order]
        first = np.ones(lin.size, bool)
        first[1:] = lin[1:] != lin[:-1]
        start = np.nonzero(first)[0]
        end = np.append(start[1:], lin.size)
        for s0, e0 in zip(start, end):
            k = s0
            while not first_wins and k + 1 < e0 and ez[k + 1] == ez[s0]:
                k += 1
            frame[:, lin[s0]] = colors[ep[k]]
            dbuf[lin[s0]] = ez[s0]
    return frame.reshape(1, 3, h, w), (dbuf > 0).astype(np.float32).reshape(1, 1, h, w), dbuf.reshape(1, 1, h, w)


def pc_bilinear(points, colors, weights, T, K, h, w, depth_threshold=100000.0):
    """render_from_camera(use_bilinear=True) for one camera (:199-250) + _splat_points_bilinear_efficient (:351-443).
    Sequential float32 scatter in the reference's order (NW, NE, SW, SE blocks of all points).  Returns (3,h,w), (1,h,w)."""
    colors, weights = np.asarray(colors, np.float32), np.asarray(weights, np.float32).reshape(-1)
    cam, proj, px, py = pc_project(points, T, K)
    vis = (cam[2] > np.float32(0.1)) & (cam[2] < np.float32(depth_threshold)) & (px >= 0) & (px < w) & (py >= 0) & (py < h)
    out = np.zeros((3, h * w), np.float32)
    wb = np.zeros(h * w, np.float32)
    if vis.any():
        x, y, d, c, wt = px[vis], py[vis], proj[2][vis], colors[vis], weights[vis]
        nd = ((d - d.min()).astype(np.float32) / np.float32(np.float32(d.max() - d.min()) + np.float32(1e-8))).astype(np.float32)
        pw = (wt * np.exp((-nd * np.float32(2)).astype(np.float32)).astype(np.float32)).astype(np.float32)
        fx0, fy0 = np.floor(x), np.floor(y)
        dx, dy = (x - fx0).astype(np.float32), (y - fy0).astype(np.float32)
        ax, ay = (np.float32(1) - dx).astype(np.float32), (np.float32(1) - dy).astype(np.float32)
        x0, y0 = fx0.astype(np.int64), fy0.astype(np.int64)
        ws_ = [((ax * ay).astype(np.float32) * pw).astype(np.float32), ((dx * ay).astype(np.float32) * pw).astype(np.float32),
               ((ax * dy).astype(np.float32) * pw).astype(np.float32), ((dx * dy).astype(np.float32) * pw).astype(np.float32)]
        for k, wk in enumerate(ws_):
            xx, yy = x0 + (k & 1), y0 + (k >> 1)
            ok = (xx >= 0) & (xx < w) & (yy >= 0) & (yy < h) & (wk > np.float32(1e-8))
            lin = (yy * w + xx)[ok]
            for ch in range(3):
                np.add.at(out[ch], lin, (c[ok, ch] * wk[ok]).astype(np.float32))
            np.add.at(wb, lin, wk[ok])
        v = wb > np.float32(1e-8)
        out[:, v] = (out[:, v] / wb[v]).astype(np.float32)
        return out.reshape(3, h, w), v.astype(np.float32).reshape(1, h, w)
    return out.reshape(3, h, w), wb.reshape(1, h, w)


def pc_disc_table(point_size):
    """The stamp of _splat_points_to_image_with_size (:281-296): offsets (dx, dy) and float32 Gaussian falloff weights."""
    radius = int(point_size / 2)
    offs, fall = [], []
    for dy in range(-radius, radius + 1):
        for dx in range(-radius, radius + 1):
            dist = (dx ** 2 + dy ** 2) ** 0.5
            if dist <= point_size / 2:
                offs.append((dx, dy))
                fall.append(np.exp(np.float32(-dist ** 2 / (2 * (point_size / 4) ** 2))))
    return offs, np.asarray(fall, np.float32)


def pc_disc(points, colors, weights, T, K, h, w, point_size=2.0, depth_threshold=100000.0):
    """render_from_camera(use_bilinear=False) for one camera (:199-256) + _splat_points_to_image_with_size (:258-349):
    sequential float32 scatter in the reference's order (point-major, offsets inside).  Returns (3,h,w), (1,h,w)."""
    colors, weights = np.asarray(colors, np.float32), np.asarray(weights, np.float32).reshape(-1)
    cam, proj, px, py = pc_project(points, T, K)
    vis = (cam[2] > np.float32(0.1)) & (cam[2] < np.float32(depth_threshold)) & (px >= 0) & (px < w) & (py >= 0) & (py < h)
    out = np.zeros((3, h * w), np.float32)
    wb = np.zeros(h * w, np.float32)
    if not vis.any():
        return out.reshape(3, h, w), wb.reshape(1, h, w)
    x, y, d, c, wt = px[vis], py[vis], proj[2][vis], colors[vis], weights[vis]
    nd = (d / np.float32(d.max() + np.float32(1e-8))).astype(np.float32)
    tw = (wt * np.exp((-nd * np.float32(2)).astype(np.float32)).astype(np.float32)).astype(np.float32)
    offs, fall = pc_disc_table(point_size)
    ex = np.stack([np.rint((x + np.float32(dx)).astype(np.float32)) for dx, _ in offs], 1).astype(np.int64)
    ey = np.stack([np.rint((y + np.float32(dy)).astype(np.float32)) for _, dy in offs], 1).astype(np.int64)
    fw = (tw[:, None] * fall[None, :]).astype(np.float32)
    inb = (ex >= 0) & (ex < w) & (ey >= 0) & (ey < h)
    lin = (ey * w + ex)[inb]
    fwv = fw[inb]
    cc = np.repeat(c[:, None, :], len(offs), 1)[inb]
    for ch in range(3):
        np.add.at(out[ch], lin, (cc[:, ch] * fwv).astype(np.float32))
    np.add.at(wb, lin, fwv)
    v = wb > np.float32(1e-6)
    out[:, v] = (out[:, v] / wb[v]).astype(np.float32)
    return out.reshape(3, h, w), v.astype(np.float32).reshape(1, h, w)
