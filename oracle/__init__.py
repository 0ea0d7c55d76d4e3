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
