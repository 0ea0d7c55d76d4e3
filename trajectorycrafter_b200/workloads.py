"""Deterministic synthetic workloads for the benchmark configs of BASELINE.json.

Host-side input builders only (numpy/torch on the CPU); nothing here computes the warp.
"""
from __future__ import annotations

import math
from typing import List, Sequence, Tuple

import numpy as np
import torch

from .poses import get_poses

# reference inference_orbits.py:248-269 — the eight (theta, phi) orbit presets
ORBIT_PRESETS: List[Tuple[float, float]] = [(0, -30), (0, 30), (30, 0), (0, -45), (0, 45), (45, 0), (0, -90), (0, 90)]


def scaled_intrinsics(h: int, w: int) -> Tuple[float, float, float, float]:
    """The reference's f=500, cx=512, cy=288 at 1024x576 (demo.py:545-547) scaled per axis to h x w."""
    return 500.0 * w / 1024.0, 500.0 * h / 576.0, 512.0 * w / 1024.0, 288.0 * h / 576.0


def synthetic_clip(n_frames: int = 49, h: int = 384, w: int = 672, seed: int = 1234):
    """Config-2 generator: smooth RGB + 0.1 uniform noise in [-1,1]; a foreground blob on a tilted
    background, drifting 2 px/frame; depth = 10000/(3900*disp) as in reference models/infer.py:86-90,
    which lands in about [2.7, 17]."""
    g = torch.Generator().manual_seed(seed)
    ys = torch.linspace(0, 1, h).view(1, h, 1)
    xs = torch.linspace(0, 1, w).view(1, 1, w)
    t = torch.arange(n_frames, dtype=torch.float32).view(n_frames, 1, 1)
    drift = 2.0 * t / w
    base = torch.stack([
        0.6 * torch.sin(6.0 * (xs - drift) + 1.0) * torch.cos(4.0 * ys),
        0.6 * torch.cos(5.0 * (xs - drift)) * torch.sin(3.0 * ys + 0.5),
        0.6 * torch.sin(3.0 * (xs - drift) + 4.0 * ys),
    ], dim=1)
    frames = (base + 0.1 * (2 * torch.rand((n_frames, 3, h, w), generator=g) - 1)).clamp(-1, 1)
    disp = 0.15 + 0.6 * torch.exp(-(((xs - drift) - 0.5) ** 2 + (ys - 0.55) ** 2) / 0.03) + 0.2 * ys
    depths = (10000.0 / (3900.0 * disp)).clamp(1e-4, 1e4).unsqueeze(1)
    return frames.float().contiguous(), depths.float().contiguous()


def gradual_units(depths: torch.Tensor, target_pose: Sequence[float], h: int, w: int,
                  radius_scale: float = 1.0, scaled_k: bool = True):
    """infer_gradual pairing (demo.py:74-90): frame i x (pose_s[i]=anchor, pose_t[i], K[i])."""
    n = depths.shape[0]
    fx, fy, cx, cy = scaled_intrinsics(h, w)
    pose_s, pose_t, K = get_poses(depths, n, camera="target", target_pose=target_pose,
                                  radius_scale=radius_scale)
    if scaled_k:
        K = torch.tensor([[fx, 0.0, cx], [0.0, fy, cy], [0.0, 0.0, 1.0]]).repeat(n, 1, 1)
    return pose_s, pose_t, K


def orbit_units(depths: torch.Tensor, h: int, w: int, d_r: float = 0.0,
                presets: Sequence[Tuple[float, float]] = ORBIT_PRESETS):
    """Config 2: every preset x every frame.  Returns T1, T2, K (n_units,...) and src_index (n_units,)
    with units ordered frame-major (the 8 poses of a source frame are adjacent, so its RGB/depth is
    fetched from HBM once and re-read from L2)."""
    n = depths.shape[0]
    per = [gradual_units(depths, (th, ph, d_r, 0.0, 0.0), h, w) for th, ph in presets]
    T1 = torch.stack([p[0] for p in per], dim=1).reshape(-1, 4, 4)
    T2 = torch.stack([p[1] for p in per], dim=1).reshape(-1, 4, 4)
    K = torch.stack([p[2] for p in per], dim=1).reshape(-1, 3, 3)
    src = torch.arange(n, dtype=torch.int32).repeat_interleave(len(presets))
    return T1.contiguous(), T2.contiguous(), K.contiguous(), src


def ring_cameras(n_views: int = 16, radius: float = 6.0, elevations=(0.0, 20.0), seed: int = 72):
    """Config 4 (run_w_cam_poses_mvtracker.py-style): look-at extrinsics [R t] on a ring around a
    scene centred at z=6 in front of the source camera, per-camera intrinsics jitter."""
    rng = np.random.RandomState(seed)
    centre = np.array([0.0, 0.0, 6.0])
    exts, ks = [], []
    for i in range(n_views):
        az = math.radians(-40.0 + 80.0 * i / max(1, n_views - 1))
        el = math.radians(elevations[i % len(elevations)])
        eye = centre + radius * np.array([math.sin(az) * math.cos(el), -math.sin(el), -math.cos(az) * math.cos(el)])
        fwd = centre - eye
        fwd /= np.linalg.norm(fwd)
        right = np.cross(np.array([0.0, -1.0, 0.0]), fwd)
        right /= np.linalg.norm(right)
        down = np.cross(fwd, right)
        R = np.stack([right, down, fwd], axis=0)          # world -> camera
        E = np.eye(4)
        E[:3, :3] = R
        E[:3, 3] = -R @ eye
        exts.append(E)
        f = rng.uniform(450.0, 650.0)
        ks.append(np.array([[f, 0, 336.0 + rng.uniform(-8, 8)], [0, f, 192.0 + rng.uniform(-8, 8)], [0, 0, 1.0]]))
    return torch.tensor(np.stack(exts), dtype=torch.float32), torch.tensor(np.stack(ks), dtype=torch.float32)
