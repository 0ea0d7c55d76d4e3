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


# ---------------------------------------------------------------------------------------------
# frame x pose pairings of the other entry points as src_index / pose-row builders
# ---------------------------------------------------------------------------------------------
def direct_pairing(video_length: int = 49, cut: int = 20):
    """infer_direct (reference demo.py:199-231): poses are generated for `cut` frames; unit i < cut warps frame 0
    with pose_t[i], unit i >= cut warps frame i - cut with the LAST pose.  Returns (src_index, pose_index), both
    (video_length,) int32: unit i = (frames[src_index[i]], pose_s[0], pose_t[pose_index[i]], K[0])."""
    i = np.arange(video_length)
    src = np.where(i < cut, 0, i - cut)
    pose = np.where(i < cut, i, cut - 1)
    return torch.tensor(src, dtype=torch.int32), torch.tensor(pose, dtype=torch.int32)


def bullet_pairing(video_length: int = 49):
    """infer_bullet (reference demo.py:325-342): the LAST frame under every pose of the trajectory."""
    return (torch.full((video_length,), video_length - 1, dtype=torch.int32),
            torch.arange(video_length, dtype=torch.int32))


def paired_units(depths: torch.Tensor, target_pose: Sequence[float], h: int, w: int, mode: str, cut: int = 20):
    """T1, T2, K, src_index of one `direct` / `bullet` run at h x w (scaled intrinsics), as the reference loops pair
    them; the centre-depth radius comes from frame 0 like demo.py:566-569."""
    n = depths.shape[0]
    n_pose = cut if mode == "direct" else n
    pose_s, pose_t, _ = get_poses(depths, n_pose, camera="target", target_pose=target_pose)
    src, pi = direct_pairing(n, cut) if mode == "direct" else bullet_pairing(n)
    fx, fy, cx, cy = scaled_intrinsics(h, w)
    K = torch.tensor([[fx, 0.0, cx], [0.0, fy, cy], [0.0, 0.0, 1.0]]).repeat(n, 1, 1)
    return pose_s[:1].repeat(n, 1, 1).contiguous(), pose_t[pi.long()].contiguous(), K, src


# ---------------------------------------------------------------------------------------------
# config 4 "_smooth": SLERP rotation + linear translation / intrinsics from the source to a target camera
# (reference run_w_cam_poses_mvtracker_smooth.py:176-256 interpolation, :258-361 quaternion helpers)
# ---------------------------------------------------------------------------------------------
def _rot_to_quat(R: torch.Tensor) -> torch.Tensor:
    # run_w_cam_poses_mvtracker_smooth.py:286-317, [w, x, y, z], largest-diagonal branch order kept
    tr = R[0, 0] + R[1, 1] + R[2, 2]
    if tr > 0:
        s = torch.sqrt(tr + 1.0) * 2
        q = [0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s]
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = torch.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2]) * 2
        q = [(R[2, 1] - R[1, 2]) / s, 0.25 * s, (R[0, 1] + R[1, 0]) / s, (R[0, 2] + R[2, 0]) / s]
    elif R[1, 1] > R[2, 2]:
        s = torch.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2]) * 2
        q = [(R[0, 2] - R[2, 0]) / s, (R[0, 1] + R[1, 0]) / s, 0.25 * s, (R[1, 2] + R[2, 1]) / s]
    else:
        s = torch.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1]) * 2
        q = [(R[1, 0] - R[0, 1]) / s, (R[0, 2] + R[2, 0]) / s, (R[1, 2] + R[2, 1]) / s, 0.25 * s]
    return torch.stack(q)


def _quat_to_rot(q: torch.Tensor) -> torch.Tensor:
    # :319-341 (normalises first)
    w_, x, y, z = q
    n = torch.sqrt(w_ * w_ + x * x + y * y + z * z)
    w_, x, y, z = w_ / n, x / n, y / n, z / n
    R = torch.zeros(3, 3, dtype=q.dtype)
    R[0, 0] = 1 - 2 * (y * y + z * z); R[0, 1] = 2 * (x * y - w_ * z); R[0, 2] = 2 * (x * z + w_ * y)
    R[1, 0] = 2 * (x * y + w_ * z); R[1, 1] = 1 - 2 * (x * x + z * z); R[1, 2] = 2 * (y * z - w_ * x)
    R[2, 0] = 2 * (x * z - w_ * y); R[2, 1] = 2 * (y * z + w_ * x); R[2, 2] = 1 - 2 * (x * x + y * y)
    return R


def _slerp(q1: torch.Tensor, q2: torch.Tensor, t: float) -> torch.Tensor:
    # :343-367: shortest path, linear blend when nearly parallel
    dot = torch.sum(q1 * q2)
    if dot < 0:
        q2, dot = -q2, -dot
    if dot > 0.9995:
        r = q1 + t * (q2 - q1)
        return r / torch.norm(r)
    th0 = torch.acos(torch.clamp(torch.abs(dot), 0, 1))
    s0 = torch.sin(th0)
    th = th0 * t
    return (torch.cos(th) - dot * torch.sin(th) / s0) * q1 + (torch.sin(th) / s0) * q2


def smooth_camera_path(src_c2w: torch.Tensor, tgt_c2w: torch.Tensor, src_K: torch.Tensor, tgt_K: torch.Tensor, n: int):
    """(n,4,4) poses and (n,3,3) intrinsics from the source camera to the target camera: SLERP of the rotation,
    linear translation and linear K with alpha = i / (n - 1) (run_w_cam_poses_mvtracker_smooth.py:176-256);
    alpha == 0 / 1 return the end rotations themselves (:269-272)."""
    poses, ks = [], []
    q1, q2 = _rot_to_quat(src_c2w[:3, :3]), _rot_to_quat(tgt_c2w[:3, :3])
    for i in range(n):
        a = i / (n - 1) if n > 1 else 0
        P = torch.zeros(4, 4, dtype=src_c2w.dtype)
        P[:3, :3] = src_c2w[:3, :3] if a == 0 else tgt_c2w[:3, :3] if a == 1 else _quat_to_rot(_slerp(q1, q2, a))
        P[:3, 3] = (1 - a) * src_c2w[:3, 3] + a * tgt_c2w[:3, 3]
        P[3, 3] = 1.0
        poses.append(P)
        ks.append((1 - a) * src_K + a * tgt_K)
    return torch.stack(poses), torch.stack(ks)


def multicam_units(n_frames: int = 49, n_views: int = 16, smooth: bool = False, h: int = 384, w: int = 672):
    """Config 4: one source view x `n_views` ring cameras x `n_frames` frames = n_views * n_frames units, frame-major.
    Extrinsics are used "as is" for both transformation1 / transformation2 like run_w_cam_poses_mvtracker.py:245-258;
    `smooth` walks from the source camera to each target over the clip (the _smooth script).
    Returns T1, T2, K1, K2 (n_units, ...) and src_index."""
    ext, ks = ring_cameras(n_views)
    fx, fy, cx, cy = scaled_intrinsics(h, w)
    src_E = torch.eye(4)
    src_K = torch.tensor([[fx, 0.0, cx], [0.0, fy, cy], [0.0, 0.0, 1.0]])
    T1 = src_E.repeat(n_frames * n_views, 1, 1)
    K1 = src_K.repeat(n_frames * n_views, 1, 1)
    T2 = torch.empty((n_frames, n_views, 4, 4))
    K2 = torch.empty((n_frames, n_views, 3, 3))
    for v in range(n_views):
        if smooth:
            p, k = smooth_camera_path(src_E, ext[v], src_K, ks[v], n_frames)
            T2[:, v], K2[:, v] = p, k
        else:
            T2[:, v], K2[:, v] = ext[v], ks[v]
    src = torch.arange(n_frames, dtype=torch.int32).repeat_interleave(n_views)
    return T1.contiguous(), T2.reshape(-1, 4, 4).contiguous(), K1.contiguous(), K2.reshape(-1, 3, 3).contiguous(), src


# ---------------------------------------------------------------------------------------------
# config 5: inference_autoregressive.py long video -- n_splits chained segments
# (reference notebooks/28_08_25_trajectories/core_autoregressive.py:26-88 deltas, :173-235 pose continuation)
# ---------------------------------------------------------------------------------------------
def autoregressive_segments(depths_per_segment, target_pose: Sequence[float], n_splits: int = 8, radius_scale: float = 1.0,
                            mode: str = "gradual"):
    """Poses of the `n_splits` chained segments.  `depths_per_segment[k]` is segment k's depth clip (the reference
    re-estimates depth from the previous segment's generated video; its centre pixel sets the radius).
    Quirks kept from the reference: the target is split into equal deltas (:36-47); every segment's trajectory runs from the
    cumulative pose to cumulative + delta on the anchor camera (generate_traj_specified :545-575); segment 0 uses its
    first pose as the source pose, segment 1 uses segment 0's pose_t as per-frame source poses (:206-216), and segments
    >= 2 REUSE segment 1's (pose_s, pose_t) unchanged (:218-223).  Returns a list of (pose_s, pose_t, K), K = f 500 /
    cx 512 / cy 288 as in :237-240 (the caller may substitute scaled intrinsics)."""
    from .poses import C2W_INIT, sphere2pose
    target = list(target_pose)
    cum = [target[i] * 0.0 for i in range(5)]
    fracs = [[target[j] * (i + 1) / n_splits for i in range(n_splits)] for j in range(5)]
    deltas = [[fr[0]] + [fr[i] - fr[i - 1] for i in range(1, n_splits)] for fr in fracs]
    out, prev_t, keep = [], None, None
    c2w_init = torch.tensor(C2W_INIT).unsqueeze(0)
    for k in range(n_splits):
        depths = depths_per_segment[k]
        n = depths.shape[0]
        radius = depths[0, 0, depths.shape[-2] // 2, depths.shape[-1] // 2].cpu() * radius_scale
        radius = min(radius, 5)
        seg = [deltas[j][k] for j in range(5)]
        ramps = [np.linspace(cum[j], cum[j] + seg[j], n) for j in range(5)]
        pose_t = torch.cat([sphere2pose(c2w_init, np.float32(th), np.float32(ph), np.float32(r), "cpu", np.float32(x), np.float32(y))
                            for th, ph, r, x, y in zip(*ramps)], dim=0)
        pose_t[:, 2, 3] = pose_t[:, 2, 3] + radius
        if k == 0 or prev_t is None:
            pose_s = pose_t[0:1].repeat(n, 1, 1)
        else:
            pose_s = prev_t
        if k == 1:
            keep = (pose_s, pose_t)
        elif k > 1:
            pose_s, pose_t = keep
        for j in range(5):
            cum[j] += seg[j]
        ret_t = pose_t
        if mode == "direct":          # :115-117: only the last target pose
            pose_t = pose_t[-1:].repeat(n, 1, 1)
        K = torch.tensor([[500.0, 0.0, 512.0], [0.0, 500.0, 288.0], [0.0, 0.0, 1.0]]).repeat(n, 1, 1)
        out.append((pose_s, pose_t, K))
        prev_t = ret_t
    return out
