"""Camera-trajectory generation that feeds the warp (host Python, kept semantically identical).

Restates reference models/utils.py:83-210 (``sphere2pose``, ``generate_traj_specified``,
``txt_interpolation``, ``generate_traj_txt``) and demo.py:538-632 (``get_poses``/``get_poses_f``).
These run once per clip on 49 4x4 matrices; they stay on the host by design (north_star keeps
"trajectory/pose generation" in Python).  The arithmetic goes through the same torch/numpy/scipy
calls as the reference so the resulting float32 poses are bit-identical (tests/golden/poses.npz).
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import numpy as np
import torch


def _axis_rotation(axis: str, angle_deg, n: int, device) -> torch.Tensor:
    """(n,4,4) float32 rotation about the x or y axis, angle in degrees (float32 like the reference)."""
    ang = torch.deg2rad(torch.tensor(angle_deg)).to(device)
    s, c = torch.sin(ang), torch.cos(ang)
    if axis == "x":      # models/utils.py:97-105
        rows = [[1, 0, 0, 0], [0, c, -s, 0], [0, s, c, 0], [0, 0, 0, 1]]
    else:                # models/utils.py:114-122
        rows = [[c, 0, s, 0], [0, 1, 0, 0], [-s, 0, c, 0], [0, 0, 0, 1]]
    return torch.tensor(rows).unsqueeze(0).repeat(n, 1, 1).to(device)


def sphere2pose(c2ws_input: torch.Tensor, theta, phi, r, device, x=None, y=None) -> torch.Tensor:
    """models/utils.py:83-131.  Translate along world z by -r (and the reference's x/y quirk:
    ``x is not None`` moves row 1 by +y, ``y is not None`` moves row 0 by -x), then rotate by
    Rx(theta) followed by Ry(phi)."""
    c2ws = c2ws_input.clone()
    c2ws[:, 2, 3] -= r
    if x is not None:
        c2ws[:, 1, 3] += y
    if y is not None:
        c2ws[:, 0, 3] -= x
    n = c2ws.shape[0]
    c2ws = torch.matmul(_axis_rotation("x", theta, n, device), c2ws)
    c2ws = torch.matmul(_axis_rotation("y", phi, n, device), c2ws)
    return c2ws


def generate_traj_specified(c2ws_anchor, theta, phi, d_r, d_x, d_y, frame: int, device) -> torch.Tensor:
    """models/utils.py:134-158: linear ramps 0 -> target for (theta, phi, r, x, y) over `frame` poses."""
    ramps = [np.linspace(0, v, frame) for v in (theta, phi, d_r, d_x, d_y)]
    out = []
    for th, ph, r, x, y in zip(*ramps):
        out.append(sphere2pose(c2ws_anchor, np.float32(th), np.float32(ph), np.float32(r), device,
                               np.float32(x), np.float32(y)))
    return torch.cat(out, dim=0)


def txt_interpolation(input_list: Sequence[float], n: int, mode: str = "smooth") -> np.ndarray:
    """models/utils.py:161-171: cubic smoothing spline ('smooth') or linear resampling to n samples."""
    from scipy.interpolate import UnivariateSpline, interp1d
    x = np.linspace(0, 1, len(input_list))
    if mode == "smooth":
        f = UnivariateSpline(x, input_list, k=3)
    elif mode == "linear":
        f = interp1d(x, input_list)
    else:
        raise KeyError(f"Invalid txt interpolation mode: {mode}")
    return f(np.linspace(0, 1, n))


def _resample_keys(keys: Sequence[float], frame: int) -> np.ndarray:
    # models/utils.py:181-200: >3 keys -> smooth spline with the end points pinned, else linear
    if len(keys) > 3:
        vals = txt_interpolation(keys, frame, mode="smooth")
        vals[0], vals[-1] = keys[0], keys[-1]
        return vals
    return txt_interpolation(keys, frame, mode="linear")


def generate_traj_txt(c2ws_anchor, phi, theta, r, frame: int, device) -> torch.Tensor:
    """models/utils.py:174-210: key-framed (phi, theta, r) trajectory from a txt file."""
    phis, thetas, rs = _resample_keys(phi, frame), _resample_keys(theta, frame), _resample_keys(r, frame)
    out = [sphere2pose(c2ws_anchor, np.float32(th), np.float32(ph), np.float32(rr), device)
           for th, ph, rr in zip(thetas, phis, rs)]
    return torch.cat(out, dim=0)


def read_traj_txt(path: str, radius) -> Tuple[list, list, list]:
    """demo.py:564-568: line 0 = theta keys, line 1 = phi keys, line 2 = r keys (scaled by radius)."""
    with open(path, "r") as fh:
        lines = fh.readlines()
    theta = [float(i) for i in lines[0].split()]
    phi = [float(i) for i in lines[1].split()]
    r = [float(i) * radius for i in lines[2].split()]
    return theta, phi, r


C2W_INIT = [[-1.0, 0.0, 0.0, 0.0], [0.0, 1.0, 0.0, 0.0], [0.0, 0.0, -1.0, 0.0], [0.0, 0.0, 0.0, 1.0]]


def get_poses(depths: torch.Tensor, num_frames: int, *, camera: str = "target",
              target_pose: Sequence[float] = (0, 0, 0, 0, 0), traj_txt: Optional[str] = None,
              radius_scale: float = 1.0, anchor_idx: int = 0, device="cpu",
              focal: float = 500.0, cx: float = 512.0, cy: float = 288.0,
              f_new: Optional[float] = None, centre_depth=None):
    """demo.py:538-586 (and :588-632 when ``f_new`` is given: focal ramps linearly to f_new).

    ``centre_depth``: skip the read-back of the centre depth pixel and use this value (``stats[2]`` of
    ``Warper.depth_postprocess``).  Returns (pose_s (n,4,4), pose_t (n,4,4), K (n,3,3)).  radius = min(centre depth of frame 0 *
    radius_scale, 5); every pose is pushed back by +radius along z; pose_s repeats the anchor pose.
    """
    if centre_depth is None:
        centre = depths[0, 0, depths.shape[-2] // 2, depths.shape[-1] // 2].cpu()
    else:
        # the centre pixel already fetched by Warper.depth_postprocess (stats[2]): the depth map is not read back; a
        # float32 0-d tensor like the reference's own value, so the two lines below are the reference's arithmetic
        centre = torch.tensor(float(centre_depth), dtype=torch.float32)
    radius = centre * radius_scale
    radius = min(radius, 5)
    if f_new is None:
        K = torch.tensor([[focal, 0.0, cx], [0.0, focal, cy], [0.0, 0.0, 1.0]]).repeat(num_frames, 1, 1).to(device)
    else:
        f_values = torch.linspace(focal, f_new, num_frames, device=device)
        K = torch.zeros((num_frames, 3, 3), device=device)
        K[:, 0, 0] = f_values
        K[:, 1, 1] = f_values
        K[:, 0, 2] = cx
        K[:, 1, 2] = cy
        K[:, 2, 2] = 1.0
    c2w_init = torch.tensor(C2W_INIT).to(device).unsqueeze(0)
    if camera == "target":
        dtheta, dphi, dr, dx, dy = target_pose
        poses = generate_traj_specified(c2w_init, dtheta, dphi, dr * radius, dx, dy, num_frames, device)
    elif camera == "traj":
        theta, phi, r = read_traj_txt(traj_txt, radius)
        poses = generate_traj_txt(c2w_init, phi, theta, r, num_frames, device)
    else:
        raise ValueError(f"unknown camera mode {camera!r}")
    poses[:, 2, 3] = poses[:, 2, 3] + radius
    pose_s = poses[anchor_idx:anchor_idx + 1].repeat(num_frames, 1, 1)
    return pose_s, poses, K
