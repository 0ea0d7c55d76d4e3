#!/usr/bin/env python
"""Golden vectors for SURVEY.md 8f N2 from the UNMODIFIED reference, container only.

Imports GlobalPointCloudWarper from /root/reference/notebooks/06_10_25_vggt/warper_point_cloud.py (device='cpu') and
executes VisualizationWarper.extract_3d_points_with_colors from the text of
/root/reference/notebooks/28_08_25_trajectories/core.py:15-79 (the module itself imports the whole diffusion stack, so only
that method is compiled, onto a subclass of the reference's own Warper).  Writes tests/golden/pointcloud_cases.npz.
TEST INFRASTRUCTURE ONLY.

    python oracle/gen_golden_pointcloud.py
"""
import importlib.util
import math
import os
import sys
import textwrap

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402

PC = "/root/reference/notebooks/06_10_25_vggt/warper_point_cloud.py"
CORE = "/root/reference/notebooks/28_08_25_trajectories/core.py"


def load_pc():
    spec = importlib.util.spec_from_file_location("_tc_reference_pc", PC)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def rot_y(a):
    c, s = math.cos(a), math.sin(a)
    return torch.tensor([[c, 0, s, 0], [0, 1, 0, 0], [-s, 0, c, 0], [0, 0, 0, 1]], dtype=torch.float32)


def main():
    torch.use_deterministic_algorithms(True)
    ref_utils = ref_loader.load()
    pc = load_pc().GlobalPointCloudWarper(device="cpu")
    # the one method of VisualizationWarper, compiled from the reference's own text onto the reference's own Warper
    lines = open(CORE).read().splitlines()[14:79]
    ns = {"torch": torch}
    exec(textwrap.dedent("\n".join(lines)), ns)
    Vis = type("VisualizationWarper", (ref_utils.Warper,), {"extract_3d_points_with_colors": ns["extract_3d_points_with_colors"]})
    vis = Vis(device="cpu")

    g = torch.Generator().manual_seed(21)
    h, w = 60, 96
    ys = torch.linspace(0, 1, h).view(1, 1, h, 1)
    xs = torch.linspace(0, 1, w).view(1, 1, 1, w)
    frame = (torch.rand((2, 3, h, w), generator=g) * 2 - 1).float()
    depth = (3.0 + 4.0 * torch.exp(-((xs - 0.5) ** 2 + (ys - 0.5) ** 2) / 0.05) + 0.5 * torch.rand((2, 1, h, w), generator=g)).float()
    depth[1, 0, :5, :7] = 0.0
    mask = (torch.rand((2, 1, h, w), generator=g) > 0.2).float()
    K = torch.tensor([[80.0, 0, 48.0], [0, 82.0, 30.0], [0, 0, 1]]).repeat(2, 1, 1)
    T1 = torch.stack([torch.eye(4), rot_y(0.1)])
    T1[1, :3, 3] = torch.tensor([0.2, -0.1, 0.3])
    out = {"frame": frame.numpy(), "depth": depth.numpy(), "mask": mask.numpy(), "K": K.numpy(), "T1": T1.numpy()}
    wp, wc = pc.lift_to_3d_pointcloud(frame, mask, depth, T1, K)
    out["lift/points"], out["lift/colors"] = wp.numpy(), wc.numpy()
    ep, ec = vis.extract_3d_points_with_colors(frame, depth, T1, K, subsample_step=4)
    out["extract/points"], out["extract/colors"] = ep.numpy(), ec.numpy()

    # a merged cloud of both images, rendered from two target cameras
    p0, c0, w0 = pc.create_pointcloud_from_image(frame[:1], mask[:1], depth[:1], T1[:1], K[:1], 1.0)
    p1, c1, w1 = pc.create_pointcloud_from_image(frame[1:], mask[1:], depth[1:], T1[1:], K[1:], 0.5)
    P, C, Wt = pc.merge_pointclouds([p0, p1], [c0, c1], [w0, w1])
    # exact ties for the z-buffer: a few duplicated points with other colours at the END of the cloud
    pick = torch.arange(500, 9000, 212)             # spread over both images: most are visible from both target cameras
    dup = P[pick].clone()
    P = torch.cat([P, dup]); C = torch.cat([C, -C[pick]]); Wt = torch.cat([Wt, Wt[pick]])
    Tt = torch.stack([rot_y(-0.15), rot_y(0.25)])
    Tt[0, :3, 3] = torch.tensor([-0.3, 0.05, 0.4]); Tt[1, :3, 3] = torch.tensor([0.5, 0.0, -0.2])
    Kt = torch.tensor([[90.0, 0, 50.0], [0, 90.0, 28.0], [0, 0, 1]]).repeat(2, 1, 1)
    out.update({"cloud/points": P.numpy(), "cloud/colors": C.numpy(), "cloud/weights": Wt.numpy(), "Tt": Tt.numpy(), "Kt": Kt.numpy()})
    for cam in range(2):
        for ps in (1, 2, 3):
            f, m, d = pc.render_pointcloud_zbuffer_vectorized_point_size(P, C, Tt[cam:cam + 1], Kt[cam:cam + 1], (h, w), ps, return_depth=True)
            out[f"zb/cam{cam}_ps{ps}/frame"], out[f"zb/cam{cam}_ps{ps}/mask"], out[f"zb/cam{cam}_ps{ps}/depth"] = f.numpy(), m.numpy(), d.numpy()
        f, m = pc.render_pointcloud_zbuffer_vectorized(P, C, Tt[cam:cam + 1], Kt[cam:cam + 1], (h, w))
        out[f"zbc/cam{cam}/frame"], out[f"zbc/cam{cam}/mask"] = f.numpy(), m.numpy()
    f, m = pc.render_from_camera(P, C, Wt, Tt, Kt, h, w)
    out["bil/frame"], out["bil/mask"] = f.numpy(), m.numpy()
    # the disc splat of render_from_camera (use_bilinear=False) and the sequential "fixed" z-buffer (a Python loop over every
    # footprint entry in the reference: seconds at this size)
    for ps in (2.0, 3.0, 5.0):
        f, m = pc.render_from_camera(P, C, Wt, Tt, Kt, h, w, point_size=ps, use_bilinear=False)
        out[f"disc/ps{ps:g}/frame"], out[f"disc/ps{ps:g}/mask"] = f.numpy(), m.numpy()
    for cam in range(2):
        for ps in (1, 2, 3):
            f, m = pc.render_pointcloud_zbuffer_vectorized_fixed(P, C, Tt[cam:cam + 1], Kt[cam:cam + 1], (h, w), ps)
            out[f"zbf/cam{cam}_ps{ps}/frame"], out[f"zbf/cam{cam}_ps{ps}/mask"] = f.numpy(), m.numpy()
    print("cloud", tuple(P.shape), "zbuffer coverage", float(out["zb/cam0_ps1/mask"].mean()), float(out["zb/cam0_ps3/mask"].mean()),
          "bilinear coverage", float(m.mean()))
    path = os.path.join(ROOT, "tests", "golden", "pointcloud_cases.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
