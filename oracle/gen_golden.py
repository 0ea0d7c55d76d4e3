"""Generates tests/golden/*.npz by running the UNMODIFIED reference (container only).

    python oracle/gen_golden.py            # writes tests/golden/warp_cases.npz, poses.npz

The reference ships no golden vectors for this path, so these fixtures — inputs plus
the outputs of /root/reference/models/utils.py::Warper on the CPU in deterministic mode — are what
pins the C oracle (tests/test_oracle_golden.py) and, through it, the CUDA path.  The prepared matrix
block M (torch CPU inverse/bmm, trajectorycrafter_b200/matrices.py) is stored with every case so
that the GPU box does not depend on its own LAPACK build reproducing the container's last ulp.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_loader  # noqa: E402
from trajectorycrafter_b200.matrices import prepare_matrices  # noqa: E402
from trajectorycrafter_b200.workloads import ring_cameras  # noqa: E402

H, W = 40, 56


def smooth_depth(h, w, lo=3.0, hi=17.0, shift=0.0):
    ys = torch.linspace(0, 1, h).view(h, 1)
    xs = torch.linspace(0, 1, w).view(1, w)
    disp = 0.15 + 0.6 * torch.exp(-(((xs - shift) - 0.5) ** 2 + (ys - 0.55) ** 2) / 0.03) + 0.2 * ys
    d = 10000.0 / (3900.0 * disp)
    return d.clamp(lo, hi).view(1, 1, h, w).float()


def rgb(h, w, g):
    ys = torch.linspace(0, 1, h).view(1, h, 1)
    xs = torch.linspace(0, 1, w).view(1, 1, w)
    base = torch.cat([0.6 * torch.sin(6 * xs + 1) * torch.cos(4 * ys), 0.6 * torch.cos(5 * xs) * torch.sin(3 * ys + .5),
                      0.6 * torch.sin(3 * xs + 4 * ys)], 0)
    return (base + 0.1 * (2 * torch.rand((3, h, w), generator=g) - 1)).clamp(-1, 1).unsqueeze(0).float()


def orbit(ref, phi, theta=0.0, r=0.0, radius=4.0):
    c2w = torch.tensor([[-1.0, 0, 0, 0], [0, 1.0, 0, 0], [0, 0, -1.0, 0], [0, 0, 0, 1.0]]).unsqueeze(0)
    src = ref.sphere2pose(c2w, np.float32(0), np.float32(0), np.float32(0), "cpu", np.float32(0), np.float32(0))
    tgt = ref.sphere2pose(c2w, np.float32(theta), np.float32(phi), np.float32(r), "cpu", np.float32(0), np.float32(0))
    src[:, 2, 3] += radius
    tgt[:, 2, 3] += radius
    return src, tgt


def K_for(h, w, f=None):
    f = f if f is not None else 500.0 * w / 1024.0
    return torch.tensor([[f, 0.0, w / 2.0], [0.0, f, h / 2.0], [0.0, 0.0, 1.0]]).unsqueeze(0)


def main():
    assert ref_loader.available(), "the reference is only mounted in the build container"
    ref = ref_loader.load()
    torch.use_deterministic_algorithms(True)      # serial index_put_: bit-reproducible oracle mode
    torch.manual_seed(0)
    g = torch.Generator().manual_seed(7)
    warper = ref.Warper(device="cpu")
    cases = {}

    def add(name, frame, mask1, depth, T1, T2, K1, K2, mask=False, twice=False):
        out = warper.forward_warp(frame, mask1, depth, T1, T2, K1, K2, mask=mask, twice=twice)
        pts = warper.compute_transformed_points(depth, T1, T2, K1, K2)
        c = {"frame": frame, "depth": depth, "T1": T1, "T2": T2, "K1": K1,
             "M": prepare_matrices(T1, T2, K1, K2), "points": pts[..., 0],
             "out_frame": out[0], "out_mask": out[1], "out_depth": out[2]}
        if mask1 is not None:
            c["mask1"] = mask1
        if K2 is not None:
            c["K2"] = K2
        if out[3] is not None:
            c["out_flow"] = out[3]
        c["flags"] = torch.tensor([int(mask), int(twice)])
        for k, v in c.items():
            cases[f"{name}/{k}"] = v.contiguous().numpy()
        print(f"{name:28s} valid={float(out[1].mean()):.3f} dtype={out[0].dtype}")

    fr = rgb(H, W, g)
    d_smooth = smooth_depth(H, W)
    K = K_for(H, W)
    # 1 smooth depth, 30 degree orbit
    s, t = orbit(ref, 30.0)
    add("smooth_phi30", fr, None, d_smooth, s, t, K, None)
    # 2 noisy depth, -45 degrees (some points go behind the camera)
    d_noisy = (d_smooth * (1 + 0.3 * (torch.rand(d_smooth.shape, generator=g) - 0.5))).float()
    s, t = orbit(ref, -45.0, radius=5.0)
    add("noisy_phi-45", fr, None, d_noisy, s, t, K, None)
    # 3 150 degrees: large behind-camera population, pixel (1,1) becomes valid
    s, t = orbit(ref, 150.0, radius=5.0)
    add("behind_phi150", fr, None, d_noisy, s, t, K, None)
    # 4 zoom: intrinsic2 = half focal
    s, t = orbit(ref, 10.0, theta=5.0)
    add("zoom_k2", fr, None, d_smooth, s, t, K, K_for(H, W, f=0.5 * 500.0 * W / 1024.0))
    # 5 random mask1 holes
    m1 = (torch.rand((1, 1, H, W), generator=g) > 0.3).float()
    s, t = orbit(ref, -20.0, theta=-10.0, r=0.5)
    add("mask1_holes", fr, m1, d_smooth, s, t, K, None)
    # 6 zero-depth holes + mask1 (core.py-style synthetic frame)
    d_holes = d_smooth * m1
    add("zero_depth_mask1", fr, m1, d_holes, s, t, K, None)
    # 7 constant far plane
    add("far_plane", fr, None, torch.full((1, 1, H, W), 1e4), s, t, K, None)
    # 8 identity warp: exact integer coordinates (floor == ceil, 4x unit weight on one cell)
    eye = torch.eye(4).unsqueeze(0)
    add("identity", fr, None, d_smooth, eye, eye.clone(), K, None)
    # 9 pure integer translation in x via intrinsics-free trick is not exact; use 90 degree edge-on fold
    s, t = orbit(ref, 90.0, radius=5.0)
    add("edge_on_phi90", fr, None, d_smooth, s, t, K, None)
    # 10 batch of two frames in one call (global soft-z max over the call)
    fr2 = torch.cat([fr, rgb(H, W, g)], 0)
    d2 = torch.cat([d_smooth, smooth_depth(H, W, shift=0.1) * 2.0], 0)
    s, t = orbit(ref, 25.0)
    add("batch2_global_max", fr2, None, d2, s.repeat(2, 1, 1), t.repeat(2, 1, 1), K.repeat(2, 1, 1), None)
    # 11 mask=True (clean_points, float64 outputs)
    s, t = orbit(ref, 30.0)
    add("clean_phi30", fr, None, d_smooth, s, t, K, None, mask=True)
    add("clean_phi-45_noisy", fr, None, d_noisy, orbit(ref, -45.0, radius=5.0)[0], orbit(ref, -45.0, radius=5.0)[1], K, None, mask=True)
    # 12 twice=True
    s, t = orbit(ref, 12.0)
    add("twice_phi12", fr, None, d_smooth, s, t, K, None, twice=True)
    # 13 general extrinsics, K2 != K1 (run_w_cam_poses_mvtracker.py style)
    ext, ks = ring_cameras(4)
    ks[:, 0, 2] = W / 2.0 + (ks[:, 0, 2] - 336.0)
    ks[:, 1, 2] = H / 2.0 + (ks[:, 1, 2] - 192.0)
    ks[:, 0, 0] *= W / 672.0
    ks[:, 1, 1] *= W / 672.0
    d_gt = (2.0 + 10.0 * torch.rand((1, 1, H, W), generator=g)).float()
    d_gt = torch.nn.functional.avg_pool2d(d_gt, 5, 1, 2)
    add("ring_cam_1", fr, None, d_gt, ext[0:1], ext[1:2], ks[0:1], ks[1:2])
    add("ring_cam_3", fr, None, d_gt, ext[0:1], ext[3:4], ks[0:1], ks[3:4])
    # 14 odd size (not a multiple of the tile)
    ho, wo = 37, 53
    add("odd_size", rgb(ho, wo, g), None, smooth_depth(ho, wo), *orbit(ref, -30.0), K_for(ho, wo), None)

    # standalone bilinear_splatting (c=2, flow field) and clean_points
    flow = 3.0 * torch.randn((1, 2, H, W), generator=g)
    sp, spm = warper.bilinear_splatting(flow, m1, d_smooth[:, 0], flow, None, is_image=False)
    cases["splat_c2/frame"] = flow.numpy()
    cases["splat_c2/mask1"] = m1.numpy()
    cases["splat_c2/depth"] = d_smooth[:, 0].numpy()
    cases["splat_c2/out"] = sp.contiguous().numpy()
    cases["splat_c2/out_mask"] = spm.contiguous().numpy()
    fmask = (torch.rand((1, 1, H, W), generator=g) > 0.2).float()
    sp, spm = warper.bilinear_splatting(fr, None, d_noisy[:, 0], flow, fmask, is_image=True)
    cases["splat_c3_img/frame"] = fr.numpy()
    cases["splat_c3_img/depth"] = d_noisy[:, 0].numpy()
    cases["splat_c3_img/flow"] = flow.numpy()
    cases["splat_c3_img/flow_mask"] = fmask.numpy()
    cases["splat_c3_img/out"] = sp.contiguous().numpy()
    cases["splat_c3_img/out_mask"] = spm.contiguous().numpy()
    m2 = (torch.rand((1, 1, H, W), generator=g) > 0.05).float()
    cf, cm = warper.clean_points(fr.clone(), m2.clone())
    cases["clean_c3/frame"] = fr.numpy()
    cases["clean_c3/mask2"] = m2.numpy()
    cases["clean_c3/out"] = cf.numpy()
    cases["clean_c3/out_mask"] = cm.numpy()
    cf, cm = warper.clean_points(d_smooth.clone(), m2.clone())
    cases["clean_c1/frame"] = d_smooth.numpy()
    cases["clean_c1/mask2"] = m2.numpy()
    cases["clean_c1/out"] = cf.numpy()
    cases["clean_c1/out_mask"] = cm.numpy()

    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    np.savez_compressed(os.path.join(out_dir, "warp_cases.npz"), **cases)

    # pose generators (models/utils.py:83-210, demo.py:538-632)
    poses = {}
    c2w = torch.tensor([[-1.0, 0, 0, 0], [0, 1.0, 0, 0], [0, 0, -1.0, 0], [0, 0, 0, 1.0]]).unsqueeze(0)
    poses["sphere2pose"] = ref.sphere2pose(c2w, np.float32(12.5), np.float32(-30.0), np.float32(0.7), "cpu",
                                           np.float32(0.1), np.float32(-0.2)).numpy()
    poses["sphere2pose_noxy"] = ref.sphere2pose(c2w, np.float32(3.0), np.float32(40.0), np.float32(-0.3), "cpu").numpy()
    poses["traj_specified"] = ref.generate_traj_specified(c2w, 10.0, -30.0, 0.3 * 4.2, 0.1, -0.05, 49, "cpu").numpy()
    for name in ("loop1", "loop2"):
        with open(f"/root/reference/test/trajs/{name}.txt") as fh:
            lines = fh.readlines()
        theta = [float(i) for i in lines[0].split()]
        phi = [float(i) for i in lines[1].split()]
        r = [float(i) * 3.7 for i in lines[2].split()]
        poses[f"traj_txt_{name}"] = ref.generate_traj_txt(c2w, phi, theta, r, 49, "cpu").numpy()
        poses[f"traj_txt_{name}_keys"] = np.array([len(theta), len(phi), len(r)])
        poses[f"traj_txt_{name}_theta"] = np.array(theta)
        poses[f"traj_txt_{name}_phi"] = np.array(phi)
        poses[f"traj_txt_{name}_r"] = np.array(r)
    np.savez_compressed(os.path.join(out_dir, "poses.npz"), **poses)
    for f in ("warp_cases.npz", "poses.npz"):
        print(f, os.path.getsize(os.path.join(out_dir, f)) // 1024, "KiB")


if __name__ == "__main__":
    main()
