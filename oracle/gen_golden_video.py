#!/usr/bin/env python
"""Golden vectors from the reference's OWN fixture clip, container only.

test/videos/p7.mp4 (stride 2 as run.sh), frames 0 / 24 / 48 of a 49-frame gradual trajectory
`--target_pose 0 -30 0.3 0 0`, resized to 72x128 (INTER_AREA) to keep the fixture small; DepthCrafter cannot run
here, so depth is the deterministic proxy of SURVEY §8d (`disp = 0.1 + 0.8 * blurred luma`, then the reference's
own post-processing models/infer.py:86-90: `depth = clip(10000 / (3900 * disp), 1e-4, 1e4)`); intrinsics are the
reference's f=500, cx=512, cy=288 @1024x576 scaled per axis.  Poses come from the reference's own generators
(models/utils.py sphere2pose / generate_traj_specified through oracle/ref_loader.py), outputs from the UNMODIFIED
reference Warper on the CPU in deterministic mode.  Writes tests/golden/video_cases.npz.  TEST INFRASTRUCTURE ONLY.

    python oracle/gen_golden_video.py
"""
import os
import sys

import cv2
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402
from trajectorycrafter_b200.matrices import prepare_matrices  # noqa: E402

CLIP = "/root/reference/test/videos/p7.mp4"
H, W, N, STRIDE, PICK = 72, 128, 49, 2, (0, 24, 48)


def read_frames():
    cap = cv2.VideoCapture(CLIP)
    frames, i = [], 0
    while len(frames) < N:
        ok, bgr = cap.read()
        if not ok:
            break
        if i % STRIDE == 0:
            rgb = cv2.cvtColor(bgr, cv2.COLOR_BGR2RGB)
            frames.append(cv2.resize(rgb, (W, H), interpolation=cv2.INTER_AREA))
        i += 1
    cap.release()
    assert len(frames) == N, f"{CLIP}: only {len(frames)} frames at stride {STRIDE}"
    return np.stack(frames).astype(np.float32) / 255.0          # (N,H,W,3) in [0,1]


def main():
    assert ref_loader.available(), "the reference is only mounted in the build container"
    ref = ref_loader.load()
    torch.use_deterministic_algorithms(True)
    vid = read_frames()
    luma = (0.299 * vid[..., 0] + 0.587 * vid[..., 1] + 0.114 * vid[..., 2]).astype(np.float32)
    disp = np.stack([0.1 + 0.8 * cv2.GaussianBlur(l, (0, 0), 3.0) for l in luma]).astype(np.float32)
    depth = np.clip(10000.0 / (3900.0 * disp), 1e-4, 1e4).astype(np.float32)                # models/infer.py:86-90
    frames = torch.from_numpy(vid).permute(0, 3, 1, 2).contiguous() * 2.0 - 1.0             # demo.py:70-72
    depths = torch.from_numpy(depth).unsqueeze(1)
    # demo.py:538-586 with the reference's own pose functions
    radius = min(depths[0, 0, H // 2, W // 2] * 1.0, 5)
    c2w_init = torch.tensor([[-1.0, 0, 0, 0], [0, 1.0, 0, 0], [0, 0, -1.0, 0], [0, 0, 0, 1.0]]).unsqueeze(0)
    poses = ref.generate_traj_specified(c2w_init, 0.0, -30.0, 0.3 * radius, 0.0, 0.0, N, "cpu")
    poses[:, 2, 3] = poses[:, 2, 3] + radius
    pose_s = poses[0:1].repeat(N, 1, 1)
    K = torch.tensor([[500.0 * W / 1024.0, 0.0, 512.0 * W / 1024.0], [0.0, 500.0 * H / 576.0, 288.0 * H / 576.0],
                      [0.0, 0.0, 1.0]]).repeat(N, 1, 1)
    warper = ref.Warper(device="cpu")
    out = {}
    for i in PICK:
        o = warper.forward_warp(frames[i:i + 1], None, depths[i:i + 1], pose_s[i:i + 1], poses[i:i + 1], K[i:i + 1], None, False,
                                twice=False)                                                # demo.py:78-90
        c = {"frame": frames[i:i + 1], "depth": depths[i:i + 1], "T1": pose_s[i:i + 1], "T2": poses[i:i + 1], "K1": K[i:i + 1],
             "M": prepare_matrices(pose_s[i:i + 1], poses[i:i + 1], K[i:i + 1], None),
             "out_frame": o[0], "out_mask": o[1], "out_depth": o[2], "out_flow": o[3]}
        for k, v in c.items():
            out[f"p7_frame{i:02d}/{k}"] = v.contiguous().numpy()
        print(f"p7 frame {i:2d}: valid={float(o[1].mean()):.3f} depth range {float(depths[i].min()):.2f}..{float(depths[i].max()):.2f}")
    path = os.path.join(ROOT, "tests", "golden", "video_cases.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
