#!/usr/bin/env python
"""BASELINE config 1 at FULL size, container only: frames 0 / 24 / 48 of the reference's own fixture clip
test/videos/p7.mp4 (stride 2 as run.sh) at 384x672, the run.sh preset `--target_pose 0 -30 0.3 0 0 --mode gradual`, outputs of
the UNMODIFIED reference Warper on the CPU (deterministic mode), with mask=False and mask=True as inference_orbits.py:281 sets it.

Kept small: the frames are stored as uint8 (frame = x / 255 * 2 - 1 is exact to reproduce), the proxy depth (DepthCrafter cannot
run here; SURVEY §8d's `disp = 0.1 + 0.8 * blurred luma`, then models/infer.py:86-90) is rounded to float16 BEFORE it is fed to
the reference, so the float16 copy in the fixture is the exact input; of the outputs the fixture holds SHA-256 digests of the
validity mask and of flow12 (both must be reproduced bit for bit) and 8x8 block means of colour / depth in float64 (the
full-resolution colour is checked against the oracle, which the same digests pin).  Writes tests/golden/config1_p7.npz.
TEST INFRASTRUCTURE ONLY.

    python oracle/gen_golden_config1.py
"""
import hashlib
import os
import sys

import cv2
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402

CLIP = "/root/reference/test/videos/p7.mp4"
H, W, N, STRIDE, PICK = 384, 672, 49, 2, (0, 24, 48)


def digest(a: np.ndarray) -> np.ndarray:
    """SHA-256 of the array's bytes with -0.0 folded into +0.0 (torch.where / clamp do not keep the sign of zero)."""
    a = np.ascontiguousarray(a) + np.zeros((), dtype=a.dtype)
    return np.frombuffer(hashlib.sha256(a.tobytes()).digest(), dtype=np.uint8).copy()


def block_mean(a: np.ndarray, k: int = 8) -> np.ndarray:
    n, c, h, w = a.shape
    return a.astype(np.float64).reshape(n, c, h // k, k, w // k, k).mean(axis=(3, 5))


def read_frames():
    cap = cv2.VideoCapture(CLIP)
    frames, i = [], 0
    while len(frames) < N:
        ok, bgr = cap.read()
        if not ok:
            break
        if i % STRIDE == 0:
            frames.append(cv2.resize(cv2.cvtColor(bgr, cv2.COLOR_BGR2RGB), (W, H), interpolation=cv2.INTER_AREA))
        i += 1
    cap.release()
    assert len(frames) == N
    return np.stack(frames)                                      # (N,H,W,3) uint8


def inputs_from_fixture(frames_u8: np.ndarray, depth_f16: np.ndarray):
    """(frames, depths) float32 tensors exactly as the generator fed them to the reference (shared with the tests)."""
    frames = torch.from_numpy(frames_u8.astype(np.float32) / 255.0).permute(0, 3, 1, 2).contiguous() * 2.0 - 1.0      # demo.py:70-72
    depths = torch.from_numpy(depth_f16.astype(np.float32)).unsqueeze(1).contiguous()
    return frames, depths


def main():
    assert ref_loader.available(), "the reference is only mounted in the build container"
    ref = ref_loader.load()
    torch.use_deterministic_algorithms(True)
    vid = read_frames()
    v = vid.astype(np.float32) / 255.0
    luma = (0.299 * v[..., 0] + 0.587 * v[..., 1] + 0.114 * v[..., 2]).astype(np.float32)
    disp = np.stack([0.1 + 0.8 * cv2.GaussianBlur(l, (0, 0), 5.0) for l in luma]).astype(np.float32)
    depth16 = np.clip(10000.0 / (3900.0 * disp), 1e-4, 1e4).astype(np.float16)                # models/infer.py:86-90, then float16
    frames, depths = inputs_from_fixture(vid, depth16)
    radius = min(depths[0, 0, H // 2, W // 2] * 1.0, 5)                                       # demo.py:566-569
    c2w_init = torch.tensor([[-1.0, 0, 0, 0], [0, 1.0, 0, 0], [0, 0, -1.0, 0], [0, 0, 0, 1.0]]).unsqueeze(0)
    poses = ref.generate_traj_specified(c2w_init, 0.0, -30.0, 0.3 * radius, 0.0, 0.0, N, "cpu")
    poses[:, 2, 3] = poses[:, 2, 3] + radius
    pose_s = poses[0:1].repeat(N, 1, 1)
    K = torch.tensor([[500.0 * W / 1024.0, 0.0, 512.0 * W / 1024.0], [0.0, 500.0 * H / 576.0, 288.0 * H / 576.0],
                      [0.0, 0.0, 1.0]]).repeat(N, 1, 1)
    warper = ref.Warper(device="cpu")
    pick = list(PICK)
    out = {"frames_u8": vid[pick], "depth_f16": depth16[pick], "pick": np.array(pick), "T1": pose_s[pick].numpy(), "T2": poses[pick].numpy(),
           "K": K[pick].numpy()}
    for j, i in enumerate(pick):
        o = warper.forward_warp(frames[i:i + 1], None, depths[i:i + 1], pose_s[i:i + 1], poses[i:i + 1], K[i:i + 1], None, False, twice=False)
        oc = warper.forward_warp(frames[i:i + 1], None, depths[i:i + 1], pose_s[i:i + 1], poses[i:i + 1], K[i:i + 1], None, True, twice=False)
        f, m, d, fl = [t.contiguous().numpy() for t in o]
        out[f"u{j}/mask_sha"] = digest(m)
        out[f"u{j}/flow_sha"] = digest(fl)
        out[f"u{j}/valid"] = np.array(int(m.sum()))
        out[f"u{j}/frame_b8"] = block_mean(f)
        out[f"u{j}/depth_b8"] = block_mean(d)
        out[f"u{j}/clean_mask_sha"] = digest(oc[1].contiguous().numpy())
        out[f"u{j}/clean_frame_b8"] = block_mean(oc[0].contiguous().numpy())
        print(f"p7 frame {i:2d}: valid {float(m.mean()):.3f} (clean {float(oc[1].mean()):.3f}), depth {float(depths[i].min()):.2f}..{float(depths[i].max()):.2f}")
    path = os.path.join(ROOT, "tests", "golden", "config1_p7.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
