#!/usr/bin/env python
"""Golden vectors for the depth post-processing lines, container only.

The reference has no function for them: they are the tail of DepthCrafterDemo.infer (models/infer.py:79-92), which cannot run
here as a whole (the DepthCrafter pipeline needs diffusers + checkpoints).  This script executes THOSE LINES verbatim (copied
below between the markers, numpy + torch exactly as the reference calls them) on seeded stand-ins for the pipeline's
`res` array, and the radius lines of demo.py:539-543, and stores inputs + outputs in tests/golden/depth_cases.npz.
TEST INFRASTRUCTURE ONLY.

    python oracle/gen_golden_depth.py
"""
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def reference_lines(res, near, far, radius_scale):
    # ---- models/infer.py:79-92, verbatim
    res = res.sum(-1) / res.shape[-1]
    depths = (res - res.min()) / (res.max() - res.min())
    depths = torch.from_numpy(depths).unsqueeze(1)
    depths *= 3900
    depths[depths < 1e-5] = 1e-5
    depths = 10000.0 / depths
    depths = depths.clip(near, far)
    # ---- demo.py:539-543, verbatim
    radius = depths[0, 0, depths.shape[-2] // 2, depths.shape[-1] // 2].cpu() * radius_scale
    radius = min(radius, 5)
    return depths, radius


def main():
    rng = np.random.RandomState(7)
    out = {}
    cases = [("small", (3, 20, 36, 3), 1e-4, 1e4, 1.0), ("clip", (2, 48, 64, 3), 5.0, 60.0, 0.7), ("odd", (5, 37, 53, 3), 1e-4, 1e4, 2.5),
             ("video", (7, 96, 160, 3), 1e-4, 1e4, 1.0)]
    for name, shape, near, far, scale in cases:
        ys = np.linspace(0, 1, shape[1], dtype=np.float32)[None, :, None, None]
        xs = np.linspace(0, 1, shape[2], dtype=np.float32)[None, None, :, None]
        res = (0.2 + 0.5 * np.exp(-((xs - 0.5) ** 2 + (ys - 0.6) ** 2) / 0.05) + 0.2 * ys
               + 0.02 * rng.rand(*shape).astype(np.float32)).astype(np.float32)
        if name == "clip":
            res[0, :4, :4] = res.min()          # exercises the 1e-5 floor and the far clip
        depths, radius = reference_lines(res.copy(), near, far, scale)
        out[f"{name}/res"] = res
        out[f"{name}/params"] = np.array([near, far, scale], dtype=np.float64)
        out[f"{name}/depths"] = depths.numpy()
        out[f"{name}/radius"] = np.array(float(radius), dtype=np.float32)
        print(name, shape, "depth", float(depths.min()), float(depths.max()), "radius", float(radius))
    path = os.path.join(ROOT, "tests", "golden", "depth_cases.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
