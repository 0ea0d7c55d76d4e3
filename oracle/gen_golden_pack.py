#!/usr/bin/env python
"""Golden vectors for the resize + pack step that follows the warp loop.

The reference has no function for it: the lines are inline in demo.py:91-100,117-120 (and repeated in every
`infer_*` method).  This script runs those torch calls verbatim on seeded inputs in the build container and
writes tests/golden/pack_cases.npz.  TEST INFRASTRUCTURE ONLY.

    python oracle/gen_golden_pack.py
"""
import os
import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = {                      # name: (frames, H, W, h, w)
    "down_1p5": (4, 36, 64, 24, 42),          # the production ratio 576x1024 -> 384x672 in miniature (plus a ragged width)
    "down_prod_ratio": (2, 72, 128, 48, 84),  # exactly 1.5 / 1.5238 like 576x1024 -> 384x672
    "up": (3, 20, 30, 32, 48),
    "same": (2, 16, 24, 16, 24),
    "half": (2, 32, 48, 16, 24),
    "double": (2, 12, 18, 24, 36),
    "odd": (1, 37, 53, 19, 31),
}


def reference_lines(warped, masks, size):
    cond_video = (warped + 1.0) / 2.0                                                   # demo.py:91
    cond_masks = masks                                                                  # demo.py:92
    cond_video = F.interpolate(cond_video, size=size, mode="bilinear", align_corners=False)   # :97-99
    cond_masks = F.interpolate(cond_masks, size=size, mode="nearest")                   # :100
    cond_video = cond_video.permute(1, 0, 2, 3).unsqueeze(0)                            # :119
    cond_masks = (1.0 - cond_masks.permute(1, 0, 2, 3).unsqueeze(0)) * 255.0            # :120
    return cond_video.contiguous(), cond_masks.contiguous()


def main():
    torch.manual_seed(20251104)
    torch.set_num_threads(1)
    out = {}
    for name, (n, H, W, h, w) in CASES.items():
        warped = (torch.rand(n, 3, H, W) * 2 - 1).clamp(-1, 1)
        masks = (torch.rand(n, 1, H, W) > 0.3).float()
        warped = torch.where(masks.bool(), warped, torch.full_like(warped, -1.0))      # holes are -1 like the warp's
        v, m = reference_lines(warped, masks, (h, w))
        out[f"{name}/frame"] = warped.numpy()
        out[f"{name}/mask"] = masks.numpy()
        out[f"{name}/size"] = np.array([h, w], np.int32)
        out[f"{name}/video"] = v.numpy()
        out[f"{name}/masks"] = m.numpy()
    path = os.path.join(ROOT, "tests", "golden", "pack_cases.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes; torch", torch.__version__)


if __name__ == "__main__":
    main()
