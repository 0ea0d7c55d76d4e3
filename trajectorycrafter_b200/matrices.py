"""Host-side preparation of the per-unit matrix block consumed by libtcwarp (include/tcwarp.h).

The reference computes ``T = bmm(T2, inv(T1))`` and ``inv(K1)`` with torch on the CPU
(reference models/utils.py:366-368, 382).  Those results feed ``floor``/``ceil`` decisions, so a
device-side closed-form inverse (last-ulp differences) would flip ~1e-4 of the integer pixel
indices.  The host therefore runs the *same torch CPU calls* on the 4x4/3x3
inputs (microseconds for hundreds of units) and uploads 30 floats per unit.
"""
from __future__ import annotations

from typing import Optional

import torch

M_FLOATS = 30

# The reference's own loop calls forward_warp once per frame with the SAME source pose and intrinsics every time
# (demo.py:77-90): the two small inverses are memoised on the bytes of their input, so a repeated matrix costs a dictionary
# lookup instead of a LAPACK call.  The cached tensor is the result of the same torch call on the same input: bit-identical.
_INV_CACHE: dict = {}
_INV_CACHE_MAX = 256


def _inv_cached(m: torch.Tensor) -> torch.Tensor:
    if m.shape[0] > 8:                       # batched submissions: one LAPACK call is cheaper than hashing
        return torch.linalg.inv(m)
    key = (tuple(m.shape), m.contiguous().numpy().tobytes())
    hit = _INV_CACHE.get(key)
    if hit is None:
        if len(_INV_CACHE) >= _INV_CACHE_MAX:
            _INV_CACHE.clear()
        hit = _INV_CACHE[key] = torch.linalg.inv(m)
    return hit


def prepare_matrices(transformation1: torch.Tensor, transformation2: torch.Tensor,
                     intrinsic1: torch.Tensor, intrinsic2: Optional[torch.Tensor] = None) -> torch.Tensor:
    """(n,4,4),(n,4,4),(n,3,3),(n,3,3)|None -> (n,30) float32 CPU tensor [K1^-1 | T[:3] | K2]."""
    t1 = transformation1.detach().to("cpu", torch.float32)
    t2 = transformation2.detach().to("cpu", torch.float32)
    k1 = intrinsic1.detach().to("cpu", torch.float32)
    k2 = k1.clone() if intrinsic2 is None else intrinsic2.detach().to("cpu", torch.float32)
    n = t1.shape[0]
    assert t1.shape == (n, 4, 4) and t2.shape == (n, 4, 4)
    assert k1.shape == (n, 3, 3) and k2.shape == (n, 3, 3)
    rel = torch.bmm(t2, _inv_cached(t1))             # models/utils.py:366-368
    k1_inv = _inv_cached(k1)                         # models/utils.py:382
    out = torch.empty((n, M_FLOATS), dtype=torch.float32)
    out[:, 0:9] = k1_inv.reshape(n, 9)
    out[:, 9:21] = rel[:, :3, :].reshape(n, 12)
    out[:, 21:30] = k2.reshape(n, 9)
    return out.contiguous()
