"""CPU: host-side logic — pose generators vs reference goldens, matrix block, sharding, workloads."""
import os

import numpy as np
import pytest
import torch

from trajectorycrafter_b200 import (generate_traj_specified, generate_traj_txt, get_poses, prepare_matrices,
                                    sphere2pose)
from trajectorycrafter_b200.sharding import local_sources, shard_bounds, shard_counts
from trajectorycrafter_b200.workloads import ORBIT_PRESETS, orbit_units, ring_cameras, synthetic_clip

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "poses.npz")
C2W = torch.tensor([[-1.0, 0, 0, 0], [0, 1.0, 0, 0], [0, 0, -1.0, 0], [0, 0, 0, 1.0]]).unsqueeze(0)


def test_pose_generators_match_reference_bitwise():
    g = np.load(GOLD)
    got = sphere2pose(C2W, np.float32(12.5), np.float32(-30.0), np.float32(0.7), "cpu", np.float32(0.1), np.float32(-0.2))
    assert np.array_equal(got.numpy(), g["sphere2pose"])
    got = sphere2pose(C2W, np.float32(3.0), np.float32(40.0), np.float32(-0.3), "cpu")
    assert np.array_equal(got.numpy(), g["sphere2pose_noxy"])
    got = generate_traj_specified(C2W, 10.0, -30.0, 0.3 * 4.2, 0.1, -0.05, 49, "cpu")
    assert got.shape == (49, 4, 4) and np.array_equal(got.numpy(), g["traj_specified"])
    for name in ("loop1", "loop2"):
        got = generate_traj_txt(C2W, list(g[f"traj_txt_{name}_phi"]), list(g[f"traj_txt_{name}_theta"]),
                                list(g[f"traj_txt_{name}_r"]), 49, "cpu")
        assert np.array_equal(got.numpy(), g[f"traj_txt_{name}"]), name


def test_sphere2pose_does_not_mutate_input_and_keeps_xy_quirk():
    before = C2W.clone()
    out = sphere2pose(C2W, np.float32(0), np.float32(0), np.float32(0), "cpu", np.float32(0.5), np.float32(0.25))
    assert torch.equal(C2W, before)
    # reference quirk (models/utils.py:89-92): `x is not None` moves row 1 by +y, `y is not None` moves row 0 by -x
    assert out[0, 1, 3] == 0.25 and out[0, 0, 3] == -0.5


def test_get_poses_shapes_and_anchor():
    depths = torch.full((49, 1, 32, 48), 3.0)
    pose_s, pose_t, K = get_poses(depths, 49, target_pose=(0, -30, 0.3, 0, 0))
    assert pose_s.shape == pose_t.shape == (49, 4, 4) and K.shape == (49, 3, 3)
    assert torch.equal(pose_s[7], pose_t[0])            # anchor pose repeated
    assert K[0, 0, 0] == 500 and K[0, 0, 2] == 512 and K[0, 1, 2] == 288
    depths[:] = 100.0
    _, pose_far, _ = get_poses(depths, 49, target_pose=(0, 0, 0, 0, 0))
    assert pose_far[0, 2, 3] == 5.0                      # radius = min(depth * scale, 5)
    _, _, Kz = get_poses(depths, 49, target_pose=(0, 0, 0, 0, 0), f_new=250.0)
    assert Kz[0, 0, 0] == 500 and Kz[-1, 1, 1] == 250


def test_prepare_matrices_layout():
    torch.manual_seed(0)
    T1 = torch.eye(4).repeat(3, 1, 1)
    T1[:, :3, 3] = torch.randn(3, 3)
    T2 = torch.eye(4).repeat(3, 1, 1)
    T2[:, :3, 3] = torch.randn(3, 3)
    K = torch.tensor([[500.0, 0, 512], [0, 500.0, 288], [0, 0, 1]]).repeat(3, 1, 1)
    M = prepare_matrices(T1, T2, K)
    assert M.shape == (3, 30) and M.dtype == torch.float32 and M.is_contiguous()
    assert torch.equal(M[:, 0:9].reshape(3, 3, 3), torch.linalg.inv(K))
    assert torch.equal(M[:, 9:21].reshape(3, 3, 4), torch.bmm(T2, torch.linalg.inv(T1))[:, :3])
    assert torch.equal(M[:, 21:30].reshape(3, 3, 3), K)
    K2 = K * 0.5
    assert torch.equal(prepare_matrices(T1, T2, K, K2)[:, 21:30].reshape(3, 3, 3), K2)


@pytest.mark.parametrize("n,world", [(49, 8), (392, 8), (784, 4), (5, 8), (0, 2), (7, 1)])
def test_shard_bounds_cover_exactly(n, world):
    spans = [shard_bounds(n, world, r) for r in range(world)]
    assert spans[0][0] == 0 and spans[-1][1] == n
    assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
    counts = shard_counts(n, world)
    assert sum(counts) == n and max(counts) - min(counts) <= 1


def test_local_sources_remap():
    src = torch.arange(49, dtype=torch.int32).repeat_interleave(8)
    lo, hi = shard_bounds(392, 8, 3)
    needed, local = local_sources(src, lo, hi)
    assert torch.equal(needed[local.long()].to(torch.int32), src[lo:hi])
    assert len(needed) <= 8


def test_workload_generators_are_deterministic():
    f1, d1 = synthetic_clip(3, 24, 40, seed=1234)
    f2, d2 = synthetic_clip(3, 24, 40, seed=1234)
    assert torch.equal(f1, f2) and torch.equal(d1, d2)
    assert f1.shape == (3, 3, 24, 40) and d1.shape == (3, 1, 24, 40)
    assert float(f1.abs().max()) <= 1.0 and 2.0 < float(d1.min()) and float(d1.max()) < 18.0
    T1, T2, K, src = orbit_units(d1, 24, 40)
    assert T1.shape == (3 * len(ORBIT_PRESETS), 4, 4) and src.tolist()[:9] == [0] * 8 + [1]
    ext, ks = ring_cameras(16)
    assert ext.shape == (16, 4, 4) and ks.shape == (16, 3, 3)
    R = ext[:, :3, :3]
    assert torch.allclose(R @ R.transpose(1, 2), torch.eye(3).expand(16, 3, 3), atol=1e-5)


def test_memoised_inverses_are_the_same_torch_results():
    """prepare_matrices caches the inverses of repeated small matrices (the reference's per-frame loop, demo.py:77-90): the cached
    result must be the tensor torch.linalg.inv returns for that input, batched calls bypass the cache, and the cache stays bounded."""
    import torch
    from trajectorycrafter_b200 import matrices
    g = torch.Generator().manual_seed(4)
    matrices._INV_CACHE.clear()
    K = torch.tensor([[500.0, 0, 336.0], [0, 500.0, 192.0], [0, 0, 1]])[None]
    for i in range(5):
        T1 = torch.eye(4)[None] + 0.1 * torch.rand((1, 4, 4), generator=g)
        T2 = torch.eye(4)[None] + 0.1 * torch.rand((1, 4, 4), generator=g)
        for _ in range(2):                                   # second round: served from the cache
            M = matrices.prepare_matrices(T1, T2, K, None)
            assert torch.equal(M[:, 9:21], torch.bmm(T2, torch.linalg.inv(T1))[:, :3].reshape(1, 12))
            assert torch.equal(M[:, 0:9], torch.linalg.inv(K).reshape(1, 9))
    assert len(matrices._INV_CACHE) == 6                     # five source poses + one intrinsic matrix
    Tb = torch.eye(4).repeat(12, 1, 1) + 0.1 * torch.rand((12, 4, 4), generator=g)
    before = len(matrices._INV_CACHE)
    Mb = matrices.prepare_matrices(Tb, Tb.flip(0), K.repeat(12, 1, 1))
    assert len(matrices._INV_CACHE) == before                # more than 8 matrices: one LAPACK call, no hashing
    assert torch.equal(Mb[3:4], matrices.prepare_matrices(Tb[3:4], Tb.flip(0)[3:4], K))      # batched == one by one
    for i in range(matrices._INV_CACHE_MAX + 10):
        matrices._inv_cached(torch.eye(3)[None] * (i + 2))
    assert len(matrices._INV_CACHE) <= matrices._INV_CACHE_MAX
