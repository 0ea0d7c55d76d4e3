"""CPU: properties of the oracle that hold at any size (the GPU suite checks the same ones on the CUDA path at full size):
geometry-only mask and flow, colour-channel permutation equivariance, affine mixing of colours on valid pixels,
holes exactly -1 / 0, and splitting a batch into per-unit calls."""
import numpy as np
import pytest
import torch

import oracle
from trajectorycrafter_b200 import prepare_matrices
from trajectorycrafter_b200.workloads import orbit_units, synthetic_clip


@pytest.fixture(scope="module")
def small():
    frames, depths = synthetic_clip(3, 48, 80, seed=11)
    T1, T2, K, src = orbit_units(depths, 48, 80, presets=[(0, -45), (20, 30)])
    return frames.numpy(), depths.numpy(), prepare_matrices(T1, T2, K).numpy(), src.numpy()


def test_mask_and_flow_depend_on_geometry_only(small):
    frames, depths, M, src = small
    a = oracle.forward_warp(frames, None, depths, M, src_index=src)
    b = oracle.forward_warp(np.ascontiguousarray(frames[:, ::-1]) * 0.5, None, depths, M, src_index=src)
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[3], b[3]) and np.array_equal(a[2], b[2])


def test_channel_permutation_equivariance_and_holes(small):
    frames, depths, M, src = small
    f, m, d, fl, flag = oracle.forward_warp(frames, None, depths, M, src_index=src)
    perm = [2, 0, 1]
    f2 = oracle.forward_warp(np.ascontiguousarray(frames[:, perm]), None, depths, M, src_index=src)[0]
    assert np.array_equal(f2, f[:, perm])
    hole = m[:, 0] == 0
    assert hole.any() and (~hole).any()
    assert np.all(f[:, 0][hole] == -1.0) and np.all(d[:, 0][hole] == 0.0) and np.abs(f).max() <= 1.0
    assert np.array_equal(m > 0, d > 0)


def test_affine_mixing_of_colours(small):
    """warp(a*c1 + (1-a)*c2) == a*warp(c1) + (1-a)*warp(c2) on valid pixels: the weights depend on geometry only."""
    frames, depths, M, src = small
    c2 = np.ascontiguousarray(frames[:, ::-1]) * 0.7
    a = 0.25
    mix = (a * frames + (1 - a) * c2).astype(np.float32)
    w1 = oracle.forward_warp(frames, None, depths, M, src_index=src)
    w2 = oracle.forward_warp(c2, None, depths, M, src_index=src)[0]
    wm = oracle.forward_warp(mix, None, depths, M, src_index=src)[0]
    valid = np.broadcast_to(w1[1] > 0, wm.shape)
    assert np.abs(wm - (a * w1[0] + (1 - a) * w2))[valid].max() <= 2e-6


def test_batch_equals_per_unit_calls(small):
    frames, depths, M, src = small
    whole = oracle.forward_warp(frames, None, depths, M, src_index=src, per_unit_max=True)
    for u in (0, M.shape[0] - 1):
        one = oracle.forward_warp(frames[src[u]:src[u] + 1], None, depths[src[u]:src[u] + 1], M[u:u + 1], per_unit_max=True)
        for k in range(4):
            assert np.array_equal(one[k][0], whole[k][u])


def test_mask1_zero_removes_contributions(small):
    frames, depths, M, src = small
    zero = np.zeros((frames.shape[0], 1) + frames.shape[2:], np.float32)
    f, m, d, fl, flag = oracle.forward_warp(frames, zero, depths, M, src_index=src)
    assert float(m.max()) == 0.0 and np.all(f == -1.0) and np.all(d == 0.0)
    ref_flow = oracle.forward_warp(frames, None, depths, M, src_index=src)[3]
    assert np.array_equal(fl, ref_flow)          # flow12 does not depend on mask1 (models/utils.py:278)
