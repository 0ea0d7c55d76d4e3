"""Resize + pack step after the warp loop (reference demo.py:91-100, 117-120).

CPU: the C oracle against goldens made by running the reference's torch lines (oracle/gen_golden_pack.py).
GPU: the CUDA kernel (through the C ABI) against the goldens and, at production size, against the oracle.
Tolerance: colour 5e-7 absolute on values in [0,1] (ATen's kernel may contract a product into an FMA: <= 2 ulp);
the packed mask ((1 - nearest) * 255) is exact."""
import os

import numpy as np
import pytest

import oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pack_cases.npz")
TOL = 5e-7


def cases():
    g = np.load(GOLD)
    return g, sorted({k.split("/")[0] for k in g.files})


NAMES = cases()[1]


@pytest.mark.parametrize("name", NAMES)
def test_oracle_matches_reference_lines(name):
    g, _ = cases()
    v, m = oracle.resize_pack(g[f"{name}/frame"], g[f"{name}/mask"], g[f"{name}/size"])
    assert v.shape == g[f"{name}/video"].shape and m.shape == g[f"{name}/masks"].shape
    assert np.abs(v - g[f"{name}/video"]).max() <= TOL
    assert np.array_equal(m, g[f"{name}/masks"])
    assert set(np.unique(m)) <= {0.0, 255.0}


def test_oracle_identity_size_is_exact():
    rng = np.random.default_rng(3)
    f = rng.uniform(-1, 1, (2, 3, 9, 13)).astype(np.float32)
    k = (rng.uniform(size=(2, 1, 9, 13)) > 0.5).astype(np.float32)
    v, m = oracle.resize_pack(f, k, (9, 13))
    assert np.array_equal(v[0], ((f + np.float32(1)) / np.float32(2)).transpose(1, 0, 2, 3))
    assert np.array_equal(m[0], ((1 - k) * 255).transpose(1, 0, 2, 3))


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_cuda_matches_goldens(name):
    import torch
    from trajectorycrafter_b200 import Warper
    g, _ = cases()
    w = Warper(device="cuda:0")
    v, m = w.pack_condition(torch.from_numpy(g[f"{name}/frame"]), torch.from_numpy(g[f"{name}/mask"]), g[f"{name}/size"])
    assert tuple(v.shape) == g[f"{name}/video"].shape and tuple(m.shape) == g[f"{name}/masks"].shape
    assert np.abs(v.cpu().numpy() - g[f"{name}/video"]).max() <= TOL
    assert np.array_equal(m.cpu().numpy(), g[f"{name}/masks"])


@pytest.mark.gpu
def test_cuda_production_size_vs_oracle():
    """49 x 576x1024 -> 384x672 (demo.py:545-547 warp resolution -> sample_size), fed by the warp itself."""
    import torch
    from trajectorycrafter_b200 import Warper
    from trajectorycrafter_b200.workloads import synthetic_clip
    torch.manual_seed(5)
    n, H, W = 6, 576, 1024
    frames = (torch.rand(n, 3, H, W) * 2 - 1)
    masks = (torch.rand(n, 1, H, W) > 0.2).float()
    frames = torch.where(masks.bool(), frames, torch.full_like(frames, -1.0))
    w = Warper(device="cuda:0")
    v, m = w.pack_condition(frames.cuda(), masks.cuda(), (384, 672))
    rv, rm = oracle.resize_pack(frames.numpy(), masks.numpy(), (384, 672))
    assert np.abs(v.cpu().numpy() - rv).max() <= TOL
    assert np.array_equal(m.cpu().numpy(), rm)
    # size-independent property: a constant image stays constant, a fully valid mask packs to zeros
    c = torch.full((2, 3, H, W), 0.25, device="cuda")
    one = torch.ones((2, 1, H, W), device="cuda")
    v2, m2 = w.pack_condition(c, one, (384, 672))
    assert float((v2 - 0.625).abs().max()) <= TOL and float(m2.abs().max()) == 0.0


@pytest.mark.gpu
def test_cuda_argument_checks():
    import torch
    from trajectorycrafter_b200 import Warper
    w = Warper(device="cuda:0")
    with pytest.raises(AssertionError):
        w.pack_condition(torch.zeros(2, 3, 8, 8), torch.zeros(2, 1, 8, 9), (4, 4))
    with pytest.raises(RuntimeError):
        w.pack_condition(torch.zeros(2, 3, 8, 8, dtype=torch.float64), torch.zeros(2, 1, 8, 8, dtype=torch.float64), (4, 4))
