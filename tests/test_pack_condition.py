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


# ---------------------------------------------------------------------------------------------
# the fused form: warp loop + pack lines in one submission (tcw_forward_warp_pack)
# ---------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(8, 576, 1024, 384, 672, True), (5, 94, 150, 61, 97, True), (3, 72, 128, 48, 84, False),
                                   (2, 64, 96, 64, 96, True)])
def test_fused_warp_and_pack_equals_the_two_calls(shape):
    """The reference's loop (demo.py:77-100, 117-120) at its own resolutions, at an odd size, with a call-wide normaliser and with
    an identity resize: one submission == forward_warp_batch followed by pack_condition, bit for bit, on two calls in a row;
    a small case is also checked against the CPU oracle chain."""
    import torch
    from trajectorycrafter_b200 import Warper
    from trajectorycrafter_b200.workloads import orbit_units, synthetic_clip
    nf, H, W, h, w, per_frame = shape
    frames, depths = synthetic_clip(nf, H, W, seed=11)
    presets = [(0, 30), (10, -45)]
    T1, T2, K, src = orbit_units(depths, H, W, presets=presets)
    wp = Warper(device="cuda:0")
    f_d, d_d, s_d = frames.cuda(), depths.cuda(), src.cuda()
    warped, mask, _, _ = wp.forward_warp_batch(f_d, d_d, T1, T2, K, src_index=s_d, per_frame_max=per_frame)
    v_ref, m_ref = wp.pack_condition(warped, mask, (h, w))
    for _ in range(2):
        v, m = wp.forward_warp_pack(f_d, d_d, T1, T2, K, None, (h, w), src_index=s_d, per_frame_max=per_frame)
        assert v.shape == (1, 3, len(presets) * nf, h, w) and m.shape == (1, 1, len(presets) * nf, h, w)
        assert torch.equal(v, v_ref), "fused condition video differs from warp + pack"
        assert torch.equal(m, m_ref), "fused condition masks differ from warp + pack"
    if H * W <= 94 * 150:
        from trajectorycrafter_b200 import prepare_matrices
        M = prepare_matrices(T1, T2, K)
        ref = oracle.forward_warp(frames, None, depths, M, src_index=src.numpy(), per_unit_max=per_frame)
        ov, om = oracle.resize_pack(ref[0], ref[1], (h, w))
        assert np.array_equal(m.cpu().numpy(), om)
        assert np.abs(v.cpu().numpy() - ov).max() <= 1e-5


@pytest.mark.gpu
def test_fused_pack_rejects_bad_arguments():
    import torch
    from trajectorycrafter_b200 import Warper, _lib
    from trajectorycrafter_b200.workloads import orbit_units, synthetic_clip
    frames, depths = synthetic_clip(2, 40, 56, seed=1)
    T1, T2, K, src = orbit_units(depths, 40, 56, presets=[(0, 30)])
    wp = Warper(device="cuda:0")
    with pytest.raises(_lib.TcwError):       # the flag is for tcw_workspace_bytes / tcw_forward_warp_pack only
        import ctypes
        out = [torch.empty((2, c, 40, 56), device="cuda:0") for c in (3, 1, 1, 2)]
        M = torch.zeros((2, 30), device="cuda:0")
        ws = torch.empty(1 << 24, dtype=torch.uint8, device="cuda:0")
        p = lambda t: ctypes.c_void_p(t.data_ptr())      # noqa: E731
        rc = wp._lib.tcw_forward_warp(wp._context(), p(frames.cuda()), None, p(depths.cuda()), None, p(M), 2, 2, 40, 56, _lib.TCW_PACK,
                                      p(out[0]), p(out[1]), p(out[2]), p(out[3]), None, p(ws), ws.numel(), None)
        _lib.check(rc, "tcw_forward_warp")
    with pytest.raises(AssertionError):
        wp.forward_warp_pack(frames, depths, T1, T2, K, None, (20, 28), src_index=torch.tensor([0, 5]))
