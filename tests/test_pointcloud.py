"""the fork's GlobalPointCloudWarper (reference notebooks/06_10_25_vggt/warper_point_cloud.py) and
VisualizationWarper.extract_3d_points_with_colors (notebooks/28_08_25_trajectories/core.py:15-79).
CPU: the numpy oracle against outputs of the unmodified reference (oracle/gen_golden_pointcloud.py).
GPU: the CUDA kernels through the host mirror against the same fixtures, and properties at 1 M points."""
import os

import numpy as np
import pytest
import torch

import oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pointcloud_cases.npz")
H, W = 60, 96
BIL_TOL = 1e-5          # depth weights go through exp(): SLEEF in the reference, CUDA's expf here


@pytest.fixture(scope="module")
def z():
    return np.load(GOLD)


def test_oracle_lift_and_extract(z):
    K, T1 = torch.from_numpy(z["K"]), torch.from_numpy(z["T1"])
    pts, cols = oracle.pc_lift(z["frame"], z["mask"], z["depth"], torch.linalg.inv(T1).numpy(), torch.linalg.inv(K).numpy())
    assert np.array_equal(pts, z["lift/points"]) and np.array_equal(cols, z["lift/colors"])
    st = 4
    ps, cs = oracle.pc_lift(z["frame"][:, :, ::st, ::st], None, z["depth"][:, :, ::st, ::st], z["T1"], torch.linalg.inv(K).numpy(), step=st)
    valid = z["depth"][:, 0, ::st, ::st] > 0
    assert np.array_equal(ps[valid], z["extract/points"]) and np.array_equal(cs[valid], z["extract/colors"])


@pytest.mark.parametrize("cam", [0, 1])
def test_oracle_renderers(z, cam):
    P, C, Wt = z["cloud/points"], z["cloud/colors"], z["cloud/weights"]
    for ps in (1, 2, 3):
        f, m, d = oracle.pc_zbuffer(P, C, z["Tt"][cam], z["Kt"][cam], H, W, ps)
        assert np.array_equal(f, z[f"zb/cam{cam}_ps{ps}/frame"]), "z-buffer winners must be bit-exact"
        assert np.array_equal(m, z[f"zb/cam{cam}_ps{ps}/mask"]) and np.array_equal(d, z[f"zb/cam{cam}_ps{ps}/depth"])
    f, m, _ = oracle.pc_zbuffer(P, C, z["Tt"][cam], z["Kt"][cam], H, W, 1, clamp=True)
    assert np.array_equal(f, z[f"zbc/cam{cam}/frame"]) and np.array_equal(m, z[f"zbc/cam{cam}/mask"])
    f, m = oracle.pc_bilinear(P, C, Wt, z["Tt"][cam], z["Kt"][cam], H, W)
    assert np.array_equal(m[None], z["bil/mask"][cam:cam + 1])
    assert np.abs(f - z["bil/frame"][cam]).max() <= BIL_TOL
    for ps in (1, 2, 3):
        f, m, _ = oracle.pc_zbuffer(P, C, z["Tt"][cam], z["Kt"][cam], H, W, ps, clamp=ps <= 1, first_wins=True)
        assert np.array_equal(f, z[f"zbf/cam{cam}_ps{ps}/frame"]), "sequential z-buffer: the first of equal depths stays"
        assert np.array_equal(m, z[f"zbf/cam{cam}_ps{ps}/mask"])
    for ps in (2.0, 3.0, 5.0):
        f, m = oracle.pc_disc(P, C, Wt, z["Tt"][cam], z["Kt"][cam], H, W, ps)
        assert np.array_equal(m[None], z[f"disc/ps{ps:g}/mask"][cam:cam + 1])
        assert np.abs(f - z[f"disc/ps{ps:g}/frame"][cam]).max() <= BIL_TOL


def test_first_and_last_winner_rules_differ_on_the_duplicated_points(z):
    """The fixture ends with duplicated points in other colours: the two z-buffer rules (last / first of equal depths) must
    disagree in some pixels and nowhere in the mask."""
    for cam in (0, 1):
        for ps in (2, 3):
            last, first = z[f"zb/cam{cam}_ps{ps}/frame"], z[f"zbf/cam{cam}_ps{ps}/frame"]
            assert (last != first).any(axis=1).sum() >= 10
            assert np.array_equal(z[f"zb/cam{cam}_ps{ps}/mask"], z[f"zbf/cam{cam}_ps{ps}/mask"])


# ---------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def pcw():
    assert torch.cuda.is_available(), "these tests need the B200"
    from trajectorycrafter_b200 import GlobalPointCloudWarper
    return GlobalPointCloudWarper(device="cuda:0")


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


@pytest.mark.gpu
def test_gpu_lift_extract_bit_exact(pcw, z):
    pts, cols = pcw.lift_to_3d_pointcloud(T(z["frame"]), T(z["mask"]), T(z["depth"]), T(z["T1"]), T(z["K"]))
    assert np.array_equal(pts.cpu().numpy(), z["lift/points"]) and np.array_equal(cols.cpu().numpy(), z["lift/colors"])
    ep, ec = pcw.extract_3d_points_with_colors(T(z["frame"]), T(z["depth"]), T(z["T1"]), T(z["K"]), subsample_step=4)
    assert np.array_equal(ep.cpu().numpy(), z["extract/points"]) and np.array_equal(ec.cpu().numpy(), z["extract/colors"])
    p0, c0, w0 = pcw.create_pointcloud_from_image(T(z["frame"][:1]), T(z["mask"][:1]), T(z["depth"][:1]), T(z["T1"][:1]), T(z["K"][:1]), 1.0)
    assert p0.shape[0] == int(z["mask"][0].sum()) and w0.shape == (p0.shape[0], 1)


@pytest.mark.gpu
@pytest.mark.parametrize("cam", [0, 1])
def test_gpu_zbuffer_bit_exact(pcw, z, cam):
    P, C = T(z["cloud/points"]), T(z["cloud/colors"])
    Tt, Kt = T(z["Tt"][cam:cam + 1]), T(z["Kt"][cam:cam + 1])
    for ps in (1, 2, 3):
        f, m, d = pcw.render_pointcloud_zbuffer_vectorized_point_size(P, C, Tt, Kt, (H, W), ps, return_depth=True)
        assert np.array_equal(f.cpu().numpy(), z[f"zb/cam{cam}_ps{ps}/frame"]), f"point_size {ps}: winners must be bit-exact"
        assert np.array_equal(m.cpu().numpy(), z[f"zb/cam{cam}_ps{ps}/mask"]) and np.array_equal(d.cpu().numpy(), z[f"zb/cam{cam}_ps{ps}/depth"])
    f, m = pcw.render_pointcloud_zbuffer_vectorized(P, C, Tt, Kt, (H, W))
    assert np.array_equal(f.cpu().numpy(), z[f"zbc/cam{cam}/frame"]) and np.array_equal(m.cpu().numpy(), z[f"zbc/cam{cam}/mask"])


@pytest.mark.gpu
def test_gpu_render_from_camera(pcw, z):
    f, m = pcw.render_from_camera(T(z["cloud/points"]), T(z["cloud/colors"]), T(z["cloud/weights"]), T(z["Tt"]), T(z["Kt"]), H, W)
    assert np.array_equal(m.cpu().numpy(), z["bil/mask"]), "coverage mask must be bit-exact"
    err = np.abs(f.cpu().numpy() - z["bil/frame"]).max()
    print(f"render_from_camera: colour max |d| {err:.2e}")
    assert err <= BIL_TOL
    e0, e1 = pcw.render_from_camera(None, None, None, T(z["Tt"]), T(z["Kt"]), H, W)
    assert e0.shape == (2, 3, H, W) and float(e0.abs().max()) == 0.0 and float(e1.max()) == 0.0


@pytest.mark.gpu
@pytest.mark.parametrize("cam", [0, 1])
def test_gpu_zbuffer_fixed_bit_exact(pcw, z, cam):
    P, C = T(z["cloud/points"]), T(z["cloud/colors"])
    Tt, Kt = T(z["Tt"][cam:cam + 1]), T(z["Kt"][cam:cam + 1])
    for ps in (1, 2, 3):
        f, m = pcw.render_pointcloud_zbuffer_vectorized_fixed(P, C, Tt, Kt, (H, W), ps)
        assert np.array_equal(f.cpu().numpy(), z[f"zbf/cam{cam}_ps{ps}/frame"]), f"point_size {ps}: first-of-equal winners must be bit-exact"
        assert np.array_equal(m.cpu().numpy(), z[f"zbf/cam{cam}_ps{ps}/mask"])


@pytest.mark.gpu
def test_gpu_render_from_camera_disc(pcw, z):
    for ps in (2.0, 3.0, 5.0):
        f, m = pcw.render_from_camera(T(z["cloud/points"]), T(z["cloud/colors"]), T(z["cloud/weights"]), T(z["Tt"]), T(z["Kt"]), H, W,
                                      point_size=ps, use_bilinear=False)
        assert np.array_equal(m.cpu().numpy(), z[f"disc/ps{ps:g}/mask"]), "coverage mask must be bit-exact"
        err = np.abs(f.cpu().numpy() - z[f"disc/ps{ps:g}/frame"]).max()
        print(f"render_from_camera(use_bilinear=False, point_size={ps:g}): colour max |d| {err:.2e}")
        assert err <= BIL_TOL
    with pytest.raises(RuntimeError):
        pcw.render_from_camera(T(z["cloud/points"]), T(z["cloud/colors"]), T(z["cloud/weights"]), T(z["Tt"]), T(z["Kt"]), H, W,
                               point_size=30.0, use_bilinear=False)          # 709 offsets: more than the table holds


@pytest.mark.gpu
def test_gpu_zbuffer_one_million_points(pcw):
    """1 M points at 576x1024: reproducible bit for bit, against the oracle, and order-independent (a shuffled cloud with
    distinct depths renders the same image)."""
    g = torch.Generator().manual_seed(9)
    n = 1_000_000
    P = torch.rand((n, 3), generator=g) * torch.tensor([8.0, 5.0, 6.0]) - torch.tensor([4.0, 2.5, 0.0]) + torch.tensor([0.0, 0.0, 4.0])
    C = torch.rand((n, 3), generator=g) * 2 - 1
    Tt = torch.eye(4)[None].clone()
    Tt[0, :3, 3] = torch.tensor([0.1, -0.05, 0.2])
    Kt = torch.tensor([[500.0, 0, 512.0], [0, 500.0, 288.0], [0, 0, 1]])[None]
    a = pcw.render_pointcloud_zbuffer_vectorized_point_size(P, C, Tt, Kt, (576, 1024), 2, return_depth=True)
    b = pcw.render_pointcloud_zbuffer_vectorized_point_size(P, C, Tt, Kt, (576, 1024), 2, return_depth=True)
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    f, m, d = oracle.pc_zbuffer(P.numpy(), C.numpy(), Tt[0].numpy(), Kt[0].numpy(), 576, 1024, 2)
    assert np.array_equal(a[0].cpu().numpy(), f) and np.array_equal(a[2].cpu().numpy(), d)
    perm = torch.randperm(n, generator=g)
    c = pcw.render_pointcloud_zbuffer_vectorized_point_size(P[perm], C[perm], Tt, Kt, (576, 1024), 1, return_depth=True)
    e = pcw.render_pointcloud_zbuffer_vectorized_point_size(P, C, Tt, Kt, (576, 1024), 1, return_depth=True)
    assert torch.equal(c[2], e[2])                                   # the depth image never depends on the order
    same = (c[0] == e[0]).all(dim=1).float().mean()
    assert float(same) > 0.9999                                     # colours only where two points share a depth exactly
