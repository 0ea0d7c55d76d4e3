"""GPU: every BASELINE.json config and every flag of the path at size (CUDA through the C ABI vs the reference-pinned
fixtures and the CPU oracle).  One test per config is named test_config{1..5}_*."""
import numpy as np
import pytest
import torch

import oracle
from _util import (assert_colour_close, assert_depth_close, bit_equal, block_mean, load_config1, parity_report, sha_digest)
from trajectorycrafter_b200 import Warper, prepare_matrices
from trajectorycrafter_b200.workloads import (ORBIT_PRESETS, autoregressive_segments, multicam_units, orbit_units, paired_units,
                                              scaled_intrinsics, synthetic_clip)

pytestmark = pytest.mark.gpu
H, W = 384, 672


@pytest.fixture(scope="module")
def warper():
    assert torch.cuda.is_available(), "these tests need the B200"
    return Warper(device="cuda:0")


@pytest.fixture(scope="module")
def clip():
    return synthetic_clip(49, H, W, seed=1234)


def npy(t):
    return None if t is None else t.detach().cpu().numpy()


def check(got, ref, what, tol=1e-5):
    gf, gm, gd, gfl = [npy(t) for t in got]
    rf, rm, rd, rfl = ref[:4]
    assert bit_equal(gm.astype(rm.dtype), rm), f"{what}: validity mask must be bit-exact"
    assert bit_equal(gfl, rfl), f"{what}: flow12 must be bit-exact"
    count = oracle.contrib_count(rfl)
    assert_colour_close(gf, rf, tol=tol, count=count, what=what + " colour")
    assert_depth_close(gd, rd, rm, rtol=tol, count=count, what=what + " depth")
    print(what, parity_report(gf, rf, rm, rd, count, got_depth=gd))


def test_config1_p7_clip_fullsize(warper):
    """Config 1: frames 0 / 24 / 48 of the reference's own test/videos/p7.mp4 at 384x672, run.sh preset, against the digests of
    the UNMODIFIED reference's outputs (tests/golden/config1_p7.npz) and against the oracle at full resolution; mask=False and
    mask=True, through the device entry and through the host-buffer entry."""
    frames, depths, z = load_config1()
    T1, T2, K = torch.from_numpy(z["T1"]), torch.from_numpy(z["T2"]), torch.from_numpy(z["K"])
    M = prepare_matrices(T1, T2, K, None)
    ref = oracle.forward_warp(frames, None, depths, M, per_unit_max=True)
    got = warper.forward_warp_batch(frames, depths, T1, T2, K)
    check(got, ref, "config1")
    gf, gm, gd, gfl = [npy(t) for t in got]
    for j in range(3):
        assert np.array_equal(sha_digest(gm[j:j + 1]), z[f"u{j}/mask_sha"]), "mask differs from the reference's"
        assert np.array_equal(sha_digest(gfl[j:j + 1]), z[f"u{j}/flow_sha"]), "flow12 differs from the reference's"
        assert np.abs(block_mean(gf[j:j + 1]) - z[f"u{j}/frame_b8"]).max() <= 1e-5
        one = warper.forward_warp(frames[j:j + 1], None, depths[j:j + 1], T1[j:j + 1], T2[j:j + 1], K[j:j + 1], None, mask=True)
        assert one[0].dtype == torch.float64
        assert np.array_equal(sha_digest(npy(one[1])), z[f"u{j}/clean_mask_sha"]), "mask=True mask differs from the reference's"
        assert np.abs(block_mean(npy(one[0])) - z[f"u{j}/clean_frame_b8"]).max() <= 1e-5
    host = warper.forward_warp_host(frames.contiguous(), depths.contiguous(), M, mask=True)
    assert host[4] == 0 and host[0].dtype == torch.float64
    for j in range(3):
        assert np.array_equal(sha_digest(host[1][j:j + 1].numpy()), z[f"u{j}/clean_mask_sha"])


def test_config2_mask_true_all_presets(warper, clip):
    """Config 2 as inference_orbits.py:281 runs it: mask=True for all eight presets at 384x672 (frame 24) -- the fused
    clean_points against the oracle's; float64 outputs, cleaned pixels exactly -1."""
    frames, depths = clip
    T1, T2, K, src = orbit_units(depths, H, W)
    pick = torch.arange(8) + 8 * 24
    M = prepare_matrices(T1[pick], T2[pick], K[pick])
    got = warper.forward_warp_matrices(frames, depths, M, src_index=src[pick], mask=True)
    ref = oracle.forward_warp(frames, None, depths, M, src_index=src[pick].numpy(), per_unit_max=True)
    gf, gm = npy(got[0]), npy(got[1])
    for j in range(8):
        cf, cm = oracle.clean_points(ref[0][j:j + 1], ref[1][j:j + 1])
        assert bit_equal(gm[j:j + 1], cm), f"preset {j}: cleaned mask must be bit-exact"
        assert_colour_close(gf[j:j + 1], cf)
        assert np.all(gf[j:j + 1][np.broadcast_to(cm == 0, cf.shape)] == -1.0)
    assert bit_equal(npy(got[3]), ref[3])


@pytest.mark.parametrize("mode", ["direct", "bullet"])
def test_config2_direct_and_bullet_pairings(warper, clip, mode):
    """The frame x pose pairings of infer_direct (demo.py:203-231, cut=20) and infer_bullet (:329-342) as src_index of one
    submission, all 49 units at 384x672, against the oracle."""
    frames, depths = clip
    T1, T2, K, src = paired_units(depths, (0, 30, 0.3, 0, 0), H, W, mode)
    pick = torch.tensor([0, 7, 19, 20, 33, 48])
    M = prepare_matrices(T1, T2, K)
    got = warper.forward_warp_matrices(frames, depths, M, src_index=src)
    ref = oracle.forward_warp(frames, None, depths, M[pick], src_index=src[pick].numpy(), per_unit_max=True)
    check([t[pick.to(t.device)] for t in got], ref, f"{mode} pairing")


def test_config3_1080p_clip(warper):
    """Config 3: 1080p, single (0,30,0.3) trajectory, K = f 1500 / 960 / 540; frames 0 and 40 of a 49-frame ramp."""
    frames, depths = synthetic_clip(49, 1080, 1920, seed=5)
    T1, T2, K, src = orbit_units(depths, 1080, 1920, d_r=0.3, presets=[(0, 30)])
    K[:] = torch.tensor([[1500.0, 0, 960], [0, 1500.0, 540], [0, 0, 1]])
    pick = torch.tensor([0, 40])
    M = prepare_matrices(T1[pick], T2[pick], K[pick])
    got = warper.forward_warp_matrices(frames, depths, M, src_index=src[pick])
    ref = oracle.forward_warp(frames, None, depths, M, src_index=src[pick].numpy(), per_unit_max=True)
    check(got, ref, "config3")


@pytest.mark.parametrize("smooth", [False, True], ids=["as_is", "smooth"])
def test_config4_multicam(warper, clip, smooth):
    """Config 4: 16 ring cameras x 49 frames = 784 units in one submission (K_t != K_s; `smooth` = SLERP pose + linear K
    walk of run_w_cam_poses_mvtracker_smooth.py:176-256); 20 units spread over frames and views against the oracle."""
    frames, depths = clip
    d = (depths / depths.max() * 10.0 + 2.0).contiguous()              # GT-like depth in [2, 12]
    T1, T2, K1, K2, src = multicam_units(49, 16, smooth=smooth, h=H, w=W)
    M = prepare_matrices(T1, T2, K1, K2)
    got = warper.forward_warp_matrices(frames, d, M, src_index=src, return_depth=True, return_flow=True)
    pick = torch.arange(0, 784, 41)
    ref = oracle.forward_warp(frames, None, d, M[pick], src_index=src[pick].numpy(), per_unit_max=True)
    check([t[pick.to(t.device)] for t in got], ref, "config4" + ("_smooth" if smooth else ""))


def test_config5_autoregressive_chain(warper):
    """Config 5: 8 chained 49-frame segments at 384x672 (inference_autoregressive.py: target (0, 90, 1, 0, 0) in 8 equal
    deltas, pose continuation of core_autoregressive.py:173-235 incl. the segments >= 2 quirk), 8 dependent batches;
    frames 0 / 24 / 48 of every segment against the oracle."""
    clips = [synthetic_clip(49, H, W, seed=1234 + k) for k in range(8)]
    segs = autoregressive_segments([c[1] for c in clips], (0, 90, 1.0, 0, 0), n_splits=8)
    fx, fy, cx, cy = scaled_intrinsics(H, W)
    K = torch.tensor([[fx, 0.0, cx], [0.0, fy, cy], [0.0, 0.0, 1.0]]).repeat(49, 1, 1)
    assert torch.equal(segs[2][0], segs[1][0]) and torch.equal(segs[5][1], segs[1][1])      # the quirk
    pick = torch.tensor([0, 24, 48])
    for k, ((frames, depths), (pose_s, pose_t, _)) in enumerate(zip(clips, segs)):
        M = prepare_matrices(pose_s, pose_t, K)
        got = warper.forward_warp_matrices(frames, depths, M)
        ref = oracle.forward_warp(frames[pick], None, depths[pick], M[pick], per_unit_max=True)
        check([t[pick.to(t.device)] for t in got], ref, f"config5 segment {k}")


def test_host_entry_mask1_and_mask_true(warper, clip):
    """forward_warp_host with mask1 holes (the only shipped caller with mask1 is core.py:146-156) and with mask=True."""
    frames, depths = clip
    T1, T2, K, src = orbit_units(depths[:3], H, W, presets=ORBIT_PRESETS[:2])
    M = prepare_matrices(T1, T2, K)
    g = torch.Generator().manual_seed(3)
    mask1 = (torch.rand((3, 1, H, W), generator=g) > 0.3).float().contiguous()
    host = warper.forward_warp_host(frames[:3].contiguous(), depths[:3].contiguous(), M, src_index=src, mask1=mask1)
    assert host[4] == 0
    ref = oracle.forward_warp(frames[:3], mask1, depths[:3], M, src_index=src.numpy(), per_unit_max=True)
    check(host[:4], ref, "host entry + mask1")
    hc = warper.forward_warp_host(frames[:3].contiguous(), depths[:3].contiguous(), M[:1].contiguous(), src_index=src[:1], mask=True,
                                  return_depth=False, return_flow=False)
    ref1 = oracle.forward_warp(frames[:3], None, depths[:3], M[:1], src_index=src[:1].numpy(), per_unit_max=True)
    cf, cm = oracle.clean_points(ref1[0], ref1[1])
    assert hc[0].dtype == torch.float64 and hc[2] is None and hc[3] is None
    assert bit_equal(hc[1].numpy(), cm)
    assert_colour_close(hc[0].numpy(), cf)


def test_twice_fullsize(warper, clip):
    """forward_warp(twice=True) (models/utils.py:296-347) at 384x672 against the oracle's chain of the same five splats.
    The backward splats read a flow that is itself a splat result, so last-bit differences of the first stage can move a
    backward sample across a cell border: the test prints how many mask pixels differ and the worst colour error where the
    masks agree instead of assuming the single-splat bound."""
    frames, depths = clip
    T1, T2, K, _ = orbit_units(depths[:1], H, W, presets=[(0, 12)])
    got = warper.forward_warp(frames[:1], None, depths[:1], T1, T2, K, None, mask=False, twice=True)
    assert got[3] is None
    M = prepare_matrices(T1, T2, K)
    f1, m1, d1, fl, _ = oracle.forward_warp(frames[:1], None, depths[:1], M, per_unit_max=False)
    zt = np.ascontiguousarray(oracle.project(depths[:1], M)[..., 2])                    # trans_depth1 (models/utils.py:276)
    ones = np.ones_like(m1)
    wfl, _, _ = oracle.splat(fl, ones, zt, fl, None, is_image=False)
    tf, tm, _ = oracle.splat(f1, m1, np.ascontiguousarray(d1[:, 0]), -wfl, None, is_image=True)
    td, _, _ = oracle.splat(d1, m1, np.ascontiguousarray(d1[:, 0]), -wfl, None, is_image=False)
    gf, gm, gd = npy(got[0]), npy(got[1]), npy(got[2])
    mism = int((gm != tm).sum())
    agree = np.broadcast_to(gm == tm, tf.shape)
    err = np.abs(gf.astype(np.float64) - tf)[agree].max()
    print(f"twice at 384x672: mask mismatches {mism} of {gm.size}, colour max |d| where masks agree {err:.3e}")
    assert mism <= 4, "twice: validity masks differ"
    assert err <= 2e-5
    valid = (gm > 0) & (tm > 0)
    assert (np.abs(gd - td)[valid] <= 2e-5 * np.abs(td[valid])).all()


def test_u8_mask_into_exchange_buffer(warper, clip):
    """The fused exchange writes colour f32 + mask u8 (13 B/px) straight into the destination rank's buffer: here the
    owner side on one GPU (tcw_ipc_alloc'd memory, slices at a unit offset), against the plain outputs bit for bit."""
    from trajectorycrafter_b200.sharding import PeerBuffer, packed_buffer, packed_views
    frames, depths = clip
    T1, T2, K, src = orbit_units(depths[:4], H, W, presets=ORBIT_PRESETS[:2])
    M = prepare_matrices(T1, T2, K)
    plain = warper.forward_warp_matrices(frames[:4], depths[:4], M, src_index=src)
    buf = PeerBuffer(warper, 12, H, W, dst=0)
    try:
        col, msk = buf.views(3, 11)
        got = warper.forward_warp_matrices(frames[:4], depths[:4], M, src_index=src, out=(col, msk, None, None), mask_u8=True)
        torch.cuda.synchronize()
        assert got[1].dtype == torch.uint8
        assert torch.equal(buf.colour[3:11], plain[0]) and torch.equal(buf.mask[3:11].float(), plain[1])
    finally:
        buf.close()
    packed = packed_buffer(8, H, W, warper.device)
    pc, pm = packed_views(packed, 8, H, W)
    warper.forward_warp_matrices(frames[:4], depths[:4], M, src_index=src, out=(pc, pm, None, None), mask_u8=True)
    assert torch.equal(pc, plain[0]) and torch.equal(pm.float(), plain[1])


@pytest.mark.parametrize("name", ["small", "clip", "odd", "video"])
def test_depth_postprocess_on_device(warper, name):
    """tcw_depth_post against the reference's own lines (models/infer.py:79-92, demo.py:539-543) bit for bit;
    the radius reaches the pose generator through 16 pinned bytes and an event, the depth map never leaves the device."""
    import os
    from trajectorycrafter_b200.poses import get_poses
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "depth_cases.npz"))
    near, far, scale = [float(v) for v in z[name + "/params"]]
    depths, stats = warper.depth_postprocess(torch.from_numpy(z[name + "/res"]), near, far, scale)
    stats.ready.synchronize()                      # 16 bytes, not the stream
    assert depths.is_cuda and bit_equal(npy(depths), z[name + "/depths"])
    assert float(stats[3]) == float(z[name + "/radius"])
    od, ost = oracle.depth_post(z[name + "/res"], near, far, scale)
    assert np.array_equal(stats.numpy(), ost)
    n = depths.shape[0]
    a = get_poses(depths.cpu(), n, target_pose=(0, 30, 0.3, 0, 0), radius_scale=scale)
    b = get_poses(None, n, target_pose=(0, 30, 0.3, 0, 0), radius_scale=scale, centre_depth=float(stats[2]))
    for x, y in zip(a, b):
        assert torch.equal(x, y), "poses from the device-side centre depth differ from the read-back path"


def test_src_index_is_checked_on_the_device(warper, clip):
    """a CUDA src_index is range-checked by the kernels that read it (no host cache that an address-reusing
    allocator could fool): the same storage, first in range, then out of range -- the second call must raise and must not
    read out of bounds (the bad unit reads source 0)."""
    frames, depths = clip
    T1, T2, K, _ = orbit_units(depths[:2], H, W, presets=ORBIT_PRESETS[:1])
    M = prepare_matrices(T1, T2, K)
    idx = torch.tensor([0, 1], dtype=torch.int32, device=warper.device)
    ok = warper.forward_warp_matrices(frames[:2], depths[:2], M, src_index=idx)
    idx.copy_(torch.tensor([0, 7], dtype=torch.int32))                 # same address, same version-less storage, now out of range
    with pytest.raises(IndexError):
        warper.forward_warp_matrices(frames[:2], depths[:2], M, src_index=idx)
    flag = torch.zeros(1, dtype=torch.int32, device=warper.device)
    got = warper.forward_warp_matrices(frames[:2], depths[:2], M, src_index=idx, check_range=flag)
    assert int(flag.item()) & 4
    assert torch.equal(got[1][0], ok[1][0]) and torch.equal(got[3][0], ok[3][0])      # the in-range unit is untouched


@pytest.mark.parametrize("mode", ["window", "atomic"])
@pytest.mark.parametrize("shape", [(75, 131), (76, 131), (75, 130), (33, 65)])
def test_odd_sizes_vs_oracle(warper, shape, mode):
    """Odd heights / widths (h * w not a multiple of 4): every mode, with mask1 holes and a camera that puts part of the
    image behind itself.  Regression for a misaligned accumulator lane of the float-reduction mode found by the sanitizer run."""
    h, w = shape
    frames, depths = synthetic_clip(2, h, w, seed=3)
    T1, T2, K, src = orbit_units(depths, h, w, presets=[(0, 60), (25, 170)])
    M = prepare_matrices(T1, T2, K)
    g = torch.Generator().manual_seed(1)
    mask1 = (torch.rand((2, 1, h, w), generator=g) > 0.2).float()
    got = warper.forward_warp_matrices(frames, depths, M, src_index=src, mask1=mask1, accumulation=mode)
    ref = oracle.forward_warp(frames, mask1, depths, M, src_index=src.numpy(), per_unit_max=True)
    check(got, ref, f"odd {h}x{w} {mode}")


@pytest.mark.gpu
def test_scatter_destinations_equal_the_direct_call():
    """tcw_forward_warp_scatter (multi-GPU sharding by balanced unit subsets): a subset of units written to its slots of a
    larger output equals the same units of one whole-clip call, bit for bit; the other slots are untouched; a bad
    destination index is flagged, not written out of bounds."""
    import torch
    from trajectorycrafter_b200 import Warper, prepare_matrices
    from trajectorycrafter_b200.sharding import balanced_units, sources_of
    from trajectorycrafter_b200.workloads import orbit_units, synthetic_clip
    nf, h, w = 7, 96, 160
    frames, depths = synthetic_clip(nf, h, w, seed=5)
    T1, T2, K, src = orbit_units(depths, h, w, presets=[(0, 30), (10, -45), (-20, 90)])
    M = prepare_matrices(T1, T2, K)
    n = M.shape[0]
    wp = Warper(device="cuda:0")
    f_d, d_d = frames.cuda(), depths.cuda()
    ref = wp.forward_warp_matrices(f_d, d_d, M, src_index=src.cuda())
    world = 3
    outs = [torch.full((n, c, h, w), 7.0, device="cuda:0") for c in (3, 1, 1, 2)]
    outs[1] = torch.full((n, 1, h, w), 7, dtype=torch.uint8, device="cuda:0")
    for rank in range(world):
        ids = balanced_units(n, nf, world, rank)
        needed, lsrc = sources_of(src, ids)
        # the rank holds only the source frames its units read
        wp.forward_warp_matrices(f_d[needed.cuda()].contiguous(), d_d[needed.cuda()].contiguous(), M[ids], src_index=lsrc.cuda(),
                                 out=tuple(outs), mask_u8=True, dst_index=ids.cuda())
        done = torch.cat([balanced_units(n, nf, world, r) for r in range(rank + 1)]).cuda()
        rest = torch.ones(n, dtype=torch.bool, device="cuda:0")
        rest[done] = False
        assert torch.equal(outs[0][done], ref[0][done]) and torch.equal(outs[1][done].float(), ref[1][done])
        assert torch.equal(outs[2][done], ref[2][done]) and torch.equal(outs[3][done], ref[3][done])
        assert bool((outs[0][rest] == 7.0).all()) and bool((outs[1][rest] == 7).all())
    # float mask, no depth / flow outputs, host index vector
    o2 = (torch.zeros((n, 3, h, w), device="cuda:0"), torch.zeros((n, 1, h, w), device="cuda:0"), None, None)
    ids = balanced_units(n, nf, world, 1)
    wp.forward_warp_matrices(f_d, d_d, M[ids], src_index=src[ids], out=o2, dst_index=ids)
    assert torch.equal(o2[0][ids.cuda()], ref[0][ids.cuda()]) and torch.equal(o2[1][ids.cuda()], ref[1][ids.cuda()])
    # a destination outside the outputs: flagged (IndexError through the status word), nothing written past the tensors
    bad = ids.clone().cuda()
    bad[0] = n + 5
    with pytest.raises(IndexError):
        wp.forward_warp_matrices(f_d, d_d, M[ids], src_index=src[ids].cuda(), out=o2, dst_index=bad)
    with pytest.raises(AssertionError):
        wp.forward_warp_matrices(f_d, d_d, M[ids], src_index=src[ids], out=o2, dst_index=torch.zeros_like(ids))


@pytest.mark.gpu
def test_twice_batched_call(warper):
    """forward_warp(twice=True) with b = 3 (one reference call with a batch: call-wide soft-z normaliser in every one of the
    five splats, models/utils.py:296-347, :492) against the oracle chain."""
    from trajectorycrafter_b200.workloads import synthetic_clip
    h, w = 96, 160
    frames, depths = synthetic_clip(3, h, w, seed=9)
    T1, T2, K, _ = orbit_units(depths, h, w, presets=[(0, 12)])
    got = warper.forward_warp(frames, None, depths, T1, T2, K, None, mask=False, twice=True)
    M = prepare_matrices(T1, T2, K)
    f1, m1, d1, fl, _ = oracle.forward_warp(frames, None, depths, M, per_unit_max=False)
    zt = np.ascontiguousarray(oracle.project(depths, M)[..., 2])
    wfl, _, _ = oracle.splat(fl, np.ones_like(m1), zt, fl, None, is_image=False)
    tf, tm, _ = oracle.splat(f1, m1, np.ascontiguousarray(d1[:, 0]), -wfl, None, is_image=True)
    td, _, _ = oracle.splat(d1, m1, np.ascontiguousarray(d1[:, 0]), -wfl, None, is_image=False)
    gf, gm, gd = npy(got[0]), npy(got[1]), npy(got[2])
    assert gf.shape == (3, 3, h, w) and got[3] is None
    mism = int((gm != tm).sum())
    agree = np.broadcast_to(gm == tm, tf.shape)
    err = np.abs(gf.astype(np.float64) - tf)[agree].max()
    print(f"twice, b = 3 at {h}x{w}: mask mismatches {mism} of {gm.size}, colour max |d| where masks agree {err:.3e}")
    assert mism <= 2 and err <= 2e-5
    valid = (gm > 0) & (tm > 0)
    assert (np.abs(gd - td)[valid] <= 2e-5 * np.abs(td[valid])).all()


@pytest.mark.gpu
def test_infinite_depth_pixels_are_dropped():
    """Non-finite depth (ADVICE, round 1).  A pixel with depth = +inf projects to NaN coordinates: it adds nothing, its neighbours
    are untouched, its flow12 is NaN, and no NaN reaches colour / mask / depth -- the oracle's behaviour, pinned here for the CUDA
    path.  (Documented deviation, DESIGN.md section 1: in the reference such a pixel's z' = inf - inf = NaN poisons
    log_depth.max(), models/utils.py:492, and with it every weight of the call: its whole frame comes out invalid.  The
    reference's own depth stage clips to [1e-4, 1e4], models/infer.py:90, so its pipelines never produce one.)
    A huge finite depth (1e30) behaves like the reference in both."""
    import torch
    from trajectorycrafter_b200 import Warper, prepare_matrices
    from trajectorycrafter_b200.workloads import orbit_units, synthetic_clip
    h, w = 72, 100
    frames, depths = synthetic_clip(2, h, w, seed=2)
    T1, T2, K, src = orbit_units(depths, h, w, presets=[(5, 25), (-10, -60)])
    depths = depths.clone()
    depths[0, 0, 5, 7] = float("inf"); depths[0, 0, 20, 30] = float("inf"); depths[1, 0, 21, 31] = 1e30; depths[1, 0, 60, 90] = float("inf")
    M = prepare_matrices(T1, T2, K)
    wp = Warper(device="cuda:0")
    for mode in ("window", "atomic"):
        got = [npy(t) for t in wp.forward_warp_matrices(frames, depths, M, src_index=src, accumulation=mode)]
        ref = oracle.forward_warp(frames, None, depths, M, src_index=src.numpy(), per_unit_max=True)
        assert not np.isnan(got[0]).any() and not np.isnan(got[1]).any() and not np.isnan(got[2]).any(), mode
        assert np.array_equal(got[1], ref[1]), f"{mode}: validity mask"
        assert np.array_equal(np.isnan(got[3]), np.isnan(ref[3])) and np.isnan(ref[3]).sum() >= 6, f"{mode}: NaN flow exactly where the oracle has it"
        fin = ~np.isnan(ref[3])
        assert np.array_equal(got[3][fin], ref[3][fin]), f"{mode}: flow12 elsewhere bit-exact"
        count = oracle.contrib_count(np.nan_to_num(ref[3], nan=1e6))
        assert_colour_close(got[0], ref[0], count=count, what=f"inf depth {mode}")
        assert_depth_close(got[2], ref[2], ref[1], count=count, what=f"inf depth {mode}")
        assert float(ref[1].mean()) > 0.5
