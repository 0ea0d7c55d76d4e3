"""The allocation-free entry point ``tcw_pc_render_zbuffer_static`` (include/tcwarp.h; no counterpart in the reference):
its z-buffer is a file-scope device array inside libtcwarp.so instead of a caller
workspace.  Called through the C-ABI and checked against the existing entry ``tcw_pc_render_zbuffer`` on two calls in a row
(the second one smaller and from another camera, so stale contents of the shared array would show).

Collected last on purpose (``test_zz_``).  Passes alone in a fresh process and after the
other tests (it owns its context and inputs; nothing is carried over but the library's static array, which every call clears)."""
import ctypes

import numpy as np
import pytest
import torch

from trajectorycrafter_b200 import _lib


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _camera(tx, fx, cx, cy):
    T = np.eye(4, dtype=np.float32)
    T[:3, 3] = tx
    K = np.array([[fx, 0, cx], [0, fx, cy], [0, 0, 1]], np.float32)
    cam = np.ascontiguousarray(np.concatenate([T[:3].reshape(-1), K.reshape(-1)]).astype(np.float32))
    return cam, cam.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


@pytest.mark.gpu
def test_static_zbuffer_matches_workspace_entry_twice_and_allocates_nothing():
    assert torch.cuda.is_available(), "this test needs the B200"
    lib = _lib.load()
    dev = torch.device("cuda:0")
    ctx = ctypes.c_void_p()
    _lib.check(lib.tcw_create(0, ctypes.byref(ctx)), "tcw_create")
    try:
        g = torch.Generator().manual_seed(77)
        n = 300_000
        P = (torch.rand((n, 3), generator=g) * torch.tensor([8.0, 5.0, 6.0]) - torch.tensor([4.0, 2.5, 0.0]) + torch.tensor([0.0, 0.0, 4.0])).to(dev)
        C = (torch.rand((n, 3), generator=g) * 2 - 1).to(dev)
        # every device buffer of both calls exists before the first call: nothing below allocates through torch either
        cases = [(576, 1024, 2, _camera([0.1, -0.05, 0.2], 500.0, 512.0, 288.0)),
                 (240, 320, 3, _camera([-0.4, 0.1, -0.3], 260.0, 160.0, 120.0))]
        bufs = []
        for h, w, _, _ in cases:
            mk = lambda c: torch.full((1, c, h, w), 7.0, dtype=torch.float32, device=dev)      # noqa: E731
            bufs.append(([mk(3), mk(1), mk(1)], [mk(3), mk(1), mk(1)], torch.empty(h * w * 8, dtype=torch.uint8, device=dev)))
        stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        torch.cuda.synchronize()
        with torch.cuda.device(dev):
            for i, ((h, w, ps, (keep, cam)), (ref, got, ws)) in enumerate(zip(cases, bufs)):
                _lib.check(lib.tcw_pc_render_zbuffer(ctx, _ptr(P), _ptr(C), cam, n, h, w, ps, _ptr(ref[0]), _ptr(ref[1]), _ptr(ref[2]),
                                                     _ptr(ws), ws.numel(), stream), "tcw_pc_render_zbuffer")
                torch.cuda.synchronize()
                free_before = torch.cuda.mem_get_info(dev)[0]
                _lib.check(lib.tcw_pc_render_zbuffer_static(ctx, _ptr(P), _ptr(C), cam, n, h, w, ps, _ptr(got[0]), _ptr(got[1]), _ptr(got[2]),
                                                            stream), "tcw_pc_render_zbuffer_static")
                torch.cuda.synchronize()
                free_after = torch.cuda.mem_get_info(dev)[0]
                # integer / index work: the same kernels on the same inputs -- bit-equal, the bar of tests/test_pointcloud.py
                for name, a, b in zip(("frame", "mask", "depth"), ref, got):
                    assert np.array_equal(a.cpu().numpy(), b.cpu().numpy()), f"call {i}: {name} differs from tcw_pc_render_zbuffer"
                assert float(ref[1].mean()) > 0.2, "the case must actually render something"
                if i == 1:      # (the first call may load the kernels' module lazily)
                    assert free_after >= free_before, "the static entry point must not allocate device memory"
            # without a depth output, and beyond the stated maximum
            h, w, ps, (keep, cam) = cases[1]
            ref, got, ws = bufs[1]
            got[0].fill_(7.0)
            _lib.check(lib.tcw_pc_render_zbuffer_static(ctx, _ptr(P), _ptr(C), cam, n, h, w, ps, _ptr(got[0]), _ptr(got[1]), None, stream),
                       "tcw_pc_render_zbuffer_static")
            torch.cuda.synchronize()
            assert torch.equal(got[0], ref[0])
            rc = lib.tcw_pc_render_zbuffer_static(ctx, _ptr(P), _ptr(C), cam, n, 1025, 1024, ps, _ptr(got[0]), _ptr(got[1]), None, stream)
            assert rc != 0 and b"TCW_ZB_STATIC_MAX_PIXELS" in lib.tcw_last_error()
            assert 1024 * 1024 == _lib.TCW_ZB_STATIC_MAX_PIXELS
    finally:
        lib.tcw_destroy(ctx)
