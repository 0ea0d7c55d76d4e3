"""``Warper`` — host-side mirror of the reference class (reference models/utils.py:213-682).

Same constructor, method names, positional order, kwargs, return tuples and error behaviour as the
reference; every tensor operation of the hot path is executed by libtcwarp.so (hand-written
sm_100a CUDA behind include/tcwarp.h).  PyTorch is used for device memory, streams and the two
tiny CPU matrix inverses only.  There is no CPU path: ``Warper(device='cpu')`` raises.

Documented deviations from the reference:
  * ``get_device``: the reference maps every string except ``'cpu'``/``'gpuN'`` — including the
    ``'cuda:0'`` all its callers pass — to the CPU (models/utils.py:669-682).  Here ``'gpuN'``,
    ``'cuda'`` and ``'cuda:N'`` all select that GPU and ``'cpu'`` is an error.
  * outputs are contiguous NCHW tensors on the GPU (the reference returns non-contiguous views).
"""
from __future__ import annotations

import ctypes
from pathlib import Path
from typing import Optional, Tuple

import numpy
import torch

from . import _lib
from .matrices import prepare_matrices


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


class Warper:
    def __init__(self, resolution: tuple = None, device: str = 'gpu0'):
        self.resolution = resolution
        self.device = self.get_device(device)
        self.dtype = torch.float32
        self._lib = _lib.load()            # raises if libtcwarp.so has not been built
        self._ctx = None
        self._workspace = None
        return

    # ------------------------------------------------------------------ plumbing
    def _context(self) -> ctypes.c_void_p:
        if self._ctx is None:
            if not torch.cuda.is_available():
                raise _lib.TcwError("no CUDA device is available and this path has no CPU fallback")
            ctx = ctypes.c_void_p()
            idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
            _lib.check(self._lib.tcw_create(idx, ctypes.byref(ctx)), "tcw_create")
            self._ctx = ctx
        return self._ctx

    def __del__(self):
        try:
            if getattr(self, "_ctx", None) is not None:
                self._lib.tcw_destroy(self._ctx)
                self._ctx = None
        except Exception:
            pass

    def _get_workspace(self, n_units: int, h: int, w: int, flags: int) -> torch.Tensor:
        need = ctypes.c_size_t(0)
        _lib.check(self._lib.tcw_workspace_bytes(self._context(), n_units, h, w, flags, ctypes.byref(need)),
                   "tcw_workspace_bytes")
        if self._workspace is None or self._workspace.numel() < need.value:
            self._workspace = torch.empty(need.value, dtype=torch.uint8, device=self.device)
        return self._workspace

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def _dev(self, t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        if t is None:
            return None
        return t.to(self.device).to(self.dtype).contiguous()

    def launch_count(self) -> int:
        """Kernels launched by libtcwarp for this Warper so far (bench.py's gpu_launches)."""
        return int(self._lib.tcw_launch_count(self._context()))

    def profile(self, enable: bool) -> None:
        """Record CUDA events around every kernel launch of this Warper (bench.py's roofline leg)."""
        _lib.check(self._lib.tcw_profile_enable(self._context(), int(enable)), "tcw_profile_enable")

    def profile_read(self):
        """-> {kernel: (total_ms, launches, units)} since the last read; synchronises."""
        ms = (ctypes.c_double * 4)()
        n = (ctypes.c_longlong * 4)()
        units = (ctypes.c_double * 4)()
        _lib.check(self._lib.tcw_profile_read(self._context(), ms, n, units), "tcw_profile_read")
        names = ("prepass", "splat", "normalise", "wmax")
        return {k: (ms[i], int(n[i]), units[i]) for i, k in enumerate(names)}

    # ------------------------------------------------------------------ batched fast entry
    def forward_warp_batch(self, frames: torch.Tensor, depths: torch.Tensor,
                           transformation1: torch.Tensor, transformation2: torch.Tensor,
                           intrinsic1: torch.Tensor, intrinsic2: Optional[torch.Tensor] = None,
                           mask1: Optional[torch.Tensor] = None, src_index: Optional[torch.Tensor] = None,
                           per_frame_max: bool = True, mask: bool = False,
                           return_depth: bool = True, return_flow: bool = True, check_range: bool = True,
                           accumulation: str = "window"):
        """All (source frame, pose) units of a clip in ONE submission.

        Equivalent to looping the reference's ``forward_warp(frames[s:s+1], mask1[s:s+1],
        depths[s:s+1], T1[u:u+1], T2[u:u+1], K1[u:u+1], K2[u:u+1], mask)`` over units u with
        s = src_index[u] (demo.py:77-90, :203-231, :329-342) and concatenating the results — when
        ``per_frame_max`` is True.  With False the soft-z normaliser is taken over the whole call
        like one reference call with b = n_units (models/utils.py:492).
        frames (n_src,3,h,w), depths (n_src,1,h,w), matrices (n_units,...).
        Returns (warped (n,3,h,w), mask (n,1,h,w), depth (n,1,h,w)|None, flow12 (n,2,h,w)|None);
        warped/mask are float64 when ``mask=True`` (reference quirk, models/utils.py:618-626).
        ``accumulation``: "window" (default: every target cell is summed in a fixed order, results are bitwise
        reproducible) or "atomic" (float reductions in L2, last bits vary from run to run).
        """
        n = transformation1.shape[0]
        M = prepare_matrices(transformation1, transformation2, intrinsic1, intrinsic2)
        assert M.shape[0] == n
        return self.forward_warp_matrices(frames, depths, M, mask1=mask1, src_index=src_index,
                                          per_frame_max=per_frame_max, mask=mask, return_depth=return_depth,
                                          return_flow=return_flow, check_range=check_range,
                                          accumulation=accumulation)

    def forward_warp_matrices(self, frames: torch.Tensor, depths: torch.Tensor, M: torch.Tensor,
                              mask1: Optional[torch.Tensor] = None, src_index: Optional[torch.Tensor] = None,
                              per_frame_max: bool = True, mask: bool = False,
                              return_depth: bool = True, return_flow: bool = True, check_range=True,
                              accumulation: str = "window", out=None, mask_u8: bool = False, dst_index=None):
        """``forward_warp_batch`` on an already prepared (n_units,30) matrix block
        (``prepare_matrices``): [K1^-1 | rows 0..2 of T2*T1^-1 | K2] per unit.

        ``check_range`` True: raise AssertionError like the reference if a colour left [-1.1,1.1]
        (one host sync); False: skip; an int32 CUDA tensor: OR the flag into it, no sync (check it
        later).  ``out`` may carry preallocated (frame, mask, depth|None, flow|None) CUDA tensors -- they may live in a peer
        GPU's memory (``sharding.PeerBuffer``).  ``mask_u8``: the validity mask is written as uint8 (default mode only).
        ``dst_index`` (n_units,) int tensor: unit u is written to unit slot ``dst_index[u]`` of the ``out`` tensors, which then hold
        ``out[0].shape[0]`` >= n units (``tcw_forward_warp_scatter``; default accumulation, ``mask=False``, ``out`` required)."""
        n_src, c, h, w = frames.shape
        n = M.shape[0]
        assert c == 3 and depths.shape == (n_src, 1, h, w) and M.shape == (n, 30)
        if mask1 is not None:
            assert mask1.shape == (n_src, 1, h, w)
        if src_index is None:
            assert n == n_src, "src_index is required when the number of poses differs from the number of frames"
        else:
            assert src_index.shape == (n,)
            # a host index vector is checked here; a CUDA one is checked by the kernels that read it (an out-of-range
            # entry reads source 0 and sets TCW_STATUS_SRC_INDEX in the status word) -- no device sync per submission
            if not src_index.is_cuda:
                assert int(src_index.min()) >= 0 and int(src_index.max()) < n_src, "src_index out of range"
        # tensors already resident on the device are used in place (no copy, no sync): a pageable host->device copy
        # per call blocks the host until the previous submission has finished
        frames_d, depths_d, mask1_d = self._dev(frames), self._dev(depths), self._dev(mask1)
        M_d = M.to(self.device, torch.float32, non_blocking=True).contiguous()
        si_d = None if src_index is None else src_index.to(self.device, torch.int32).contiguous()
        mode = {"window": 0, "atomic": _lib.TCW_ATOMIC_F32}[accumulation]
        flags = (_lib.TCW_PER_UNIT_MAX if per_frame_max else 0) | (_lib.TCW_CLEAN if mask else 0) | mode \
            | (_lib.TCW_MASK_U8 if mask_u8 else 0)
        odt = torch.float64 if mask else torch.float32
        mdt = torch.uint8 if mask_u8 else odt
        assert not (mask_u8 and (mask or mode)), "mask_u8 needs the default accumulation and mask=False"
        n_dst = n                      # units the output tensors hold
        di_d = None
        if dst_index is not None:
            assert out is not None and not mask and not mode, "dst_index needs preallocated outputs, the default accumulation and mask=False"
            assert dst_index.shape == (n,)
            n_dst = out[0].shape[0]
            if not dst_index.is_cuda:      # (a CUDA index vector is checked by the kernels: TCW_STATUS_SRC_INDEX)
                assert int(dst_index.min()) >= 0 and int(dst_index.max()) < n_dst and len(set(dst_index.tolist())) == n, \
                    "dst_index must hold distinct unit slots of the outputs"
            di_d = dst_index.to(self.device, torch.int32).contiguous()
        if out is not None:
            out_frame, out_mask, out_depth, out_flow = out
            assert out_frame.shape == (n_dst, 3, h, w) and out_frame.dtype == odt and out_frame.is_contiguous()
            assert out_mask.shape == (n_dst, 1, h, w) and out_mask.dtype == mdt and out_mask.is_contiguous()
            assert out_depth is None or (out_depth.shape == (n_dst, 1, h, w) and out_depth.is_contiguous())
            assert out_flow is None or (out_flow.shape == (n_dst, 2, h, w) and out_flow.is_contiguous())
        else:
            out_frame = torch.empty((n, 3, h, w), dtype=odt, device=self.device)
            out_mask = torch.empty((n, 1, h, w), dtype=mdt, device=self.device)
            out_depth = torch.empty((n, 1, h, w), dtype=torch.float32, device=self.device) if return_depth else None
            out_flow = torch.empty((n, 2, h, w), dtype=torch.float32, device=self.device) if return_flow else None
        lazy_flag = isinstance(check_range, torch.Tensor)
        if lazy_flag:
            flag = check_range
            assert flag.dtype == torch.int32 and flag.is_cuda
        else:
            flag = torch.zeros(1, dtype=torch.int32, device=self.device) if check_range else None
        ws = self._get_workspace(n, h, w, flags)
        with torch.cuda.device(self.device):
            if di_d is None:
                rc = self._lib.tcw_forward_warp(self._context(), _ptr(frames_d), _ptr(mask1_d), _ptr(depths_d),
                                                _ptr(si_d), _ptr(M_d), n, n_src, h, w, flags,
                                                _ptr(out_frame), _ptr(out_mask), _ptr(out_depth), _ptr(out_flow),
                                                _ptr(flag), _ptr(ws), ws.numel(), self._stream())
            else:
                rc = self._lib.tcw_forward_warp_scatter(self._context(), _ptr(frames_d), _ptr(mask1_d), _ptr(depths_d),
                                                        _ptr(si_d), _ptr(M_d), _ptr(di_d), n, n_src, n_dst, h, w, flags,
                                                        _ptr(out_frame), _ptr(out_mask), _ptr(out_depth), _ptr(out_flow),
                                                        _ptr(flag), _ptr(ws), ws.numel(), self._stream())
        _lib.check(rc, "tcw_forward_warp")
        if flag is not None and not lazy_flag:
            self.check_status(int(flag.item()))
        return out_frame, out_mask, out_depth, out_flow

    @staticmethod
    def check_status(status: int) -> None:
        """Raise for the status bits a call ORs into its flag word (include/tcwarp.h TCW_STATUS_*)."""
        if status & _lib.TCW_STATUS_SRC_INDEX:
            raise IndexError("src_index holds a value outside [0, n_src) (or dst_index one outside the output tensors)")
        if status & _lib.TCW_STATUS_BOUNDS:      # only a -DTCW_BOUNDS_CHECK=1 build can raise it
            raise RuntimeError("bounds-checked build: a window / slab access left its allocation")
        # the reference asserts -1.1 <= warped <= 1.1 before clipping (models/utils.py:580-581)
        assert not (status & _lib.TCW_STATUS_RANGE), "warped colour left [-1.1, 1.1]"

    def forward_warp_host(self, frames, depths, M, src_index=None, mask1=None, per_frame_max: bool = True,
                          mask: bool = False, outputs=None, return_depth: bool = True, return_flow: bool = True,
                          mask_u8: bool = False):
        """Host-buffer entry (``tcw_forward_warp_host``): CPU float32 tensors in (ideally pinned),
        CPU tensors out; uploads, kernels and downloads are pipelined inside the C library.
        ``M`` is the prepared (n,30) matrix block.  ``outputs`` may carry preallocated pinned
        (frame, mask, depth, flow) tensors.  ``return_depth`` / ``return_flow`` False skip those downloads
        (every shipped caller drops them: ``warped, mask, _, _``), ``mask_u8`` returns the mask as uint8.
        Returns (frame, mask, depth, flow, status)."""
        n_src, _, h, w = frames.shape
        n = M.shape[0]
        odt = torch.float64 if mask else torch.float32
        assert not (mask and mask_u8)
        if outputs is None:
            outputs = (torch.empty((n, 3, h, w), dtype=odt),
                       torch.empty((n, 1, h, w), dtype=torch.uint8 if mask_u8 else odt),
                       torch.empty((n, 1, h, w), dtype=torch.float32) if return_depth else None,
                       torch.empty((n, 2, h, w), dtype=torch.float32) if return_flow else None)
        of, om, od, ofl = outputs
        assert om.dtype == (torch.uint8 if mask_u8 else odt)
        for t in (frames, depths, M):
            assert t.device.type == "cpu" and t.dtype == torch.float32 and t.is_contiguous()
        si = None if src_index is None else src_index.to(torch.int32).contiguous()
        flags = (_lib.TCW_PER_UNIT_MAX if per_frame_max else 0) | (_lib.TCW_CLEAN if mask else 0) \
            | (_lib.TCW_HOST_MASK_U8 if mask_u8 else 0)
        if mask1 is not None:
            assert mask1.device.type == "cpu" and mask1.dtype == torch.float32 and mask1.is_contiguous()
        flag = ctypes.c_int(0)
        rc = self._lib.tcw_forward_warp_host(self._context(), _ptr(frames), _ptr(mask1), _ptr(depths), _ptr(si),
                                             _ptr(M), n, n_src, h, w, flags, _ptr(of), _ptr(om), _ptr(od), _ptr(ofl),
                                             ctypes.byref(flag))
        _lib.check(rc, "tcw_forward_warp_host")
        return of, om, od, ofl, flag.value

    # ------------------------------------------------------------------ reference API
    def forward_warp(
        self,
        frame1: torch.Tensor,
        mask1: Optional[torch.Tensor],
        depth1: torch.Tensor,
        transformation1: torch.Tensor,
        transformation2: torch.Tensor,
        intrinsic1: torch.Tensor,
        intrinsic2: Optional[torch.Tensor],
        mask=False,
        twice=False,
    ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
        """Drop-in for reference models/utils.py:220-347 (same arguments, shapes and returns).

        frame1 (b,3,h,w) in [-1,1]; mask1 (b,1,h,w)|None; depth1 (b,1,h,w); transformation1/2
        (b,4,4); intrinsic1/2 (b,3,3).  Returns (warped_frame2, mask2, warped_depth2, flow12), or
        the twice-warped triple and ``None`` when ``twice=True``.
        """
        if self.resolution is not None:
            assert frame1.shape[2:4] == self.resolution
        b, c, h, w = frame1.shape
        assert frame1.shape == (b, 3, h, w)
        assert mask1 is None or mask1.shape == (b, 1, h, w)
        assert depth1.shape == (b, 1, h, w)
        assert transformation1.shape == (b, 4, 4)
        assert transformation2.shape == (b, 4, 4)
        assert intrinsic1.shape == (b, 3, 3)
        assert intrinsic2 is None or intrinsic2.shape == (b, 3, 3)
        if mask and b != 1:
            # the reference's clean_points only works for b == 1 (models/utils.py:595-597)
            raise RuntimeError("forward_warp(mask=True) requires batch size 1, as in the reference")

        if not twice:
            # one call with b images: the soft-z max is over the whole call (models/utils.py:492)
            return self.forward_warp_batch(frame1, depth1, transformation1, transformation2, intrinsic1,
                                           intrinsic2, mask1=mask1, per_frame_max=False, mask=mask)

        # twice=True (models/utils.py:296-347): forward splat, then splat back along the warped flow.  The matrices are
        # prepared once, the two range checks (:580-581) share one status word that the host reads once, at the end.
        frame1_d, depth1_d = self._dev(frame1), self._dev(depth1)
        mask1_d = self._dev(mask1) if mask1 is not None else torch.ones((b, 1, h, w), device=self.device)
        M = prepare_matrices(transformation1, transformation2, intrinsic1, intrinsic2)
        status = torch.zeros(1, dtype=torch.int32, device=self.device)
        warped_frame2, mask2, warped_depth2, flow12 = self.forward_warp_matrices(
            frame1_d, depth1_d, M, mask1=mask1, per_frame_max=False, mask=False, check_range=status)
        trans_depth1 = self._project(depth1_d, M)[..., 2].contiguous()
        if mask:
            warped_frame2, mask2 = self.clean_points(warped_frame2, mask2)
            warped_depth2, _ = self.clean_points(warped_depth2, mask2)
        warped_flow, _ = self.bilinear_splatting(flow12, mask1_d, trans_depth1, flow12, None, is_image=False)
        back_flow = -warped_flow
        twice_warped_frame1, twice_warped_mask1 = self.bilinear_splatting(
            warped_frame2, mask2, warped_depth2.squeeze(1), back_flow, None, is_image=True, _status=status)
        twice_warped_depth1, _ = self.bilinear_splatting(
            warped_depth2, mask2, warped_depth2.squeeze(1), back_flow, None, is_image=False)
        if mask:
            twice_warped_frame1, twice_warped_mask1 = self.clean_points(twice_warped_frame1, twice_warped_mask1)
            twice_warped_depth1, _ = self.clean_points(twice_warped_depth1, twice_warped_mask1)
        self.check_status(int(status.item()))
        return twice_warped_frame1, twice_warped_mask1, twice_warped_depth1, None

    def compute_transformed_points(
        self,
        depth1: torch.Tensor,
        transformation1: torch.Tensor,
        transformation2: torch.Tensor,
        intrinsic1: torch.Tensor,
        intrinsic2: Optional[torch.Tensor],
    ):
        """Drop-in for models/utils.py:350-420: (b,1,h,w) depth -> (b,h,w,3,1) projected points
        (all three components 1000 where the transformed point is behind the camera)."""
        if self.resolution is not None:
            assert depth1.shape[2:4] == self.resolution
        M = prepare_matrices(transformation1, transformation2, intrinsic1, intrinsic2)
        return self._project(self._dev(depth1), M).unsqueeze(-1)

    def _project(self, depth_d: torch.Tensor, M: torch.Tensor) -> torch.Tensor:
        """(b,1,h,w) device depth + prepared (b,30) matrices -> (b,h,w,3) projected points."""
        b, _, h, w = depth_d.shape
        M_d = M.to(self.device, torch.float32, non_blocking=True).contiguous()
        out = torch.empty((b, h, w, 3), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            rc = self._lib.tcw_project(self._context(), _ptr(depth_d), None, _ptr(M_d), b, b, h, w,
                                       _ptr(out), self._stream())
        _lib.check(rc, "tcw_project")
        return out

    def bilinear_splatting(
        self,
        frame1: torch.Tensor,
        mask1: Optional[torch.Tensor],
        depth1: torch.Tensor,
        flow12: torch.Tensor,
        flow12_mask: Optional[torch.Tensor],
        is_image: bool = False,
        _status: Optional[torch.Tensor] = None,
    ) -> Tuple[torch.Tensor, torch.Tensor]:
        """Drop-in for models/utils.py:422-583.  frame1 (b,c,h,w), c in 1..4; mask1 (b,1,h,w)|None;
        depth1 (b,h,w) (as the reference actually passes it; (b,1,h,w) is accepted too);
        flow12 (b,2,h,w); flow12_mask (b,1,h,w)|None.  Returns (warped (b,c,h,w), mask (b,1,h,w)).
        ``_status`` (not in the reference): an int32 CUDA tensor the range flag is ORed into instead of being read back here."""
        if self.resolution is not None:
            assert frame1.shape[2:4] == self.resolution
        b, c, h, w = frame1.shape
        assert flow12.shape == (b, 2, h, w)
        assert depth1.numel() == b * h * w
        odt = torch.promote_types(frame1.dtype, torch.float32) if frame1.dtype == torch.float64 else torch.float32
        frame_d, flow_d = self._dev(frame1), self._dev(flow12)
        depth_d = self._dev(depth1.reshape(b, h, w))
        mask1_d, fmask_d = self._dev(mask1), self._dev(flow12_mask)
        flags = _lib.TCW_IS_IMAGE if is_image else 0
        out = torch.empty((b, c, h, w), dtype=torch.float32, device=self.device)
        out_mask = torch.empty((b, 1, h, w), dtype=torch.float32, device=self.device)
        flag = _status if _status is not None else torch.zeros(1, dtype=torch.int32, device=self.device)
        ws = self._get_workspace(b, h, w, flags)
        with torch.cuda.device(self.device):
            rc = self._lib.tcw_splat(self._context(), _ptr(frame_d), _ptr(mask1_d), _ptr(depth_d), _ptr(flow_d),
                                     _ptr(fmask_d), b, c, h, w, flags, _ptr(out), _ptr(out_mask), _ptr(flag),
                                     _ptr(ws), ws.numel(), self._stream())
        _lib.check(rc, "tcw_splat")
        if is_image and _status is None:
            self.check_status(int(flag.item()))       # models/utils.py:580-581
        if odt == torch.float64:               # float64 inputs (after clean_points) stay float64 in the reference
            out, out_mask = out.double(), out_mask.double()
        return out, out_mask

    def clean_points(self, warped_frame2, mask2):
        """Drop-in for models/utils.py:585-626: removes a 2-pixel rim around every hole (5x5 dilation
        of the hole mask).  b must be 1; outputs are float64 like the reference's."""
        if warped_frame2.shape[0] != 1:
            raise RuntimeError("clean_points requires batch size 1, as in the reference")
        _, c, h, w = warped_frame2.shape
        frame_d, mask_d = self._dev(warped_frame2), self._dev(mask2)
        out = torch.empty((1, c, h, w), dtype=torch.float64, device=self.device)
        out_mask = torch.empty((1, 1, h, w), dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            rc = self._lib.tcw_clean_points(self._context(), _ptr(frame_d), _ptr(mask_d), c, h, w,
                                            _ptr(out), _ptr(out_mask), self._stream())
        _lib.check(rc, "tcw_clean_points")
        return out, out_mask

    def pack_condition(self, warped_frames, masks, sample_size):
        """The lines every reference pipeline runs right after its warp loop (demo.py:91-100, 117-120), as one
        kernel: ``cond_video = interpolate((warped + 1) / 2, sample_size, 'bilinear', align_corners=False)`` laid
        out ``(1, 3, F, h, w)`` and ``cond_masks = (1 - interpolate(masks, sample_size, 'nearest')) * 255`` laid out
        ``(1, 1, F, h, w)`` -- the two tensors the diffusion pipeline consumes.

        warped_frames (F,3,H,W) float32 in [-1,1]; masks (F,1,H,W) float32 (``forward_warp_batch`` outputs).
        The reference's ``mask=True`` variant hands float64 tensors to these lines; that is not covered here."""
        assert warped_frames.dim() == 4 and warped_frames.shape[1] == 3, "warped_frames must be (F,3,H,W)"
        n, _, H, W = warped_frames.shape
        assert masks.shape == (n, 1, H, W), "masks must be (F,1,H,W)"
        if warped_frames.dtype != torch.float32 or masks.dtype != torch.float32:
            raise RuntimeError("pack_condition takes float32 tensors (the mask=False path of the reference)")
        h, w = int(sample_size[0]), int(sample_size[1])
        frame_d, mask_d = self._dev(warped_frames), self._dev(masks)
        video = torch.empty((1, 3, n, h, w), dtype=torch.float32, device=self.device)
        cmask = torch.empty((1, 1, n, h, w), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            rc = self._lib.tcw_resize_pack(self._context(), _ptr(frame_d), _ptr(mask_d), n, H, W, h, w,
                                           _ptr(video), _ptr(cmask), self._stream())
        _lib.check(rc, "tcw_resize_pack")
        return video, cmask

    def forward_warp_pack(self, frames, depths, transformation1, transformation2, intrinsic1, intrinsic2=None, sample_size=(384, 672),
                          mask1=None, src_index=None, per_frame_max: bool = True, check_range=True):
        """The reference's warp loop and the lines behind it in ONE submission (demo.py:77-100, 117-120): equal, bit for bit, to
        ``pack_condition(*forward_warp_batch(...)[:2], sample_size)``, without the full-resolution colour / mask / depth / flow ever
        being written out (``tcw_forward_warp_pack``).  Returns (cond_video (1,3,n,h,w), cond_masks (1,1,n,h,w)).
        ``check_range`` as in ``forward_warp_matrices``."""
        n_src, c, H, W = frames.shape
        M = prepare_matrices(transformation1, transformation2, intrinsic1, intrinsic2)
        n = M.shape[0]
        assert c == 3 and depths.shape == (n_src, 1, H, W)
        assert mask1 is None or mask1.shape == (n_src, 1, H, W)
        if src_index is None:
            assert n == n_src, "src_index is required when the number of poses differs from the number of frames"
        elif not src_index.is_cuda:
            assert src_index.shape == (n,) and int(src_index.min()) >= 0 and int(src_index.max()) < n_src, "src_index out of range"
        h, w = int(sample_size[0]), int(sample_size[1])
        frames_d, depths_d, mask1_d = self._dev(frames), self._dev(depths), self._dev(mask1)
        M_d = M.to(self.device, torch.float32, non_blocking=True).contiguous()
        si_d = None if src_index is None else src_index.to(self.device, torch.int32).contiguous()
        flags = _lib.TCW_PER_UNIT_MAX if per_frame_max else 0
        video = torch.empty((1, 3, n, h, w), dtype=torch.float32, device=self.device)
        cmask = torch.empty((1, 1, n, h, w), dtype=torch.float32, device=self.device)
        lazy_flag = isinstance(check_range, torch.Tensor)
        flag = check_range if lazy_flag else (torch.zeros(1, dtype=torch.int32, device=self.device) if check_range else None)
        ws = self._get_workspace(n, H, W, flags | _lib.TCW_PACK)
        with torch.cuda.device(self.device):
            rc = self._lib.tcw_forward_warp_pack(self._context(), _ptr(frames_d), _ptr(mask1_d), _ptr(depths_d), _ptr(si_d), _ptr(M_d),
                                                 n, n_src, H, W, h, w, flags, _ptr(video), _ptr(cmask), _ptr(flag),
                                                 _ptr(ws), ws.numel(), self._stream())
        _lib.check(rc, "tcw_forward_warp_pack")
        if flag is not None and not lazy_flag:
            self.check_status(int(flag.item()))
        return video, cmask

    def depth_postprocess(self, res, near: float = 0.0001, far: float = 10000.0, radius_scale: float = 1.0):
        """The lines that end the reference's depth stage (models/infer.py:79-92) plus the trajectory radius of
        demo.py:539-543, on the device: channel mean, min / max normalisation over the whole clip, ``* 3900``, floor at 1e-5,
        ``10000 / x``, clip -- bit-identical to the numpy / torch lines.

        res (n,h,w,3) (the pipeline's channels-last output) or (n,h,w) float32, on the host or the device.
        Returns (depths (n,1,h,w) CUDA float32, stats): ``stats`` is a 4-float PINNED host tensor (min, max, centre depth,
        radius) filled by an asynchronous copy, and ``stats.ready`` the CUDA event behind it -- ``stats.ready.synchronize()``
        waits for these 16 bytes only, never for the stream (the warp of an earlier clip keeps running).  Pass
        ``centre_depth=float(stats[2])`` to ``poses.get_poses`` instead of letting it read the depth map back (``stats[3]`` is
        the radius itself for callers that build poses elsewhere)."""
        if res.dim() == 3:
            res = res.unsqueeze(-1)
        n, h, w, c = res.shape
        res_d = self._dev(res.contiguous())
        depths = torch.empty((n, 1, h, w), dtype=torch.float32, device=self.device)
        stats_d = torch.empty(4, dtype=torch.float32, device=self.device)
        stats = torch.empty(4, dtype=torch.float32).pin_memory()
        with torch.cuda.device(self.device):
            rc = self._lib.tcw_depth_post(self._context(), _ptr(res_d), n, h, w, c, float(near), float(far), float(radius_scale),
                                          _ptr(depths), _ptr(stats_d), self._stream())
            _lib.check(rc, "tcw_depth_post")
            stats.copy_(stats_d, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
        stats.ready = ev
        return depths, stats

    @staticmethod
    def create_grid(b, h, w):
        """models/utils.py:628-636: (b,2,h,w) int64 pixel grid, channel 0 = x, channel 1 = y."""
        ys, xs = torch.meshgrid(torch.arange(0, h), torch.arange(0, w), indexing="ij")
        return torch.stack([xs, ys], dim=0)[None].repeat([b, 1, 1, 1])

    @staticmethod
    def read_image(path: Path) -> torch.Tensor:
        """models/utils.py:638-641 (skimage.io.imread there; PIL here — same RGB uint8 array)."""
        import PIL.Image
        return numpy.array(PIL.Image.open(path.as_posix()))

    @staticmethod
    def read_depth(path: Path) -> torch.Tensor:
        """models/utils.py:643-654: .png / .npy / .npz('depth')."""
        if path.suffix == '.png':
            import PIL.Image
            depth = numpy.array(PIL.Image.open(path.as_posix()))
        elif path.suffix == '.npy':
            depth = numpy.load(path.as_posix())
        elif path.suffix == '.npz':
            with numpy.load(path.as_posix()) as depth_data:
                depth = depth_data['depth']
        else:
            raise RuntimeError(f'Unknown depth format: {path.suffix}')
        return depth

    @staticmethod
    def camera_intrinsic_transform(capture_width=1920, capture_height=1080, patch_start_point: tuple = (0, 0)):
        """models/utils.py:656-666: 4x4 intrinsics, f=2100, principal point at the (cropped) centre."""
        start_y, start_x = patch_start_point
        k = numpy.eye(4)
        k[0, 0] = k[1, 1] = 2100
        k[0, 2] = capture_width / 2.0 - start_x
        k[1, 2] = capture_height / 2.0 - start_y
        return k

    @staticmethod
    def get_device(device: str):
        """'gpuN' (reference spelling), 'cuda' or 'cuda:N' -> torch.device('cuda', N).

        Deviation from models/utils.py:668-682, which silently falls back to the CPU for anything
        but 'gpuN' (and so runs every shipped caller on the CPU): this implementation has no CPU
        path, so 'cpu' — or a CUDA request without a GPU — raises instead of falling back."""
        if isinstance(device, torch.device):
            device = str(device)
        if device.startswith('gpu'):
            dev = torch.device(f'cuda:{int(device[3:])}')
        elif device.startswith('cuda'):
            dev = torch.device(device if ':' in device else 'cuda:0')
        else:
            raise ValueError(f"Warper device {device!r}: this implementation is CUDA-only (no CPU fallback)")
        return dev
