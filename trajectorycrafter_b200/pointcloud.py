"""Host-side mirror of the fork's ``GlobalPointCloudWarper`` (reference notebooks/06_10_25_vggt/warper_point_cloud.py) over
libtcwarp.so.  Same method names, argument order and return shapes; tensors live on the CUDA device.
No CPU path: a missing library raises (``_lib.load``).

Built: every method of the class -- ``lift_to_3d_pointcloud`` / ``create_pointcloud_from_image`` / ``merge_pointclouds`` /
``downsample_pointcloud``, ``render_from_camera`` (bilinear and disc splats) and the three z-buffer renderers.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Tuple

import torch

from . import _lib
from .warper import Warper, _ptr


class GlobalPointCloudWarper:
    def __init__(self, resolution: tuple = None, device: str = 'cuda', max_points: int = 1000000):
        self.max_points = max_points
        self._w = Warper(resolution=None, device=device)
        self.device = self._w.device
        self.dtype = torch.float32
        self.resolution = resolution
        self._ws = None

    # ---- plumbing
    def _workspace(self, nbytes: int) -> torch.Tensor:
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        return self._ws

    def _dev(self, t):
        return None if t is None else t.to(self.device, self.dtype).contiguous()

    @staticmethod
    def _camera(transformation: torch.Tensor, intrinsic: torch.Tensor):
        """21 host floats: rows 0..2 of the (4,4) extrinsic, then the (3,3) intrinsic."""
        cam = torch.cat([transformation.detach().to("cpu", torch.float32)[:3].reshape(-1),
                         intrinsic.detach().to("cpu", torch.float32).reshape(-1)]).contiguous()
        return cam, cam.numpy().ctypes.data_as(ctypes.POINTER(ctypes.c_float))

    # ---- reference :12-77
    def lift_to_3d_pointcloud(self, frame1: torch.Tensor, mask1: Optional[torch.Tensor], depth1: torch.Tensor,
                              transformation1: torch.Tensor, intrinsic1: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """Lift 2D image points to 3D world coordinates with colours: (b,h,w,3), (b,h,w,3), both multiplied by mask1.
        The two small inverses are taken on the host with the reference's own torch CPU calls (DESIGN.md "H2")."""
        if self.resolution is not None:
            assert frame1.shape[2:4] == self.resolution
        b, c, h, w = frame1.shape
        assert c == 3 and depth1.shape == (b, 1, h, w)
        kinv = torch.linalg.inv(intrinsic1.detach().to("cpu", torch.float32))
        tinv = torch.linalg.inv(transformation1.detach().to("cpu", torch.float32))
        M = torch.cat([kinv.reshape(b, 9), tinv[:, :3].reshape(b, 12)], dim=1).contiguous().to(self.device)
        frame_d, mask_d, depth_d = self._dev(frame1), self._dev(mask1), self._dev(depth1)
        points = torch.empty((b, h, w, 3), dtype=torch.float32, device=self.device)
        colors = torch.empty((b, h, w, 3), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            rc = self._w._lib.tcw_pc_lift(self._w._context(), _ptr(frame_d), _ptr(mask_d), _ptr(depth_d), _ptr(M), b, h, w, 1,
                                          _ptr(points), _ptr(colors), self._w._stream())
        _lib.check(rc, "tcw_pc_lift")
        return points, colors

    def extract_3d_points_with_colors(self, frame1, depth1, transformation1, intrinsic1, subsample_step: int = 10):
        """VisualizationWarper.extract_3d_points_with_colors (reference notebooks/28_08_25_trajectories/core.py:15-79): every
        `subsample_step`-th pixel lifted with ``transformation1`` itself (not its inverse), points with depth > 0 kept.
        Returns (points (N,3), colours (N,3))."""
        b = frame1.shape[0]
        st = int(subsample_step)
        depth_sub = self._dev(depth1)[:, :, ::st, ::st].contiguous()
        colors_sub = self._dev(frame1)[:, :, ::st, ::st].contiguous()
        hs, ws_ = depth_sub.shape[-2:]
        kinv = torch.linalg.inv(intrinsic1.detach().to("cpu", torch.float32))
        t = transformation1.detach().to("cpu", torch.float32)
        M = torch.cat([kinv.reshape(b, 9), t[:, :3].reshape(b, 12)], dim=1).contiguous().to(self.device)
        points = torch.empty((b, hs, ws_, 3), dtype=torch.float32, device=self.device)
        colors = torch.empty((b, hs, ws_, 3), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            rc = self._w._lib.tcw_pc_lift(self._w._context(), _ptr(colors_sub), None, _ptr(depth_sub), _ptr(M), b, hs, ws_, st,
                                          _ptr(points), _ptr(colors), self._w._stream())
        _lib.check(rc, "tcw_pc_lift")
        valid = depth_sub[:, 0] > 0
        return points[valid], colors[valid]

    # ---- reference :79-120
    def create_pointcloud_from_image(self, frame, mask, depth, transformation, intrinsic, confidence_weight: float = 1.0):
        world_points, colors = self.lift_to_3d_pointcloud(frame, mask, depth, transformation, intrinsic)
        b, h, w, _ = world_points.shape
        if mask is None:
            mask = torch.ones(b, 1, h, w, device=self.device, dtype=self.dtype)
        valid_mask = self._dev(mask)[:, 0, :, :].bool()
        points = world_points[valid_mask]
        point_colors = colors[valid_mask]
        weights = torch.full((points.shape[0], 1), confidence_weight, device=self.device, dtype=self.dtype)
        return points, point_colors, weights

    # ---- reference :122-136
    def merge_pointclouds(self, point_clouds: List[torch.Tensor], colors_list: List[torch.Tensor], weights_list: List[torch.Tensor]):
        if not point_clouds:
            return None, None, None
        return torch.cat(point_clouds, dim=0), torch.cat(colors_list, dim=0), torch.cat(weights_list, dim=0)

    # ---- reference :138-156 (random: confidence-weighted sampling without replacement, torch's generator)
    def downsample_pointcloud(self, points, colors, weights, max_points: Optional[int] = None):
        if max_points is None:
            max_points = self.max_points
        if points.shape[0] <= max_points:
            return points, colors, weights
        probabilities = (weights / weights.sum()).squeeze()
        indices = torch.multinomial(probabilities, max_points, replacement=False)
        return points[indices], colors[indices], weights[indices]

    # ---- reference :158-256
    def render_from_camera(self, points, colors, weights, transformation, intrinsic, target_height: int, target_width: int,
                           depth_threshold: float = 100000.0, point_size: float = 2.0, use_bilinear: bool = True):
        """(b,3,h,w) image and (b,1,h,w) mask of the point cloud seen from each of the b cameras (depth-weighted bilinear
        splat; pixels nothing reached stay 0)."""
        b = transformation.shape[0]
        h, w = int(target_height), int(target_width)
        if points is None or points.shape[0] == 0:
            return (torch.zeros(b, 3, h, w, device=self.device), torch.zeros(b, 1, h, w, device=self.device))
        pts, col, wts = self._dev(points), self._dev(colors), self._dev(weights).reshape(-1)
        out = torch.empty((b, 3, h, w), dtype=torch.float32, device=self.device)
        msk = torch.empty((b, 1, h, w), dtype=torch.float32, device=self.device)
        ws = self._workspace(256 + h * w * 16)
        if not use_bilinear:
            offs, fall = self._disc(point_size)
        for i in range(b):
            keep, cam = self._camera(transformation[i], intrinsic[i])
            with torch.cuda.device(self.device):
                if use_bilinear:
                    rc = self._w._lib.tcw_pc_render_bilinear(self._w._context(), _ptr(pts), _ptr(col), _ptr(wts), cam, pts.shape[0],
                                                             h, w, float(depth_threshold), _ptr(out[i]), _ptr(msk[i]), _ptr(ws),
                                                             ws.numel(), self._w._stream())
                else:
                    rc = self._w._lib.tcw_pc_render_disc(self._w._context(), _ptr(pts), _ptr(col), _ptr(wts), cam, pts.shape[0], h, w,
                                                         float(depth_threshold),
                                                         offs.numpy().ctypes.data_as(ctypes.POINTER(ctypes.c_int)),
                                                         fall.numpy().ctypes.data_as(ctypes.POINTER(ctypes.c_float)), fall.numel(),
                                                         _ptr(out[i]), _ptr(msk[i]), _ptr(ws), ws.numel(), self._w._stream())
            _lib.check(rc, "tcw_pc_render_bilinear" if use_bilinear else "tcw_pc_render_disc")
        return out, msk

    @staticmethod
    def _disc(point_size: float):
        """The stamp of ``_splat_points_to_image_with_size`` (reference :281-296): integer offsets (dx, dy) within
        point_size / 2 of the centre, rows first, and their Gaussian falloff exp(-d^2 / (2 (point_size / 4)^2)) -- the exponent
        in Python doubles, the exp in float32, as the reference evaluates them."""
        radius = int(point_size / 2)
        offs, fall = [], []
        for dy in range(-radius, radius + 1):
            for dx in range(-radius, radius + 1):
                dist = (dx ** 2 + dy ** 2) ** 0.5
                if dist <= point_size / 2:
                    offs.append((dx, dy))
                    fall.append(-dist ** 2 / (2 * (point_size / 4) ** 2))
        return (torch.tensor(offs, dtype=torch.int32).reshape(-1, 2).contiguous(),
                torch.exp(torch.tensor(fall, dtype=torch.float32)).contiguous())

    def _zbuffer(self, points_3d, colors_3d, transformation_target, intrinsic_target, image_size, point_size, return_depth,
                 first_wins: bool = False):
        h, w = int(image_size[0]), int(image_size[1])
        pts, col = self._dev(points_3d), self._dev(colors_3d)
        frame = torch.empty((1, 3, h, w), dtype=torch.float32, device=self.device)
        mask = torch.empty((1, 1, h, w), dtype=torch.float32, device=self.device)
        depth = torch.empty((1, 1, h, w), dtype=torch.float32, device=self.device) if return_depth else None
        ws = self._workspace(h * w * 8)
        keep, cam = self._camera(transformation_target[0], intrinsic_target[0])
        with torch.cuda.device(self.device):
            entry = self._w._lib.tcw_pc_render_zbuffer_first if first_wins else self._w._lib.tcw_pc_render_zbuffer
            rc = entry(self._w._context(), _ptr(pts), _ptr(col), cam, pts.shape[0], h, w, int(point_size),
                       _ptr(frame), _ptr(mask), _ptr(depth), _ptr(ws), ws.numel(), self._w._stream())
        _lib.check(rc, "tcw_pc_render_zbuffer")
        return (frame, mask, depth) if return_depth else (frame, mask)

    # ---- reference :445-519: one pixel per point, the rounded pixel clamped into the image
    def render_pointcloud_zbuffer_vectorized(self, points_3d, colors_3d, transformation_target, intrinsic_target,
                                             image_size: tuple = (576, 1024)):
        return self._zbuffer(points_3d, colors_3d, transformation_target, intrinsic_target, image_size, 0, False)

    # ---- reference :522-646: point_size x point_size footprint, nearest point per pixel, holes -1
    def render_pointcloud_zbuffer_vectorized_point_size(self, points_3d, colors_3d, transformation_target, intrinsic_target,
                                                        image_size: tuple = (576, 1024), point_size: int = 1,
                                                        return_depth: bool = False):
        assert int(point_size) >= 1
        return self._zbuffer(points_3d, colors_3d, transformation_target, intrinsic_target, image_size, int(point_size), return_depth)

    # ---- reference :650-765: the sequential loop with a strict `<` (the first of equal depths stays); point_size <= 1 clamps
    def render_pointcloud_zbuffer_vectorized_fixed(self, points_3d, colors_3d, transformation_target, intrinsic_target,
                                                   image_size: tuple = (576, 1024), point_size: int = 1):
        return self._zbuffer(points_3d, colors_3d, transformation_target, intrinsic_target, image_size, int(point_size), False,
                             first_wins=True)
