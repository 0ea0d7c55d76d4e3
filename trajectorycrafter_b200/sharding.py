"""Sharding of warp units across ranks and the single exchange step of the path.

Units (source frame, target pose) are independent: each rank warps a contiguous
block of units with no data-path collective.  The one exchange the pipeline needs is assembling
the warped condition video (colour + validity mask) on the denoiser's rank — a gather over
NVLink 5 / NVSwitch via NCCL (``torch.distributed``; gloo in the CPU tests).
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n_units: int, world_size: int, rank: int) -> Tuple[int, int]:
    """Contiguous, balanced block [lo, hi) of units for `rank` (the first n % world ranks get one extra)."""
    assert 0 <= rank < world_size and n_units >= 0
    base, rem = divmod(n_units, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_counts(n_units: int, world_size: int) -> List[int]:
    return [shard_bounds(n_units, world_size, r)[1] - shard_bounds(n_units, world_size, r)[0]
            for r in range(world_size)]


def balanced_units(n_units: int, group_size: int, world_size: int, rank: int, rotate: bool = True) -> torch.Tensor:
    """Unit ids (ascending, int64) of `rank` when every GROUP of `group_size` consecutive units (one trajectory of a clip:
    the units of one camera path cost alike, different paths do not -- a +-90 degree orbit costs 1.8x a small one) is cut
    into `world_size` contiguous frame blocks and rank r takes one block of EVERY group.  With `rotate` the block taken from
    group g is block (r + g) mod world, so the ragged remainders even out (8 groups of 49 on 8 ranks: 49 units each).
    A trailing partial group is cut the same way.  To be used with ``forward_warp_matrices(..., dst_index=ids, out=...)``."""
    assert 0 <= rank < world_size and n_units >= 0 and group_size > 0
    ids = []
    for g, g0 in enumerate(range(0, n_units, group_size)):
        gs = min(group_size, n_units - g0)
        lo, hi = shard_bounds(gs, world_size, (rank + g) % world_size if rotate else rank)
        ids.append(torch.arange(g0 + lo, g0 + hi, dtype=torch.int64))
    return torch.cat(ids) if ids else torch.zeros(0, dtype=torch.int64)


def sources_of(src_index: torch.Tensor, unit_ids: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """``local_sources`` for an arbitrary unit subset: (needed source frames (k,), src_index of the subset remapped onto them)."""
    needed, inverse = torch.unique(src_index[unit_ids].to(torch.int64), sorted=True, return_inverse=True)
    return needed, inverse.to(torch.int32)


def local_sources(src_index: torch.Tensor, lo: int, hi: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """Source frames a rank needs for its unit block, and the block's src_index remapped onto them.

    Returns (needed_sources (k,), local_src_index (hi-lo,)) so a rank uploads only its own frames
    (gradual mode) or the shared frame(s) once (bullet/orbit modes)."""
    block = src_index[lo:hi].to(torch.int64)
    needed, inverse = torch.unique(block, sorted=True, return_inverse=True)
    return needed, inverse.to(torch.int32)


def gather_condition_video(colour: torch.Tensor, mask: torch.Tensor, n_units: int, dst: int = 0,
                           group: Optional[dist.ProcessGroup] = None):
    """Assemble the warped clip on rank `dst`: colour (n_local,3,h,w) and mask (n_local,1,h,w) of every
    rank, concatenated in unit order.  The mask travels as uint8 (lossless: it is 0/1), i.e.
    13 B/px instead of 16 B/px on the wire.  Returns (colour (n,3,h,w), mask (n,1,h,w)) on `dst`,
    (None, None) elsewhere.  Ragged shards (n % world != 0) are padded to the largest shard."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    counts = shard_counts(n_units, world)
    assert colour.shape[0] == counts[rank] == mask.shape[0]
    _, _, h, w = colour.shape
    cmax = max(counts)
    mask_u8 = (mask > 0).to(torch.uint8)
    if counts[rank] < cmax:      # pad so every rank contributes the same number of bytes
        colour = torch.cat([colour, colour.new_zeros((cmax - counts[rank], 3, h, w))])
        mask_u8 = torch.cat([mask_u8, mask_u8.new_zeros((cmax - counts[rank], 1, h, w))])
    colour = colour.contiguous()
    mask_u8 = mask_u8.contiguous()
    if rank == dst:
        col_all = colour.new_empty((world * cmax, 3, h, w))
        msk_all = mask_u8.new_empty((world * cmax, 1, h, w))
        col_list = list(col_all.chunk(world))
        msk_list = list(msk_all.chunk(world))
    else:
        col_all = msk_all = None
        col_list = msk_list = None
    dist.gather(colour, col_list, dst=dst, group=group)
    dist.gather(mask_u8, msk_list, dst=dst, group=group)
    if rank != dst:
        return None, None
    if any(c != cmax for c in counts):
        keep = torch.cat([torch.arange(r * cmax, r * cmax + c) for r, c in enumerate(counts)]).to(col_all.device)
        col_all, msk_all = col_all[keep], msk_all[keep]
    return col_all, msk_all.to(colour.dtype)


# ---------------------------------------------------------------------------------------------
# The fused form of the exchange: no collective at all.  The destination rank owns one buffer laid out as
# [colour (n,3,h,w) float32 | mask (n,1,h,w) uint8] = 13 B per pixel; every other rank maps it (CUDA IPC over
# NVLink / NVSwitch) and its gather kernel stores its units' colour and mask straight into its slice.
# ---------------------------------------------------------------------------------------------
class _DevicePointer:
    """Raw device memory as a __cuda_array_interface__ object (torch.as_tensor wraps it without a copy)."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 3,
                                         "strides": None}


class PeerBuffer:
    """The condition video of one clip on the denoiser's rank, writable by every rank.

    ``PeerBuffer(warper, n_units, h, w, dst, group)`` is collective: rank `dst` allocates (``tcw_ipc_alloc``), the 64-byte
    handle travels once through ``dist.broadcast_object_list`` and the other ranks map it (``tcw_ipc_open``).
    ``views(lo, hi)`` returns (colour (hi-lo,3,h,w) float32, mask (hi-lo,1,h,w) uint8) tensors over units [lo, hi) of the
    buffer -- local memory on `dst`, peer memory elsewhere -- to be passed as ``out=`` to ``forward_warp_matrices(...,
    mask_u8=True)``.  After a barrier ``colour`` / ``mask`` on `dst` hold the assembled clip (what demo.py:117-148 feeds to
    the pipeline, before its resize)."""

    def __init__(self, warper, n_units: int, h: int, w: int, dst: int = 0, group: Optional[dist.ProcessGroup] = None):
        import ctypes
        from . import _lib
        self.warper, self.n, self.h, self.w, self.dst, self.group = warper, n_units, h, w, dst, group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.owner = self.rank == dst
        lib, ctx = warper._lib, warper._context()
        self.colour_bytes = n_units * 3 * h * w * 4
        total = self.colour_bytes + n_units * h * w
        handle = (ctypes.c_ubyte * _lib.TCW_IPC_HANDLE_BYTES)()
        ptr = ctypes.c_void_p()
        if self.owner:
            _lib.check(lib.tcw_ipc_alloc(ctx, total, ctypes.byref(ptr), handle), "tcw_ipc_alloc")
        box = [bytes(handle) if self.owner else None]
        if dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.broadcast_object_list(box, src=dst, group=group)
        if not self.owner:
            buf = (ctypes.c_ubyte * _lib.TCW_IPC_HANDLE_BYTES).from_buffer_copy(box[0])
            _lib.check(lib.tcw_ipc_open(ctx, buf, ctypes.byref(ptr)), "tcw_ipc_open")
        self.ptr = int(ptr.value)
        dev = warper.device
        self.colour = torch.as_tensor(_DevicePointer(self.ptr, (n_units, 3, h, w), "<f4"), device=dev)
        self.mask = torch.as_tensor(_DevicePointer(self.ptr + self.colour_bytes, (n_units, 1, h, w), "|u1"), device=dev)

    def views(self, lo: int, hi: int):
        return self.colour[lo:hi], self.mask[lo:hi]

    def close(self):
        if self.ptr:
            lib, ctx = self.warper._lib, self.warper._context()
            self.colour = self.mask = None
            (lib.tcw_ipc_free if self.owner else lib.tcw_ipc_close)(ctx, self.ptr)
            self.ptr = 0


def gather_packed(packed: torch.Tensor, n_units: int, h: int, w: int, dst: int = 0, group: Optional[dist.ProcessGroup] = None):
    """The collective form with ONE gather: every rank's units as one uint8 buffer [colour float32 | mask uint8]
    (13 B per pixel, written by the warp itself through ``packed_views``), equal shards.  Returns the (world, bytes) uint8
    tensor on `dst` (None elsewhere); ``unpack_gathered`` gives the views."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    assert n_units % world == 0, "gather_packed needs equal shards (pad the unit list)"
    out = packed.new_empty((world, packed.numel())) if rank == dst else None
    dist.gather(packed, list(out.unbind(0)) if rank == dst else None, dst=dst, group=group)
    return out


def packed_buffer(n_local: int, h: int, w: int, device) -> torch.Tensor:
    return torch.empty(n_local * h * w * 13, dtype=torch.uint8, device=device)


def packed_views(packed: torch.Tensor, n_local: int, h: int, w: int):
    """(colour (n,3,h,w) float32, mask (n,1,h,w) uint8) views of one rank's packed buffer."""
    cb = n_local * 3 * h * w * 4
    return packed[:cb].view(torch.float32).view(n_local, 3, h, w), packed[cb:].view(n_local, 1, h, w)


def unpack_gathered(gathered: torch.Tensor, n_units: int, h: int, w: int):
    """Per-rank (colour, mask) views of the gathered (world, bytes) buffer, in unit order."""
    world = gathered.shape[0]
    return [packed_views(gathered[r], n_units // world, h, w) for r in range(world)]
