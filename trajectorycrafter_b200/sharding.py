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
