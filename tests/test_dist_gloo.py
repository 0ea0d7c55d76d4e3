"""CPU, world_size 2 over gloo: unit sharding + the gather of the warped condition video."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from trajectorycrafter_b200.sharding import gather_condition_video, shard_bounds


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_units, results):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        h, w = 6, 10
        g = torch.Generator().manual_seed(11)
        colour_all = torch.rand((n_units, 3, h, w), generator=g) * 2 - 1
        mask_all = (torch.rand((n_units, 1, h, w), generator=g) > 0.4).float()
        lo, hi = shard_bounds(n_units, world, rank)
        col, msk = gather_condition_video(colour_all[lo:hi].clone(), mask_all[lo:hi].clone(), n_units, dst=0)
        if rank == 0:
            ok = torch.equal(col, colour_all) and torch.equal(msk, mask_all)
            results.put(("ok" if ok else "mismatch", tuple(col.shape)))
        else:
            assert col is None and msk is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def _run(n_units):
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_units, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    status, shape = q.get()
    assert status == "ok" and shape[0] == n_units


def test_gather_even_shards():
    _run(8)


def test_gather_ragged_shards():
    _run(7)


def _worker_packed(rank, world, port, n_units, results):
    from trajectorycrafter_b200.sharding import gather_packed, packed_buffer, packed_views, unpack_gathered
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        h, w = 6, 10
        g = torch.Generator().manual_seed(5)
        colour_all = torch.rand((n_units, 3, h, w), generator=g) * 2 - 1
        mask_all = (torch.rand((n_units, 1, h, w), generator=g) > 0.4).to(torch.uint8)
        lo, hi = shard_bounds(n_units, world, rank)
        packed = packed_buffer(hi - lo, h, w, "cpu")
        col, msk = packed_views(packed, hi - lo, h, w)
        col.copy_(colour_all[lo:hi])            # on the GPU the warp writes these views itself (out=, mask_u8=True)
        msk.copy_(mask_all[lo:hi])
        got = gather_packed(packed, n_units, h, w, dst=0)
        if rank == 0:
            parts = unpack_gathered(got, n_units, h, w)
            ok = torch.equal(torch.cat([p[0] for p in parts]), colour_all) and torch.equal(torch.cat([p[1] for p in parts]), mask_all)
            results.put(("ok" if ok else "mismatch", (n_units,)))
        else:
            assert got is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_single_gather_of_packed_colour_and_u8_mask():
    """The collective form of the exchange is ONE gather of 13 B per pixel (colour float32 + mask uint8 in one buffer)."""
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_packed, args=(r, 2, port, 8, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    status, shape = q.get()
    assert status == "ok"


def test_balanced_units_partition_the_clip():
    """Every unit belongs to exactly one rank; with rotation 8 groups of 49 give 49 units to each of 8 ranks, and every rank
    holds a block of every group."""
    import torch
    from trajectorycrafter_b200.sharding import balanced_units, sources_of
    for n, gs, world in ((392, 49, 8), (392, 49, 4), (392, 49, 2), (100, 49, 3), (7, 49, 8), (49, 49, 1)):
        parts = [balanced_units(n, gs, world, r) for r in range(world)]
        allu = torch.cat(parts)
        assert allu.numel() == n and torch.equal(torch.sort(allu).values, torch.arange(n))
        for p in parts:
            assert torch.equal(p, torch.sort(p).values)
    parts = [balanced_units(392, 49, 8, r) for r in range(8)]
    assert all(p.numel() == 49 for p in parts)
    assert all(len(set((p // 49).tolist())) == 8 for p in parts)
    sizes = [balanced_units(392, 49, 8, r, rotate=False).numel() for r in range(8)]
    assert max(sizes) == 56 and min(sizes) == 48      # why the blocks rotate
    src = torch.arange(392) % 49
    needed, local = sources_of(src, parts[3])
    assert torch.equal(needed[local.long()], src[parts[3]])
