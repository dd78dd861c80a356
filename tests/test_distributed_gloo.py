"""CPU tier, world_size 2 over gloo: the host side of the multi-GPU path — each rank's shard plan
and the SUM-reduce of per-rank RGBA-sum frames to rank 0 (rustraytrace_b200/distributed.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _frame(camera, spp, sample_range, mask):
    """A deterministic stand-in for one rank's render: pixel value = f(pixel, sample) summed over the
    rank's samples, zero outside its tiles — the same additivity the kernel has."""
    H, W = mask.shape
    y, x = np.mgrid[0:H, 0:W]
    out = np.zeros((H, W, 4), dtype=np.float32)
    for s in range(*sample_range):
        v = ((x * 7 + y * 13 + s * 3) % 11).astype(np.float32)
        out[..., 0] += v; out[..., 1] += 2 * v; out[..., 2] += 1; out[..., 3] += 1
    return out * mask[..., None]


def _worker(rank, world, port, mode, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import rustraytrace_b200 as R
    from rustraytrace_b200 import distributed
    F = R._ffi
    cam = R.make_camera(image_width=80, aspect_ratio=16 / 9, samples_per_pixel=10)
    plan = R.shard_plan(cam, F.INT_BOOK2, distributed.SHARD[mode], rank, world, mask=True)
    mine = torch.from_numpy(_frame(cam, 10, (plan["sample_begin"], plan["sample_end"]), plan["mask"]))
    distributed.reduce_frames(mine, dst=0)
    if rank == 0:
        full = R.shard_plan(cam, F.INT_BOOK2, F.SHARD_NONE, 0, 1, mask=True)
        want = _frame(cam, 10, (0, 10), full["mask"])
        q.put(bool(np.array_equal(mine.numpy(), want)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["sample", "tile"])
def test_two_rank_shares_reduce_to_the_full_frame(mode):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, mode, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def _gather_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import rustraytrace_b200 as R
    from rustraytrace_b200 import distributed
    F = R._ffi
    cam = R.make_camera(image_width=100, aspect_ratio=16 / 9, samples_per_pixel=4)   # 100 x 56: ragged edge tiles
    full = R.shard_plan(cam, F.INT_BOOK2, F.SHARD_NONE, 0, 1, mask=True)
    want = _frame(cam, 4, (0, 4), full["mask"])
    plan = R.shard_plan(cam, F.INT_BOOK2, F.SHARD_TILE, rank, world, mask=True)
    mine = distributed.pack_tiles_host(want * plan["mask"][..., None], rank, world)   # what the kernel's compact mode writes
    assert mine.size == plan["n_super_tiles"] * 1024 * 4
    sizes = [R.shard_plan(cam, F.INT_BOOK2, F.SHARD_TILE, r, world)["n_super_tiles"] * 4096 for r in range(world)]
    pad = torch.zeros(max(sizes)); pad[:mine.size] = torch.from_numpy(mine)
    got = [torch.zeros(max(sizes)) for _ in range(world)] if rank == 0 else None
    dist.gather(pad, got, dst=0)
    if rank == 0:
        frame = R.tile_scatter_host(cam, [g.numpy()[:n] for g, n in zip(got, sizes)])
        q.put(bool(np.array_equal(frame, want)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gather_of_owned_tiles_rebuilds_the_frame(world):
    """Tile sharding as the library exchanges it: every rank sends only its own packed super tiles,
    rank 0 scatters them (rrt_tile_scatter_host = the host mirror of tile_scatter_kernel)."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gather_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def test_combine_host_is_the_same_sum():
    from rustraytrace_b200 import distributed
    a = np.ones((2, 3, 4), dtype=np.float32); b = 2 * a
    assert np.array_equal(distributed.combine_host([a, b]), 3 * np.ones((2, 3, 4)))
