"""Glue for one-process-per-GPU launches (torchrun): torch.distributed carries the 128-byte NCCL id from
rank 0 to the other ranks, everything after that — this rank's share of the frame and the exchange of
the shares over NVLink — happens inside librrt.so (rrt_comm_init + rrt_render_collective,
rustraytrace_b200/csrc/multi.cu).

The reference has no multi-device code; its only parallel split is rayon over rows
(the_rest_of_your_life/camera.rs:70-72) and, on its GPU path, additive passes into one accum buffer
(src/cuda/mod.rs:403-432, accum.w counts samples).  Both carry over:
  * sample-range sharding: rank r renders samples [r*n/N, (r+1)*n/N) of every pixel; the RGBA-sum frames
    are SUM-reduced to rank 0 (one ncclReduce of W*H*4 f32);
  * tile sharding: rank r renders all samples of super tiles k = r (mod N), packed back to back; rank 0
    gathers every rank's OWN tiles (1/N of the frame each) and scatters them into the frame.
Every (pixel, sample) draws from Philox keyed on (seed; pixel, sample, bounce): the union of the shares
is the unsharded frame up to float summation order.

reduce_frames / render_sharded are the round-1 path (per-rank rrt_render_device + torch.distributed
reduce of zero-padded frames); they stay as a TEST HARNESS: the gloo tests run them on CPU tensors, and
the GPU tests hold the in-library exchange against them.
"""
import numpy as np

from . import _ffi as F

SHARD = {"none": F.SHARD_NONE, "tile": F.SHARD_TILE, "sample": F.SHARD_SAMPLE}


def init_comm_from_torch(ctx, group=None):
    """Create librrt's NCCL communicator across the ranks of an initialised torch.distributed group:
    rank 0 makes the id (rrt_comm_unique_id), torch broadcasts it, every rank joins (rrt_comm_init)."""
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    box = [ctx.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group)
    ctx.comm_init(box[0], rank, world)
    return rank, world


def pack_tiles_host(frame, rank, world):
    """Host mirror of the kernel's compact tile layout (tests): the super tiles of `frame` (H, W, 4) that
    belong to `rank` — tile k with k % world == rank, where row sy of the frame is numbered from column sy on
    (device_types.h super_tile_xy: sy = k // sx, column (k % sx + sy) % sx) — back to back,
    1024 RGBA pixels each, row-major inside a tile, zero where a tile hangs over the frame's edge."""
    H, W = frame.shape[:2]
    sx, sy = (W + 31) // 32, (H + 31) // 32
    ks = list(range(rank, sx * sy, world))
    out = np.zeros((len(ks), 32, 32, 4), dtype=np.float32)
    for j, k in enumerate(ks):
        ty = k // sx
        y0, x0 = ty * 32, ((k % sx + ty) % sx) * 32
        blk = frame[y0:y0 + 32, x0:x0 + 32]
        out[j, :blk.shape[0], :blk.shape[1]] = blk
    return out.reshape(-1)


def reduce_frames(accum, dst=0, group=None):
    """SUM-reduce per-rank RGBA-sum frames to `dst` (torch tensor, CPU/gloo or CUDA/nccl), in place."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return accum


def render_sharded(ctx, gscene, camera, integrator, accum, seed=1, mode="sample", rank=0, world=1, group=None):
    """Test harness: render this rank's share into `accum` (CUDA float32 tensor (H, W, 4) on ctx's device)
    with rrt_render_device and reduce to rank 0 with torch.distributed.  The library must be launching on
    the current torch stream (ctx.set_stream) so the reduce is ordered after the kernel."""
    m = SHARD[mode] if isinstance(mode, str) else int(mode)
    if world == 1:
        m = F.SHARD_NONE
    ctx.render_device(gscene, camera, integrator, accum.data_ptr(), seed=seed, shard_mode=m, rank=rank, world=world)
    return reduce_frames(accum, 0, group)


def combine_host(frames):
    """Host-side statement of the reduce (tests): float64 sum of per-rank (H, W, 4) frames."""
    tot = np.zeros_like(np.asarray(frames[0]), dtype=np.float64)
    for f in frames:
        tot += np.asarray(f, dtype=np.float64)
    return tot
