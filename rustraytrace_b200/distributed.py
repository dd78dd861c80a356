"""Multi-GPU render: one process per GPU, torch.distributed for the plumbing.

The reference has no multi-device code; its only parallel split is rayon over rows
(the_rest_of_your_life/camera.rs:70-72) and, on its GPU path, additive passes into one accum
buffer (src/cuda/mod.rs:403-432, accum.w counts samples).  Both carry over directly:

  * sample-range sharding: rank r renders samples [r*n/N, (r+1)*n/N) of every pixel into a
    full-frame RGBA-sum buffer; frames are SUM-reduced to rank 0 (one NCCL reduce of W*H*4 f32).
  * tile sharding: rank r renders all samples of super tiles k = r (mod N); every other pixel of
    its buffer is zero, so the same SUM-reduce is the gather.

Every (pixel, sample) draws from Philox keyed on (seed; pixel, sample, bounce): the union of the
shares is the unsharded frame up to float summation order.  No collective is needed inside the
path itself; the reduce is the only exchange.
"""
import numpy as np

from . import _ffi as F

SHARD = {"none": F.SHARD_NONE, "tile": F.SHARD_TILE, "sample": F.SHARD_SAMPLE}


def reduce_frames(accum, dst=0, group=None):
    """SUM-reduce per-rank RGBA-sum frames to `dst` (torch tensor, CPU/gloo or CUDA/nccl), in place."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return accum


def render_sharded(ctx, gscene, camera, integrator, accum, seed=1, mode="sample", rank=0, world=1, group=None):
    """Render this rank's share into `accum` (CUDA float32 tensor (H, W, 4) on ctx's device) and
    reduce to rank 0.  Returns accum (complete on rank 0 only).  The library must be launching on
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
