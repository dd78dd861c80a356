"""GPU tier: several GPUs behind the C ABI (include/rrt.h "several GPUs of one box", csrc/multi.cu).

On the one-GPU test box the exchange code runs with a world of one — a group over a single device, a
communicator of one rank: compact tile packing + scatter kernel, in-place reduce, stats — and is held
against the plain single-device render.  With two or more devices visible (gpurun --gpus 2) the same
tests run over two GPUs in both forms: one process driving a group (rrt_init_multi, no torch anywhere),
and one process per GPU with the NCCL id handed over a multiprocessing pipe (rrt_comm_init +
rrt_render_collective)."""
import multiprocessing as mp

import numpy as np
import pytest

import rustraytrace_b200 as R

F = R._ffi
pytestmark = pytest.mark.gpu


def n_gpus():
    import torch
    return torch.cuda.device_count()


def scene(width=200, spp=36, book="the_next_week", no=1):
    return R.BookScene(book, no, overrides=dict(image_width=width, samples_per_pixel=spp, max_depth=12))


def single(s, seed):
    c = R.Context(0)
    gs = c.scene(s.desc)
    a = c.render(gs, s.camera, s.integrator, seed=seed)
    gs.close(); c.close()
    return a


@pytest.mark.parametrize("world", [1, 2])
@pytest.mark.parametrize("mode", [F.SHARD_SAMPLE, F.SHARD_TILE])
@pytest.mark.parametrize("book,no", [("the_next_week", 1), ("the_rest_of_your_life", None)])
def test_group_context_returns_the_unsharded_frame(world, mode, book, no):
    if world > n_gpus():
        pytest.skip(f"needs {world} GPUs")
    s = scene(book=book, no=no)                      # 200 x 112 / 200 x 200: ragged 32x32 super tiles at the edges
    want = single(s, 5).astype(np.float64)
    g = R.Context(devices=list(range(world)))
    gs = g.scene(s.desc)
    got = g.render(gs, s.camera, s.integrator, seed=5, shard_mode=mode).astype(np.float64)
    st = g.stats()
    assert st["n_devices"] == world and st["paths"] == s.width * s.height * st["spp_effective"]
    assert st["kernel_launches"] == world + (1 if mode == F.SHARD_TILE else 0)
    assert st["exchange_ms"] >= 0 and st["d2h_bytes"] == got.size * 4
    assert np.array_equal(got[..., 3], want[..., 3])
    assert np.allclose(got[..., :3], want[..., :3], rtol=2e-4, atol=1e-3)
    # one-shot form and the device-output form
    again = g.render_frame(s.desc, s.camera, s.integrator, seed=5).astype(np.float64)   # RRT_SHARD_NONE -> sample ranges
    assert np.array_equal(again[..., 3], want[..., 3]) and np.allclose(again[..., :3], want[..., :3], rtol=2e-4, atol=1e-3)
    ids, t = g.probe(gs, s.camera, 64)
    assert ids.shape == (s.height, s.width)
    assert g.resolve(got.astype(np.float32), st["spp_effective"]).shape == (s.height, s.width, 3)
    gs.close(); g.close()


def test_group_context_rejects_bad_device_lists():
    for devs in ([], [0, 0], [99], list(range(65))):
        with pytest.raises(R.RrtError):
            R.Context(devices=devs)
    g = R.Context(devices=[0])
    with pytest.raises(R.RrtError):
        g.set_stream(0x1234)
    g.close()


def _rank(rank, world, conn, mode, seed, q, one_shot=False):
    import numpy as np
    import rustraytrace_b200 as R
    s = scene()
    c = R.Context(rank)
    if rank == 0:
        uid = R.Context.comm_unique_id()
        for k in conn:
            k.send(uid)
    else:
        uid = conn.recv()
    c.comm_init(uid, rank, world)
    out = np.zeros((s.height, s.width, 4), dtype=np.float32) if rank == 0 else None
    if one_shot:   # rrt_render_frame_collective: upload, share, exchange, free in one call
        c.render_frame_collective(s.desc, s.camera, s.integrator, seed=seed, shard_mode=mode, out=out)
        st = c.stats()
    else:
        gs = c.scene(s.desc)
        c.render_collective(gs, s.camera, s.integrator, seed=seed, shard_mode=mode, out=out)
        st = c.stats()
        gs.close()
    if rank == 0:
        q.put((out, st["n_devices"], st["exchange_ms"]))
    c.close()


@pytest.mark.parametrize("world", [1, 2])
@pytest.mark.parametrize("mode,one_shot", [(F.SHARD_SAMPLE, False), (F.SHARD_TILE, False), (F.SHARD_SAMPLE, True), (F.SHARD_TILE, True)])
def test_one_process_per_gpu_collective_render(world, mode, one_shot):
    if world > n_gpus():
        pytest.skip(f"needs {world} GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    pipes = [ctx.Pipe() for _ in range(world - 1)]
    procs = [ctx.Process(target=_rank, args=(0, world, [p[0] for p in pipes], mode, 9, q, one_shot))]
    procs += [ctx.Process(target=_rank, args=(r, world, pipes[r - 1][1], mode, 9, q, one_shot)) for r in range(1, world)]
    for p in procs:
        p.start()
    out, ndev, xms = q.get(timeout=180)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    want = single(scene(), 9).astype(np.float64)
    assert ndev == world and xms >= 0
    assert np.array_equal(out[..., 3].astype(np.float64), want[..., 3])
    assert np.allclose(out[..., :3], want[..., :3], rtol=2e-4, atol=1e-3)


def test_collective_render_needs_a_communicator():
    s = scene(width=64, spp=4)
    c = R.Context(0)
    gs = c.scene(s.desc)
    with pytest.raises(R.RrtError) as e:
        c.render_collective(gs, s.camera, s.integrator)
    assert "rrt_comm_init" in str(e.value)
    gs.close(); c.close()
