"""rrt_probe_static (include/rrt.h, csrc/noalloc.cu): the probe with its device scratch in a static device
array instead of cudaMalloc'ed buffers.  No counterpart in the reference.
Through the C ABI, against rrt_probe on the same scene and camera: ids and t identical
(probe-against-probe comparisons in this suite are exact: same kernel, same inputs), on two calls in a row
that use different frame sizes, so stale contents of the shared scratch would show.

Collected last on purpose."""
import numpy as np
import pytest

import rustraytrace_b200 as R

F = R._ffi
pytestmark = pytest.mark.gpu


def test_probe_static_matches_probe_on_two_calls_in_a_row():
    ctx = R.Context(0)
    big = R.BookScene("in_one_weekend", None, overrides=dict(image_width=640))
    small = R.BookScene("the_next_week", 9, overrides=dict(image_width=200))
    g_big, g_small = ctx.scene(big.desc), ctx.scene(small.desc)
    for precision in (64, 32):
        # first call fills 640x360 pixels of the scratch, the second (another scene, 200x200) reuses its start
        for s, g in ((big, g_big), (small, g_small), (big, g_big)):
            ids, t = ctx.probe(g, s.camera, precision)
            sids, st = ctx.probe_static(g, s.camera, precision)
            assert ctx.stats()["kernel_launches"] == 1
            assert (ids >= 0).any()
            assert np.array_equal(ids, sids), f"precision {precision}: {(ids != sids).sum()} ids differ"
            assert np.array_equal(t, st), f"precision {precision}: max |dt| {np.abs(t - st).max()}"
    g_big.close(); g_small.close(); ctx.close()


def test_probe_static_refuses_frames_beyond_its_scratch():
    ctx = R.Context(0)
    s = R.BookScene("the_next_week", 2, overrides=dict(image_width=64))
    g = ctx.scene(s.desc)
    cam = R.copy_camera(s.camera, image_width=2048, aspect_ratio=1.0)   # 4 Mi pixels > 1 Mi
    with pytest.raises(R.RrtError) as e:
        ctx.probe_static(g, cam, 64)
    assert e.value.status == F.ERR_UNSUPPORTED and "RRT_PROBE_STATIC_MAX_PIXELS" in str(e.value)
    cam = R.copy_camera(s.camera, image_width=1024, aspect_ratio=1.0)   # exactly the maximum
    ids, t = ctx.probe(g, cam, 64)
    sids, st = ctx.probe_static(g, cam, 64)
    assert np.array_equal(ids, sids) and np.array_equal(t, st)
    g.close(); ctx.close()
