"""Axis-aligned rectangles of a brute-force root leaf (flatten.cpp rect_axis, intersect.cuh rect_quads_hit32).

The walls, floor, ceiling and light of the Cornell boxes (the_rest_of_your_life/mod.rs:82-130,
the_next_week/mod.rs:236-295) are quads whose edges run along coordinate axes.  Quad::hit (quad.rs:64-91) for
them is t = (plane - o_a) / d_a plus two interval tests; the flattener puts them first in the leaf, grouped by the
axis of their normal.  Same hits as the general quad test up to f32 rounding on the rectangles' edges.
"""
import numpy as np
import pytest

import rustraytrace_b200 as R

F = R._ffi


def test_cornell_walls_are_rectangles(monkeypatch):
    s = R.BookScene("the_rest_of_your_life", None, overrides=dict(image_width=64))
    assert R.flatten_rects(s.desc) == (2, 3, 1)          # green + red walls, floor + ceiling + light, back wall
    R.flatten_check(s.desc)                              # every record on its quad's plane
    s8 = R.BookScene("the_next_week", 8, overrides=dict(image_width=64))
    assert R.flatten_rects(s8.desc) == (2, 3, 1)         # the boxes of the smoke scene are media boundaries, not in the leaf
    s7 = R.BookScene("the_next_week", 7, overrides=dict(image_width=64))
    assert R.flatten_rects(s7.desc) == (0, 0, 0)         # 18 quads: a tree, not a root leaf
    monkeypatch.setenv("RRT_NO_RECT_QUADS", "1")
    assert R.flatten_rects(s.desc) == (0, 0, 0)


def test_a_rectangle_after_a_coplanar_slanted_quad_keeps_the_reference_order():
    """Equal-t rule (quad.rs:72): the later quad wins.  Rectangles are tested ahead of the other quads, which keeps
    that rule when the rectangle is the earlier of two coplanar quads (a box on the floor) and would reverse it
    when it is the later one: the flattener then leaves the leaf as it is."""
    b = R.DescBuilder()
    m = b.lambertian((0.5, 0.5, 0.5))
    floor = lambda: b.quad((0.0, 0.0, 0.0), (4.0, 0.0, 0.0), (0.0, 0.0, 4.0), m)
    slanted = lambda: b.quad((1.0, 0.0, 1.0), (1.0, 0.0, 1.0), (-1.0, 0.0, 1.0), m)   # in the plane y = 0, edges off the axes
    wall = b.quad((0.0, 0.0, 4.0), (4.0, 0.0, 0.0), (0.0, 4.0, 0.0), m)
    assert R.flatten_rects(b.desc(b.list([floor(), slanted(), wall]))) == (0, 1, 1)
    assert R.flatten_rects(b.desc(b.list([slanted(), floor(), wall]))) == (0, 0, 0)


@pytest.mark.gpu
@pytest.mark.parametrize("book,scene_no", [("the_rest_of_your_life", None), ("the_next_week", 8), ("the_next_week", 5)])
def test_rectangle_test_finds_the_hits_of_the_quad_test(book, scene_no, monkeypatch):
    s = R.BookScene(book, scene_no, overrides=dict(image_width=300, samples_per_pixel=64, max_depth=12))
    ctx = R.Context(0)
    out = []
    for rects in (True, False):
        if not rects:
            monkeypatch.setenv("RRT_NO_RECT_QUADS", "1")
        gs = ctx.scene(s.desc)
        monkeypatch.delenv("RRT_NO_RECT_QUADS", raising=False)
        ids, t = ctx.probe(gs, s.camera, 32)
        frame = ctx.render(gs, s.camera, s.integrator, seed=11).astype(np.float64)
        out.append((ids, t, frame))
        gs.close()
    ctx.close()
    (ids, t, a), (pids, pt, b) = out
    assert (ids != pids).mean() <= 1e-4                  # edges of the rectangles, rounded the other way
    same = (ids == pids) & (ids >= 0)
    assert np.max(np.abs(t[same] - pt[same]) / pt[same]) <= 2e-6
    assert np.array_equal(a[..., 3], b[..., 3])
    ma, mb = a[..., :3].mean(axis=(0, 1)), b[..., :3].mean(axis=(0, 1))
    assert np.all(np.abs(ma - mb) <= 2e-3 * np.maximum(mb, 1e-6)), (ma, mb)
