"""GPU tier: randomly generated scene graphs (nested lists / BVHs / Translate / RotateY instances,
spheres, quads, boxes, exactly coincident primitives) — probe ids must equal the oracle's on every
pixel, t within 1e-5, and short renders must agree in the mean.  Covers the edge cases the book
scenes do not: empty worlds, single primitives, deep nesting, ties decided by list/BVH order.

Known limit, kept out of the generator: coincident primitives INSIDE a rotated instance.  The
reference intersects them in object space (identical t); the flattener bakes instances to world
space in f64, where the two baked planes round independently, so such a tie is decided by rounding.
"""
import numpy as np
import pytest

import oracle_ffi as O
import rustraytrace_b200 as R

F = R._ffi
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = R.Context(0)
    yield c
    c.close()


def box(b, lo, hi, mat):
    """make_box (quad.rs:112-134): six quads, fixed order, as a list."""
    (x0, y0, z0), (x1, y1, z1) = lo, hi
    dx, dy, dz = (x1 - x0, 0, 0), (0, y1 - y0, 0), (0, 0, z1 - z0)
    neg = lambda v: tuple(-c for c in v)
    qs = [b.quad((x0, y0, z1), dx, dy, mat), b.quad((x1, y0, z1), neg(dz), dy, mat), b.quad((x1, y0, z0), neg(dx), dy, mat),
          b.quad((x0, y0, z0), dz, dy, mat), b.quad((x0, y1, z1), dx, neg(dz), mat), b.quad((x0, y0, z0), dx, dz, mat)]
    return b.list(qs)


def random_scene(seed):
    rng = np.random.default_rng(seed)
    b = R.DescBuilder()
    mats = [b.lambertian(tuple(rng.uniform(0.2, 0.9, 3))) for _ in range(3)]
    mats.append(b.material(F.M_METAL, albedo=(0.8, 0.8, 0.8), fuzz=0.1))
    mats.append(b.material(F.M_DIELECTRIC, ri=1.5))

    def prims(n, allow_ties):
        out = []
        for _ in range(n):
            m = mats[rng.integers(len(mats))]
            kind = rng.integers(4)
            if kind == 0:
                out.append(b.sphere(tuple(rng.uniform(-4, 4, 3)), float(rng.uniform(0.1, 1.5)), m))
            elif kind == 1:
                out.append(b.quad(tuple(rng.uniform(-4, 4, 3)), tuple(rng.uniform(-2, 2, 3)), tuple(rng.uniform(-2, 2, 3)), m))
            elif kind == 2:
                lo = rng.uniform(-4, 3, 3)
                out.append(box(b, tuple(lo), tuple(lo + rng.uniform(0.3, 1.5, 3)), m))
            elif allow_ties and out:   # an exact duplicate of an earlier primitive: a guaranteed equal-t tie
                h = b.h[out[rng.integers(len(out))]]
                if h.kind == F.H_SPHERE:
                    out.append(b.sphere(tuple(h.a), h.c[0], m))
                elif h.kind == F.H_QUAD:
                    out.append(b.quad(tuple(h.a), tuple(h.b), tuple(h.c), m))
        return out

    top = prims(int(rng.integers(0, 12)), True)
    for _ in range(int(rng.integers(0, 4))):          # containers, possibly instanced
        inner = prims(int(rng.integers(1, 9)), False)
        node = b.bvh(inner) if rng.integers(2) else b.list(inner)
        if rng.integers(2):
            node = b.rotate_y(node, float(rng.uniform(-80, 80)))
        if rng.integers(2):
            node = b.translate(node, tuple(rng.uniform(-2, 2, 3)))
        if rng.integers(3) == 0:                       # instance of an instance
            node = b.translate(b.rotate_y(node, float(rng.uniform(-40, 40))), tuple(rng.uniform(-1, 1, 3)))
        top.append(node)
    world = b.bvh(top) if (len(top) > 2 and rng.integers(2)) else b.list(top)
    cam = R.make_camera(image_width=160, aspect_ratio=1.0, samples_per_pixel=64, max_depth=8, background=(0.6, 0.7, 0.9),
                        vfov=float(rng.uniform(40, 90)), lookfrom=tuple(rng.uniform(-1, 1, 3) + np.array([0, 0, 9.0])),
                        lookat=tuple(rng.uniform(-1, 1, 3)))
    return b, b.desc(world), cam


@pytest.mark.parametrize("seed", range(24))
def test_random_scene_probe_parity(ctx, seed):
    b, d, cam = random_scene(seed)
    gs = ctx.scene(d)
    ids, t = ctx.probe(gs, cam, 64)
    oids, ot = O.OracleScene(d).probe(cam)
    assert np.array_equal(ids, oids), f"seed {seed}: {(ids != oids).sum()} ids differ"
    hit = oids >= 0
    if hit.any():
        assert np.max(np.abs(t[hit] - ot[hit]) / ot[hit]) <= 1e-5
    assert np.all(t[~hit] == -1.0)
    gs.close()


def outsized_scene(seed):
    """Many small primitives and a few far larger ones (a ground quad or sphere, a few big spheres):
    what the flattener's front run takes out of the BVH (flatten.cpp)."""
    rng = np.random.default_rng(1000 + seed)
    b = R.DescBuilder()
    mats = [b.lambertian(tuple(rng.uniform(0.2, 0.9, 3))) for _ in range(3)]
    mats.append(b.material(F.M_METAL, albedo=(0.8, 0.8, 0.8), fuzz=0.1))
    mats.append(b.material(F.M_DIELECTRIC, ri=1.5))
    pick = lambda: mats[rng.integers(len(mats))]
    top = []
    with_quads = seed % 2 == 1
    for _ in range(int(rng.integers(24, 60))):
        c = (float(rng.uniform(-6, 6)), float(rng.uniform(0.1, 0.6)), float(rng.uniform(-6, 6)))
        if with_quads and rng.integers(3) == 0:
            top.append(b.quad(c, tuple(rng.uniform(-0.4, 0.4, 3)), tuple(rng.uniform(-0.4, 0.4, 3)), pick()))
        else:
            top.append(b.sphere(c, float(rng.uniform(0.08, 0.3)), pick()))
    if with_quads:   # a ground quad spanning far more than half of the bounds
        top.insert(int(rng.integers(len(top))), b.quad((-500.0, 0.0, -500.0), (1000.0, 0, 0), (0, 0, 1000.0), pick()))
    else:            # the book-1 ground sphere and a few large spheres among the small ones
        top.insert(int(rng.integers(len(top))), b.sphere((0.0, -1000.0, 0.0), 1000.0, pick()))
        for k in range(int(rng.integers(1, 6))):
            top.insert(int(rng.integers(len(top))), b.sphere((float(rng.uniform(-5, 5)), 1.0, float(rng.uniform(-5, 5))), float(rng.uniform(0.9, 1.6)), pick()))
    world = b.bvh(top) if rng.integers(2) else b.list(top)
    cam = R.make_camera(image_width=160, aspect_ratio=1.0, samples_per_pixel=64, max_depth=8, background=(0.6, 0.7, 0.9),
                        vfov=float(rng.uniform(30, 60)), lookfrom=(float(rng.uniform(-3, 3)), float(rng.uniform(1.5, 4)), 12.0),
                        lookat=(0.0, 0.5, 0.0))
    return b, b.desc(world), cam


@pytest.mark.parametrize("seed", range(8))
def test_outsized_primitives_probe_parity_and_front_equivalence(ctx, seed, monkeypatch):
    """Front run against the oracle's pointer-tree BVH (ids bit-exact, t within 1e-5), and against
    the same scene with every primitive kept in the tree (RRT_NO_FRONT): same hits, same frame."""
    b, d, cam = outsized_scene(seed)
    assert R.flatten_check(d)["n_nodes"] < R.flatten_check(d)["n_prims"] - 1   # something left the tree
    gs = ctx.scene(d)
    ids, t = ctx.probe(gs, cam, 64)
    oids, ot = O.OracleScene(d).probe(cam)
    assert np.array_equal(ids, oids), f"seed {seed}: {(ids != oids).sum()} ids differ"
    hit = oids >= 0
    assert hit.any() and np.max(np.abs(t[hit] - ot[hit]) / ot[hit]) <= 1e-5
    cam2 = R.copy_camera(cam, image_width=96, samples_per_pixel=32)
    a = ctx.render(gs, cam2, F.INT_BOOK2, seed=seed).astype(np.float64)
    gs.close()
    monkeypatch.setenv("RRT_NO_FRONT", "1")
    gp = ctx.scene(d)
    monkeypatch.delenv("RRT_NO_FRONT")
    pids, pt = ctx.probe(gp, cam, 64)
    q = ctx.render(gp, cam2, F.INT_BOOK2, seed=seed).astype(np.float64)
    gp.close()
    assert np.array_equal(ids, pids) and np.array_equal(t, pt)
    assert np.array_equal(a[..., 3], q[..., 3])
    la, lq = a[..., :3].sum(axis=2), q[..., :3].sum(axis=2)
    differs = np.abs(la - lq) > 1e-4 * np.maximum(lq, 1e-3)
    assert differs.mean() < 0.002, differs.mean()


@pytest.mark.parametrize("seed", [3, 7, 11, 19])
def test_random_scene_render_mean(ctx, seed):
    b, d, cam = random_scene(seed)
    cam = R.copy_camera(cam, image_width=64, samples_per_pixel=1024)
    gs = ctx.scene(d)
    acc = ctx.render(gs, cam, F.INT_BOOK2, seed=seed)
    a, _ = O.OracleScene(d).render(cam, F.INT_BOOK2, seed=100 + seed)
    g = acc[..., :3].astype(np.float64).mean(axis=(0, 1)) / 1024
    r = a.mean(axis=(0, 1))
    assert np.all(np.abs(g - r) / r < 0.01), (g, r)
    gs.close()


def test_empty_and_single_primitive_worlds(ctx):
    b = R.DescBuilder()
    d = b.desc(b.list([]))
    cam = R.make_camera(image_width=32, samples_per_pixel=4, max_depth=5, background=(0.25, 0.5, 1.0))
    gs = ctx.scene(d)
    ids, t = ctx.probe(gs, cam, 64)
    assert np.all(ids == -1) and np.all(t == -1.0)
    acc = ctx.render(gs, cam, F.INT_BOOK2)
    assert np.allclose(acc[..., :3] / 4, [0.25, 0.5, 1.0]) and np.all(acc[..., 3] == 4)
    gs.close()
    b2 = R.DescBuilder()
    s = b2.sphere((0, 0, -3), 1.0, b2.lambertian((0.5, 0.5, 0.5)))
    d2 = b2.desc(s)                                   # the world IS one primitive, not a list
    gs2 = ctx.scene(d2)
    ids, _ = ctx.probe(gs2, cam, 64)
    oids, _ = O.OracleScene(d2).probe(cam)
    assert np.array_equal(ids, oids) and (ids == s).any()
    gs2.close()


@pytest.mark.parametrize("seed", [1, 2, 5, 8, 13, 21])
def test_box_leaves_change_nothing_but_the_tree(ctx, seed, monkeypatch):
    """Axis-aligned boxes become one BVH leaf each (slab test + entry-face quad test, all six quads when
    that does not settle it).  With the grouping off (RRT_NO_BOX_LEAVES, read by the flattener) the
    same rays must find the same hits: frames agree up to float summation order, including scenes
    with glass boxes (rays starting inside a box) and metal ones (rays leaving along a face)."""
    b, d, cam = random_scene(seed)
    extra = [box(b, (-3.5 + 2.2 * k, -1.0, 1.0 + 0.3 * k), (-2.0 + 2.2 * k, 0.4 + 0.2 * k, 2.4), k % 5) for k in range(3)]
    d = b.desc(b.list([d.world] + extra))   # three more boxes: lambertian, metal, glass among them
    cam = R.copy_camera(cam, image_width=96, samples_per_pixel=64)
    gs = ctx.scene(d)
    a = ctx.render(gs, cam, F.INT_BOOK2, seed=seed).astype(np.float64)
    gs.close()
    monkeypatch.setenv("RRT_NO_BOX_LEAVES", "1")
    gq = ctx.scene(d)
    monkeypatch.delenv("RRT_NO_BOX_LEAVES")
    q = ctx.render(gq, cam, F.INT_BOOK2, seed=seed).astype(np.float64)
    gq.close()
    assert np.array_equal(a[..., 3], q[..., 3])
    la, lq = a[..., :3].sum(axis=2), q[..., :3].sum(axis=2)
    differs = np.abs(la - lq) > 1e-4 * np.maximum(lq, 1e-3)
    assert differs.mean() < 0.002, differs.mean()     # exact edge ties may pick the adjacent face
    assert np.all(np.abs(a[..., :3].mean(axis=(0, 1)) - q[..., :3].mean(axis=(0, 1))) <= 1e-3 * q[..., :3].mean(axis=(0, 1)))
