"""The uniform grids of even fields (flatten.cpp choose_grids / emit_grids, intersect.cuh grid_hit32).

An even field of equal items — the small spheres of the book-1 final scene (in_one_weekend/mod.rs:65-108), the
bouncing spheres of book 2 (the_next_week/mod.rs:96-149), the ground boxes and the sphere cloud of the book-2 final
scene (:507-605) — is walked cell by cell instead of through the BVH, and whatever is left sits in the front run.
The grid only changes WHERE the candidates come from: every test is sphere_hit32's / the box-leaf test's, so hits,
equal-t decisions and therefore whole frames must be those of the tree walk — bit for bit, except for boxes that tile
a lattice, where a ray within f32 rounding of a vertical box edge may be settled the other way round.

CPU tier: which scenes get grids, and the flattener's own check of the cell tables (every item referenced from every
cell it reaches).  GPU tier: f32 probe and deterministic renders with and without the grids; forced grids over uneven
random sphere scenes and over random box fields; the one-shot render that builds no tree.
"""
import numpy as np
import pytest

import rustraytrace_b200 as R

F = R._ffi


def random_spheres(seed, n, moving=False, spread=6.0):
    rng = np.random.default_rng(seed)
    b = R.DescBuilder()
    mats = [b.lambertian(tuple(rng.uniform(0.2, 0.9, 3))) for _ in range(3)]
    mats.append(b.material(F.M_METAL, albedo=(0.8, 0.8, 0.8), fuzz=0.1))
    mats.append(b.material(F.M_DIELECTRIC, ri=1.5))
    items = []
    for _ in range(n):
        c = tuple(rng.uniform(-spread, spread, 3) * np.array([1.0, 0.25, 1.0]))
        r = float(rng.choice([0.1, 0.2, 0.2, 0.3, 0.9]))
        m = mats[rng.integers(len(mats))]
        if moving and rng.random() < 0.5:
            c1 = tuple(np.array(c) + rng.uniform(-0.5, 0.5, 3))
            items.append(b.sphere(c, r, m, center2=c1))
        else:
            items.append(b.sphere(c, r, m))
    world = b.bvh(items) if seed % 2 else b.list(items)
    cam = R.make_camera(image_width=128, aspect_ratio=1.0, samples_per_pixel=16, max_depth=8, background=(0.6, 0.7, 0.9),
                        vfov=50.0, lookfrom=(float(rng.uniform(-3, 3)), float(rng.uniform(1.0, 4.0)), 11.0), lookat=(0.0, 0.0, 0.0))
    return b.desc(world), cam


# ---- CPU tier --------------------------------------------------------------------------------------
def test_grid_is_laid_over_even_fields_only(monkeypatch):
    for book, no, want in [("in_one_weekend", None, True), ("in_one_weekend_gpu", None, True), ("the_next_week", 1, True),
                           ("the_next_week", 2, False), ("the_next_week", 7, False), ("the_next_week", 8, False),
                           ("the_rest_of_your_life", None, False)]:
        s = R.BookScene(book, no, overrides=dict(image_width=64))
        g = R.flatten_grid(s.desc, 0)
        assert (g["res"] != (0, 0, 0)) == want, (book, no, g)
        assert R.flatten_grid(s.desc, 1)["res"] == (0, 0, 0)
        if want:
            n = R.flatten_check(s.desc)["n_prims"]       # the check walks the cell tables too
            assert g["res"][1] == 1                      # a field on the ground: one cell thick
            assert g["n_items"] == n - 4                 # everything but the ground and the three large spheres (front run)
            assert g["n_refs"] <= 4 * n and g["max_per_cell"] <= 12
    s = R.BookScene("in_one_weekend", None, overrides=dict(image_width=64))
    monkeypatch.setenv("RRT_NO_GRID", "1")
    assert R.flatten_grid(s.desc)["res"] == (0, 0, 0)


def test_book2_final_scene_is_two_grids_and_a_front_run(monkeypatch):
    """the_next_week/mod.rs:297-401: 400 ground boxes that tile a 20 x 20 lattice (one box per cell), the cluster of
    1 000 spheres of radius 10, and seven other primitives, which all fit the front run — no tree left to walk."""
    s = R.BookScene("the_next_week", 9, overrides=dict(image_width=64))
    boxes, spheres = R.flatten_grid(s.desc, 1), R.flatten_grid(s.desc, 0)
    assert boxes["res"] == (20, 1, 20) and boxes["n_items"] == 400 and boxes["n_refs"] == 400 and boxes["max_per_cell"] == 1
    assert spheres["n_items"] == 1000 and min(spheres["res"]) >= 4
    c = R.flatten_check(s.desc)
    assert c["n_prims"] == 400 * 6 + 1000 + 7
    monkeypatch.setenv("RRT_GRID_NO_TILING", "1")        # the same boxes in a padded grid: several per cell
    padded = R.flatten_grid(s.desc, 1)
    assert padded["n_items"] == 400 and padded["max_per_cell"] > 1
    R.flatten_check(s.desc)


@pytest.mark.parametrize("seed", range(6))
def test_forced_grid_passes_the_flattener_check(seed, monkeypatch):
    """Uneven fields (radii 0.1-0.9, clumps, moving spheres): not what the balance test accepts, but a forced
    grid must still reference every sphere from every cell its swept box touches."""
    d, _ = random_spheres(seed, 40 + 60 * seed, moving=bool(seed & 1))
    monkeypatch.setenv("RRT_FORCE_GRID", "1")
    g = R.flatten_grid(d)
    assert g["res"] != (0, 0, 0) and g["n_refs"] >= 40 + 60 * seed
    R.flatten_check(d)


def test_small_or_mixed_trees_keep_the_bvh():
    d, _ = random_spheres(1, 20)
    assert R.flatten_grid(d)["res"] == (0, 0, 0)        # under RRT_GRID_MIN_ITEMS
    b = R.DescBuilder()
    m = b.lambertian((0.5, 0.5, 0.5))
    items = [b.sphere((float(x), 0.2, float(z)), 0.2, m) for x in range(-6, 6) for z in range(-6, 6)]
    assert R.flatten_grid(b.desc(b.list(items)))["res"][1] == 1
    for k in range(16):   # sixteen quads next to the field: more than the front run holds, so the tree stays
        items.append(b.quad((-1.0 + k, 0.5, -1.0), (0.5, 0.0, 0.0), (0.0, 0.0, 0.5), m))
    assert R.flatten_grid(b.desc(b.list(items)))["res"] == (0, 0, 0)
    R.flatten_check(b.desc(b.list(items)))


# ---- GPU tier --------------------------------------------------------------------------------------
def _both(desc, cam, integrator, monkeypatch, force=False):
    """(probe ids, probe t, deterministic frame) with the grid and with RRT_NO_GRID."""
    monkeypatch.setenv("RRT_DETERMINISTIC", "1")
    ctx = R.Context(0)
    monkeypatch.delenv("RRT_DETERMINISTIC")
    out = []
    for grid in (True, False):
        if not grid:
            monkeypatch.setenv("RRT_NO_GRID", "1")
        elif force:
            monkeypatch.setenv("RRT_FORCE_GRID", "1")
        gs = ctx.scene(desc)
        monkeypatch.delenv("RRT_NO_GRID", raising=False)
        monkeypatch.delenv("RRT_FORCE_GRID", raising=False)
        ids, t = ctx.probe(gs, cam, 32)
        frame = ctx.render(gs, cam, integrator, seed=5)
        cells = ctx.stats()["n_grid_cells"]
        assert (cells > 0) == grid
        out.append((ids, t, frame))
        gs.close()
    ctx.close()
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("book,scene_no", [("in_one_weekend", None), ("in_one_weekend_gpu", None), ("the_next_week", 1)])
def test_grid_and_tree_render_the_same_frame_bit_for_bit(book, scene_no, monkeypatch):
    s = R.BookScene(book, scene_no, overrides=dict(image_width=200, samples_per_pixel=24, max_depth=16))
    (ids, t, a), (pids, pt, b) = _both(s.desc, s.camera, s.integrator, monkeypatch)
    assert np.array_equal(ids, pids) and np.array_equal(t, pt)
    assert np.array_equal(a, b)


@pytest.mark.gpu
def test_book2_final_scene_grids_match_the_tree(monkeypatch):
    """Boxes that tile their lattice are listed in their own cell only, so a ray within f32 rounding of a vertical box
    edge may be settled the other way round than by the tree: identical frames up to such pixels (none expected at
    this size), identical sample counts."""
    s = R.BookScene("the_next_week", 9, overrides=dict(image_width=200, samples_per_pixel=16, max_depth=12))
    (ids, t, a), (pids, pt, b) = _both(s.desc, s.camera, s.integrator, monkeypatch)
    assert (ids != pids).mean() <= 1e-4 and (t != pt).mean() <= 1e-4
    assert np.array_equal(a[..., 3], b[..., 3])
    assert (np.abs(a[..., :3] - b[..., :3]).max(axis=2) > 1e-6 * np.maximum(b[..., :3].max(axis=2), 1e-3)).mean() <= 1e-3


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(6))
def test_forced_grid_over_uneven_spheres_matches_the_tree(seed, monkeypatch):
    d, cam = random_spheres(seed, 40 + 60 * seed, moving=bool(seed & 1))
    integ = F.INT_BOOK2 if seed & 1 else F.INT_BOOK1
    (ids, t, a), (pids, pt, b) = _both(d, cam, integ, monkeypatch, force=True)
    assert (ids >= 0).any()
    assert np.array_equal(ids, pids) and np.array_equal(t, pt)
    assert np.array_equal(a, b)


@pytest.mark.gpu
def test_grid_probe_against_the_oracle():
    """The grid walk in the f32 probe against the oracle's f64 pointer tree: ids equal except on silhouettes."""
    import oracle_ffi as O
    s = R.BookScene("in_one_weekend", None, overrides=dict(image_width=400))
    ctx = R.Context(0)
    gs = ctx.scene(s.desc)
    assert ctx.stats()["n_grid_cells"] > 0
    ids, t = ctx.probe(gs, s.camera, 32)
    oids, ot = O.OracleScene(s.desc).probe(s.camera)
    gs.close(); ctx.close()
    assert (ids != oids).mean() <= 2e-5
    hit = (oids >= 0) & (ids == oids)
    rel = np.abs(t[hit] - ot[hit]) / ot[hit]
    assert np.percentile(rel, 99) <= 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("book,scene_no", [("in_one_weekend", None), ("the_next_week", 9)])
def test_one_shot_render_without_a_tree_is_the_same_frame(book, scene_no, monkeypatch):
    """rrt_render_frame uploads, renders and frees: where the scene is walked through grids it builds no BVH at all
    (flatten_scene tree_optional).  Same primitives, same cells, same frame as the scene handle that has one."""
    s = R.BookScene(book, scene_no, overrides=dict(image_width=160, samples_per_pixel=16, max_depth=12))
    monkeypatch.setenv("RRT_DETERMINISTIC", "1")
    ctx = R.Context(0)
    monkeypatch.delenv("RRT_DETERMINISTIC")
    gs = ctx.scene(s.desc)
    assert ctx.stats()["n_bvh_nodes"] > 0
    a = ctx.render(gs, s.camera, s.integrator, seed=8)
    gs.close()
    b = ctx.render_frame(s.desc, s.camera, s.integrator, seed=8)
    st = ctx.stats()
    ctx.close()
    assert st["n_grid_cells"] > 0 and st["n_bvh_nodes"] == 0
    assert np.array_equal(a, b)


def random_boxes(seed, n_boxes, n_spheres, lattice):
    """Axis-aligned boxes (make_box: six quads as a list) — on a lattice of equal footprints or scattered with random
    sizes — plus a cloud of equal spheres and a few odd primitives for the front run."""
    rng = np.random.default_rng(1000 + seed)
    b = R.DescBuilder()
    mats = [b.lambertian(tuple(rng.uniform(0.2, 0.9, 3))) for _ in range(3)] + [b.material(F.M_DIELECTRIC, ri=1.5)]

    def box(lo, hi, mat):
        (x0, y0, z0), (x1, y1, z1) = lo, hi
        dx, dy, dz = (x1 - x0, 0, 0), (0, y1 - y0, 0), (0, 0, z1 - z0)
        neg = lambda v: tuple(-c for c in v)
        return b.list([b.quad((x0, y0, z1), dx, dy, mat), b.quad((x1, y0, z1), neg(dz), dy, mat), b.quad((x1, y0, z0), neg(dx), dy, mat),
                       b.quad((x0, y0, z0), dz, dy, mat), b.quad((x0, y1, z1), dx, neg(dz), mat), b.quad((x0, y0, z0), dx, dz, mat)])
    items = []
    side = int(np.ceil(np.sqrt(n_boxes)))
    for k in range(n_boxes):
        m = mats[rng.integers(3)]
        if lattice:
            i, j = k % side, k // side
            lo = (-4.0 + i * 1.0, -1.0, -4.0 + j * 1.0)
            hi = (lo[0] + 1.0, -1.0 + float(rng.uniform(0.1, 1.0)), lo[2] + 1.0)
        else:
            lo = tuple(rng.uniform(-4, 3.5, 3) * np.array([1.0, 0.3, 1.0]))
            hi = tuple(np.array(lo) + rng.uniform(0.3, 0.6, 3))
        items.append(box(lo, hi, m))
    for _ in range(n_spheres):
        items.append(b.sphere(tuple(rng.uniform(-3, 3, 3) * np.array([1.0, 0.5, 1.0]) + np.array([0, 2.0, 0])), 0.25, mats[rng.integers(4)]))
    items.append(b.sphere((0.0, -1001.0, 0.0), 1000.0, mats[0]))
    items.append(b.quad((-2.0, 4.0, -2.0), (4.0, 0.0, 0.0), (0.0, 0.0, 4.0), mats[1]))
    cam = R.make_camera(image_width=144, aspect_ratio=1.0, samples_per_pixel=16, max_depth=8, background=(0.6, 0.7, 0.9),
                        vfov=45.0, lookfrom=(float(rng.uniform(-3, 3)), float(rng.uniform(2.0, 5.0)), 12.0), lookat=(0.0, 0.0, 0.0))
    return b.desc(b.list(items)), cam


@pytest.mark.parametrize("seed,lattice", [(0, True), (1, True), (2, False), (3, False)])
def test_box_fields_pass_the_flattener_check(seed, lattice, monkeypatch):
    d, _ = random_boxes(seed, 64 + 17 * seed, 80, lattice)
    monkeypatch.setenv("RRT_FORCE_GRID", "1")
    boxes, spheres = R.flatten_grid(d, 1), R.flatten_grid(d, 0)
    assert boxes["n_items"] == 64 + 17 * seed and spheres["n_items"] >= 80
    if lattice and seed == 0:
        assert boxes["max_per_cell"] == 1 and boxes["res"][0] == 8 and boxes["res"][2] == 8    # 64 boxes tile 8 x 8
    R.flatten_check(d)


@pytest.mark.gpu
@pytest.mark.parametrize("seed,lattice", [(0, True), (1, True), (2, False), (3, False)])
def test_box_and_sphere_grids_match_the_tree(seed, lattice, monkeypatch):
    """Scattered boxes go through the padded membership (every cell their bounds touch), lattice boxes through the
    one-box-per-cell tiles; either way the walk must find the tree's hits (lattice: up to roundings at box edges)."""
    d, cam = random_boxes(seed, 64 + 17 * seed, 80, lattice)
    (ids, t, a), (pids, pt, b) = _both(d, cam, F.INT_BOOK2, monkeypatch, force=True)
    assert (ids >= 0).any()
    if lattice:
        assert (ids != pids).mean() <= 2e-4
        assert np.array_equal(a[..., 3], b[..., 3])
        assert (np.abs(a[..., :3] - b[..., :3]).max(axis=2) > 1e-6 * np.maximum(b[..., :3].max(axis=2), 1e-3)).mean() <= 2e-3
    else:
        assert np.array_equal(ids, pids) and np.array_equal(t, pt)
        assert np.array_equal(a, b)
