"""CPU tier: the flattener's self-check (every primitive in exactly one leaf, child boxes inside their
parents, box leaves naming six quads) on randomly generated scene graphs — the generator of the GPU
fuzz tests, run without a GPU."""
import importlib.util
import os

import numpy as np
import pytest

import rustraytrace_b200 as R

F = R._ffi
_spec = importlib.util.spec_from_file_location("fuzz_gen", os.path.join(os.path.dirname(__file__), "test_fuzz_gpu.py"))


def _generator():
    # test_fuzz_gpu imports the oracle at module level; that is fine on the CPU tier (test infrastructure)
    m = importlib.util.module_from_spec(_spec)
    _spec.loader.exec_module(m)
    return m


@pytest.mark.parametrize("seed", range(40))
def test_random_scene_graph_flattens_to_a_valid_bvh(seed):
    g = _generator()
    b, d, cam = g.random_scene(seed)
    c = R.flatten_check(d)
    n_quads = sum(1 for h in b.h if h.kind == F.H_QUAD)
    n_spheres = sum(1 for h in b.h if h.kind == F.H_SPHERE)
    assert 0 <= c["n_prims"] <= n_quads + n_spheres      # unreferenced primitives are not emitted
    assert c["bvh_depth"] < 64
    feat = R.scene_features(d)
    assert not (feat & (F.SF_MEDIA | F.SF_TEXTURES | F.SF_NOISE | F.SF_MOTION))
    if c["n_prims"] <= 16:
        assert not (feat & F.SF_BOXES)                   # small scenes are one brute-force leaf


@pytest.mark.parametrize("seed", range(16))
def test_outsized_primitives_leave_the_tree_and_the_check_still_holds(seed, monkeypatch):
    """Front run (flatten.cpp): scenes with a ground quad or sphere and a few large spheres among many
    small primitives.  At most four primitives leave the tree, the self-check (every primitive reached
    exactly once) holds with and without it, and the primitive count does not change."""
    g = _generator()
    b, d, cam = g.outsized_scene(seed)
    front = R.flatten_check(d)
    monkeypatch.setenv("RRT_NO_FRONT", "1")
    plain = R.flatten_check(d)
    monkeypatch.delenv("RRT_NO_FRONT")
    assert front["n_prims"] == plain["n_prims"]
    assert plain["n_nodes"] == plain["n_prims"] - 1
    assert 1 <= plain["n_nodes"] - front["n_nodes"] <= 4
    if seed % 2 == 1:   # quads in the tree: only the ground quad (>= 50 % of the bounds) qualifies
        assert plain["n_nodes"] - front["n_nodes"] == 1


def test_box_detection_needs_exactly_the_six_faces(monkeypatch):
    g = _generator()

    def scene(mutate):
        b = R.DescBuilder()
        m = b.lambertian((0.5, 0.5, 0.5))
        filler = [b.sphere((float(i), 5.0, 0.0), 0.3, m) for i in range(20)]   # > 16 primitives: a real BVH
        bx = g.box(b, (0.0, 0.0, 0.0), (1.0, 2.0, 3.0), m)
        node = mutate(b, bx, m)
        return b, b.desc(b.list(filler + [node]))

    b, d = scene(lambda b, bx, m: bx)
    assert R.scene_features(d) & F.SF_BOXES
    b, d = scene(lambda b, bx, m: b.translate(bx, (3.0, -1.0, 2.0)))            # translations keep it axis-aligned
    assert R.scene_features(d) & F.SF_BOXES
    b, d = scene(lambda b, bx, m: b.rotate_y(bx, 15.0))                          # rotated: plain quads
    assert not (R.scene_features(d) & F.SF_BOXES)

    def open_box(b, bx, m):   # five faces and a stray quad: not a box
        (x0, y0, z0), (x1, y1, z1) = (0.0, 0.0, 0.0), (1.0, 2.0, 3.0)
        qs = [b.quad((x0, y0, z1), (x1 - x0, 0, 0), (0, y1 - y0, 0), m), b.quad((x1, y0, z1), (0, 0, z0 - z1), (0, y1 - y0, 0), m),
              b.quad((x1, y0, z0), (x0 - x1, 0, 0), (0, y1 - y0, 0), m), b.quad((x0, y0, z0), (0, 0, z1 - z0), (0, y1 - y0, 0), m),
              b.quad((x0, y1, z1), (x1 - x0, 0, 0), (0, 0, z0 - z1), m), b.quad((x0, y0 - 0.5, z0), (x1 - x0, 0, 0), (0, 0, z1 - z0), m)]
        return b.list(qs)
    b, d = scene(open_box)
    assert not (R.scene_features(d) & F.SF_BOXES)
    for bb, dd in (scene(lambda b, bx, m: bx), scene(open_box)):
        assert R.flatten_check(dd)["n_prims"] == 26
