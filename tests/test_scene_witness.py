"""librrt_host.so's scene builders (host/books.cpp) against an independent restatement of the reference's Rust
scene code (tests/scene_witness.py): same seed, same order of draws — every primitive, transform, material,
texture, Perlin table and light of the scenes behind C1-C5 (book-1 final, bouncing spheres, earth, Perlin spheres, book-2
final, book-3 Cornell box) must agree.
The oracle and the GPU both consume books.cpp's output, so this is what makes a scene-builder error visible."""
import math

import rustraytrace_b200 as R
import scene_witness as W

F = R._ffi
SEED = 0x5EED1234   # BookScene's default, the reference's GPU-path seed (src/gpu/mod.rs:199)
MAT = {F.M_LAMBERTIAN: W.LAMBERTIAN, F.M_METAL: W.METAL, F.M_DIELECTRIC: W.DIELECTRIC, F.M_DIFFUSE_LIGHT: W.DIFFUSE_LIGHT,
       F.M_ISOTROPIC: W.ISOTROPIC, F.M_EMPTY: "empty"}


def v3(d):
    return (d[0], d[1], d[2])


def walk(desc, idx):
    """The C description as the nested form of scene_witness."""
    h = desc.hittables[idx]
    def tex(t):
        if t < 0:
            return None
        x = desc.textures[t]
        if x.kind == F.T_SOLID:
            return ("solid", v3(x.color))
        if x.kind == F.T_IMAGE:
            return ("image",)
        if x.kind == F.T_CHECKER:
            return ("checker", x.scale, tex(x.even), tex(x.odd))
        if x.kind == F.T_NOISE:
            p = desc.perlins[x.perlin]
            return ("noise", x.scale, ([v3(p.randvec[i]) for i in range(256)], list(p.perm_x), list(p.perm_y), list(p.perm_z)))
        raise AssertionError(f"texture kind {x.kind}")
    def mat(m):
        x = desc.materials[m]
        return (MAT[x.kind], v3(x.albedo), x.fuzz, x.refraction_index, tex(x.texture))
    kids = lambda: [walk(desc, desc.children[h.first_child + k]) for k in range(h.n_children)]
    if h.kind == F.H_SPHERE:
        return ("sphere", v3(h.a), v3(h.b), h.c[0], mat(h.material))
    if h.kind == F.H_QUAD:
        return ("quad", v3(h.a), v3(h.b), v3(h.c), mat(h.material))
    if h.kind == F.H_LIST:
        return ("list", kids())
    if h.kind == F.H_BVH:
        return ("bvh", kids())
    if h.kind == F.H_TRANSLATE:
        return ("translate", v3(h.a), walk(desc, h.first_child))
    if h.kind == F.H_ROTATE_Y:
        return ("rotate_y", h.a[0], walk(desc, h.first_child))
    if h.kind == F.H_CONSTANT_MEDIUM:
        return ("medium", h.a[0], mat(h.material), walk(desc, h.first_child))
    raise AssertionError(f"hittable kind {h.kind}")


def same(a, b, path="world"):
    """Deep comparison; floats to 1e-12 relative (the two sides may associate a product differently)."""
    if isinstance(a, float) or isinstance(b, float):
        assert math.isclose(a, b, rel_tol=1e-12, abs_tol=1e-300), f"{path}: {a!r} != {b!r}"
    elif isinstance(a, (tuple, list)):
        assert isinstance(b, (tuple, list)) and len(a) == len(b), f"{path}: {len(a)} items against {len(b) if hasattr(b, '__len__') else b!r}"
        for k, (x, y) in enumerate(zip(a, b)):
            same(x, y, f"{path}[{k}]")
    else:
        assert a == b, f"{path}: {a!r} != {b!r}"


def same_material(got, want, path):
    """books.cpp may carry redundant fields (an albedo next to a solid texture, fuzz 0 on a Lambertian): compare what
    the material kind reads (material.rs)."""
    kind = want[0]
    assert got[0] == kind, f"{path}: {got[0]} != {kind}"
    if kind in (W.LAMBERTIAN, W.DIFFUSE_LIGHT, W.ISOTROPIC):
        gt, wt = got[4], want[4]
        if wt[0] == "solid":
            colour = gt[1] if gt is not None and gt[0] == "solid" else got[1]
            same(colour, wt[1], path + ".colour")
        else:
            same(gt, wt, path + ".texture")
    elif kind == W.METAL:
        same(got[1], want[1], path + ".albedo"); same(got[2], want[2], path + ".fuzz")
    elif kind == W.DIELECTRIC:
        same(got[3], want[3], path + ".ri")


def same_scene(got, want, path="world"):
    assert got[0] == want[0], f"{path}: {got[0]} != {want[0]}"
    kind = want[0]
    if kind in ("list", "bvh"):
        assert len(got[1]) == len(want[1]), f"{path}: {len(got[1])} children against {len(want[1])}"
        for k, (x, y) in enumerate(zip(got[1], want[1])):
            same_scene(x, y, f"{path}.{kind}[{k}]")
    elif kind in ("translate", "rotate_y"):
        same(got[1], want[1], path + "." + kind); same_scene(got[2], want[2], path + "." + kind)
    elif kind == "medium":
        same(got[1], want[1], path + ".density"); same_material(got[2], want[2], path + ".phase"); same_scene(got[3], want[3], path + ".boundary")
    else:   # sphere, quad
        same(got[1:4], want[1:4], path + "." + kind); same_material(got[4], want[4], path + ".material")


def same_camera(cam, want):
    for k, v in want.items():
        g = getattr(cam, k)
        same(v3(g) if isinstance(v, tuple) else g, v, "camera." + k)


def test_rng_witness_agrees_with_the_host_rng():
    g = W.SmallRng(0x5EED1234)
    for k in range(64):
        assert g.random_double() == F.rrth().rrth_rng_kat_seeded_f64(0x5EED1234, k)


def test_book1_final_scene_matches_the_rust_source():
    s = R.BookScene("in_one_weekend", None, seed=SEED)
    world, cam = W.book1_final(SEED)
    same_scene(walk(s.desc, s.desc.world), world)
    same_camera(s.camera, cam)
    assert len(world[1]) == 486


def test_book2_final_scene_matches_the_rust_source():
    s = R.BookScene("the_next_week", 9, seed=SEED)
    world, cam = W.book2_final(SEED)
    same_scene(walk(s.desc, s.desc.world), world)
    same_camera(s.camera, cam)


def test_book2_configs_c3_match_the_rust_source():
    """Bouncing spheres (C3a: moving spheres, checker ground), earth (C3b), Perlin spheres (C3c: one shared noise texture)."""
    for no, (world, cam) in ((1, W.book2_bouncing_spheres(SEED)), (3, W.book2_earth()), (4, W.book2_perlin_spheres(SEED))):
        s = R.BookScene("the_next_week", no, seed=SEED, overrides={})
        same_scene(walk(s.desc, s.desc.world), world, f"scene{no}")
        same_camera(s.camera, cam)


def test_book3_cornell_box_and_its_light_list_match_the_rust_source():
    s = R.BookScene("the_rest_of_your_life", None, seed=SEED, overrides={})
    world, lights, cam = W.book3_cornell()
    same_scene(walk(s.desc, s.desc.world), world)
    assert s.desc.lights >= 0
    same_scene(walk(s.desc, s.desc.lights), lights, "lights")
    same_camera(s.camera, cam)
