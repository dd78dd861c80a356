"""CPU tier: the host-side mirror of the reference's Rust surface (librrt_host.so):
src/main.rs book aliases, src/config.rs overrides, src/books scene builders, render_io.rs PPM."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_ffi as O
import rustraytrace_b200 as R

F = R._ffi


def test_xoshiro256pp_reference_vector():
    """rand 0.8.5 SmallRng on 64-bit = Xoshiro256++; its published test vector from state {1,2,3,4}."""
    want = [41943041, 58720359, 3588806011781223, 3591011842654386, 9228616714210784205, 9973669472204895162,
            14011001112246962877, 12406186145184390807, 15849039046786891736, 10450023813501588000]
    assert [F.rrth().rrth_rng_kat_state1234(i) for i in range(10)] == want
    xs = [F.rrth().rrth_rng_kat_seeded_f64(0x5EED1234, i) for i in range(2000)]
    assert all(0.0 <= x < 1.0 for x in xs) and abs(np.mean(xs) - 0.5) < 0.03
    assert xs[:3] == [F.rrth().rrth_rng_kat_seeded_f64(0x5EED1234, i) for i in range(3)]  # deterministic


def test_book_aliases_like_main_rs():
    """src/main.rs:7-12 normalises case/punctuation; :87-92 lists the aliases."""
    n = F.rrth().rrth_normalize_book
    for name in ("in_one_weekend", "InOneWeekend", "in-one-weekend", "one_weekend", "Weekend"):
        assert n(name.encode()) == F.BOOK_1, name
    for name in ("the_next_week", "TheNextWeek", "next_week", "nextweek", "NEXT"):
        assert n(name.encode()) == F.BOOK_2, name
    for name in ("the_rest_of_your_life", "rest_of_your_life", "restofyourlife", "rest", "rest-of-life"):
        assert n(name.encode()) == F.BOOK_3, name
    assert n(b"no_such_book") == 0 and n(b"book1") == 0 and n(b"") == 0
    with pytest.raises(ValueError):
        R.BookScene("no_such_book")


def test_shipped_overrides_and_application():
    """src/config.rs:50-62 ships width 2160, 5000 spp, depth 100; apply_overrides is field-wise."""
    o = R.shipped_overrides()
    assert (o.image_width, o.samples_per_pixel, o.max_depth) == (2160, 5000, 100)
    assert o.has == F.O_BITS["image_width"] | F.O_BITS["samples_per_pixel"] | F.O_BITS["max_depth"]
    s = R.BookScene("the_rest_of_your_life", overrides=o)
    assert (s.camera.image_width, s.camera.samples_per_pixel, s.camera.max_depth) == (2160, 5000, 100)
    assert s.height == 2160 and s.integrator == F.INT_BOOK3
    d = R.BookScene("the_rest_of_your_life")          # scene defaults, the_rest_of_your_life/mod.rs:142-155
    c = d.camera
    assert (c.image_width, c.samples_per_pixel, c.max_depth, c.vfov) == (600, 100, 50, 40.0)
    assert list(c.lookfrom) == [278, 278, -800] and list(c.lookat) == [278, 278, 0] and list(c.background) == [0, 0, 0]
    # book 1 has no background field to override (in_one_weekend/mod.rs:23-55)
    b1 = R.BookScene("in_one_weekend", overrides=dict(background=(1, 0, 0), vfov=33.0, lookfrom=(1, 2, 3)))
    assert b1.camera.vfov == 33.0 and list(b1.camera.lookfrom) == [1, 2, 3] and b1.integrator == F.INT_BOOK1


def test_book1_defaults_and_sphere_counts():
    """in_one_weekend/mod.rs:57-126: ground + grid (skipping those near (4,0.2,0)) + 3 big spheres."""
    s = R.BookScene("in_one_weekend")
    c = s.camera
    assert (c.image_width, c.samples_per_pixel, c.max_depth) == (1200, 10, 20) and s.height == 675
    assert (c.vfov, c.defocus_angle, c.focus_dist) == (20.0, 0.6, 10.0) and list(c.lookfrom) == [13, 2, 3]
    d = s.desc
    kinds = [d.hittables[i].kind for i in range(d.n_hittables)]
    n_sph = kinds.count(F.H_SPHERE)
    assert 400 < n_sph <= 488 and kinds.count(F.H_BVH) == 1 and d.hittables[d.world].kind == F.H_BVH
    mats = [d.materials[d.hittables[i].material].kind for i in range(d.n_hittables) if kinds[i] == F.H_SPHERE]
    frac_diffuse = mats.count(F.M_LAMBERTIAN) / len(mats)
    assert 0.7 < frac_diffuse < 0.9 and F.M_METAL in mats and F.M_DIELECTRIC in mats
    # same seed -> same scene; other seed -> other scene
    s2 = R.BookScene("in_one_weekend")
    s3 = R.BookScene("in_one_weekend", seed=99)
    a = [tuple(d.hittables[i].a) for i in range(d.n_hittables)]
    assert a == [tuple(s2.desc.hittables[i].a) for i in range(s2.desc.n_hittables)]
    assert a != [tuple(s3.desc.hittables[i].a) for i in range(s3.desc.n_hittables)]


def test_book2_scene_table():
    """the_next_week/mod.rs:68-81 scene numbers; :507-624 final scene's 11-entry top-level list."""
    final = R.BookScene("the_next_week", 9)
    assert (final.camera.image_width, final.camera.samples_per_pixel, final.camera.max_depth) == (800, 10000, 40)
    dflt = R.BookScene("the_next_week", None)
    assert (dflt.camera.image_width, dflt.camera.samples_per_pixel, dflt.camera.max_depth) == (400, 250, 4)
    d = final.desc
    top = d.hittables[d.world]
    assert top.kind == F.H_LIST and top.n_children == 11
    kinds = [d.hittables[d.children[top.first_child + k]].kind for k in range(11)]
    assert kinds == [F.H_BVH, F.H_QUAD, F.H_SPHERE, F.H_SPHERE, F.H_SPHERE, F.H_SPHERE, F.H_CONSTANT_MEDIUM,
                     F.H_CONSTANT_MEDIUM, F.H_SPHERE, F.H_SPHERE, F.H_TRANSLATE]
    assert d.n_images == 1 and d.images[0].width == 1024 and d.images[0].height == 512   # earthmap.jpg
    assert d.n_perlins == 1
    smoke = R.BookScene("the_next_week", 8)
    ks = [smoke.desc.hittables[i].kind for i in range(smoke.desc.n_hittables)]
    assert ks.count(F.H_CONSTANT_MEDIUM) == 2 and ks.count(F.H_ROTATE_Y) == 2 and ks.count(F.H_TRANSLATE) == 2
    s1 = R.BookScene("the_next_week", 1)
    assert list(s1.camera.background) == pytest.approx([0.7, 0.8, 1.0])


def test_book3_light_list_is_proxy_geometry():
    """the_rest_of_your_life/mod.rs:128-140: light quad + glass sphere proxies with EmptyMaterial."""
    s = R.BookScene("the_rest_of_your_life")
    d = s.desc
    L = d.hittables[d.lights]
    assert L.kind == F.H_LIST and L.n_children == 2
    q, sp = (d.hittables[d.children[L.first_child + k]] for k in range(2))
    assert q.kind == F.H_QUAD and list(q.a) == [343, 554, 332] and list(q.b) == [-130, 0, 0] and list(q.c) == [0, 0, -105]
    assert sp.kind == F.H_SPHERE and list(sp.a) == [190, 90, 190] and sp.c[0] == 90
    assert d.materials[q.material].kind == F.M_EMPTY and d.materials[sp.material].kind == F.M_EMPTY


def test_perlin_tables_are_permutations_and_unit_vectors():
    """perlin.rs:12-23, 72-77."""
    s = R.BookScene("the_next_week", 4)   # keep the scene alive: desc borrows its arrays
    p = s.desc.perlins[0]
    for perm in (p.perm_x, p.perm_y, p.perm_z):
        assert sorted(perm) == list(range(256))
    v = np.array([list(r) for r in p.randvec])
    assert np.allclose(np.linalg.norm(v, axis=1), 1.0)


def test_camera_frame_two_witnesses():
    """Camera::initialize (camera.rs:108-154) via the oracle vs the closed form of gpu/mod.rs:174-197."""
    s = R.BookScene("in_one_weekend", overrides=dict(image_width=400))
    ci = O.camera_internals(s.camera, F.INT_BOOK1)
    lookfrom, lookat = np.array([13.0, 2, 3]), np.zeros(3)
    h = np.tan(np.radians(20.0) / 2)
    vh = 2 * h * 10.0
    vw = vh * (400 / 225)
    w = (lookfrom - lookat) / np.linalg.norm(lookfrom - lookat)
    u = np.cross([0, 1, 0], w); u /= np.linalg.norm(u)
    v = np.cross(w, u)
    assert ci["image_height"] == 225
    assert np.allclose(ci["pixel_delta_u"], vw * u / 400, atol=1e-14)
    assert np.allclose(ci["pixel_delta_v"], -vh * v / 225, atol=1e-14)
    p00 = lookfrom - 10 * w - vw * u / 2 + vh * v / 2 + 0.5 * (vw * u / 400 - vh * v / 225)
    assert np.allclose(ci["pixel00_loc"], p00, atol=1e-12)
    r = 10 * np.tan(np.radians(0.3))
    assert np.allclose(ci["defocus_disk_u"], u * r) and np.allclose(ci["defocus_disk_v"], v * r)
    # book 3 stratification: 5000 -> 70^2, 4096 -> 64^2, 100 -> 10^2 (camera.rs:114-116)
    for spp, root in [(5000, 70), (4096, 64), (100, 10)]:
        c = R.copy_camera(s.camera, samples_per_pixel=spp)
        k = O.camera_internals(c, F.INT_BOOK3)
        assert k["sqrt_spp"] == root and k["pixel_samples_scale"] == pytest.approx(1 / root ** 2)


def test_ppm_writer_matches_render_io(tmp_path):
    """src/render_io.rs:3-30: "P3\\nW H\\n255\\n" then one "r g b" line per pixel."""
    rng = np.random.default_rng(1)
    acc = (rng.random((5, 7, 4)) * 50).astype(np.float32)
    acc[..., 3] = 50
    acc[0, 0, 0] = np.nan
    path = str(tmp_path / "o.ppm")
    assert F.rrth().rrth_write_ppm_from_accum(path.encode(), 7, 5, acc.ctypes.data, 50) == 0
    lines = open(path).read().split("\n")
    assert lines[:3] == ["P3", "7 5", "255"] and len([x for x in lines[3:] if x]) == 35
    got = np.array([[int(x) for x in ln.split()] for ln in lines[3:] if ln], dtype=np.uint8).reshape(5, 7, 3)
    assert np.array_equal(got, O.resolve_accum(acc, 50))
    p2 = str(tmp_path / "p.ppm")
    assert F.rrth().rrth_write_ppm_rgb8(p2.encode(), 7, 5, got.ctypes.data) == 0
    assert open(p2).read() == open(path).read()


def test_native_jpeg_decode_matches_libjpeg():
    """rtw_image.rs:57-67 decodes with zune-jpeg; our baseline decoder (host/jpeg_decode.cpp) is checked
    against libjpeg-turbo (PIL) on the reference's one asset: same size, texels within a few LSB."""
    from PIL import Image
    w, h = F.c_int(), F.c_int()
    assert F.rrth().rrth_load_image(b"earthmap.jpg", F.IMAGE_DIR.encode(), C.byref(w), C.byref(h), None, 0) == 0
    assert (w.value, h.value) == (1024, 512)
    buf = np.zeros((512, 1024, 3), dtype=np.uint8)
    assert F.rrth().rrth_load_image(b"earthmap.jpg", F.IMAGE_DIR.encode(), C.byref(w), C.byref(h), buf.ctypes.data, buf.nbytes) == 0
    ref = np.asarray(Image.open(os.path.join(F.IMAGE_DIR, "earthmap.jpg")).convert("RGB")).astype(int)
    d = np.abs(buf.astype(int) - ref)
    assert d.max() <= 3 and d.mean() < 0.05
    assert np.allclose(buf.mean(axis=(0, 1)), [81.10, 85.93, 104.10], atol=0.05)
    assert F.rrth().rrth_load_image(b"earthmap.jpg", F.IMAGE_DIR.encode(), None, None, buf.ctypes.data, 10) == -2
    assert F.rrth().rrth_load_image(b"no_such_image.jpg", None, None, None, None, 0) == -1


def test_jpeg_decoder_handles_subsampling_restarts_and_grey(tmp_path):
    from PIL import Image
    rng = np.random.default_rng(3)
    yy, xx = np.mgrid[0:70, 0:93]
    img = np.stack([(xx * 2) % 256, (yy * 3) % 256, ((xx + yy) * 2) % 256], axis=-1).astype(np.uint8)
    img = (img // 2 + rng.integers(0, 20, img.shape)).astype(np.uint8)
    w, h = F.c_int(), F.c_int()
    for name, kw, tol in [("s444.jpg", dict(subsampling=0), 4), ("s420.jpg", dict(subsampling=2), 64),
                          ("s422.jpg", dict(subsampling=1), 64), ("rst.jpg", dict(subsampling=0, restart_marker_blocks=3), 4)]:
        p = str(tmp_path / name)
        Image.fromarray(img).save(p, quality=92, **kw)
        buf = np.zeros((70, 93, 3), dtype=np.uint8)
        assert F.rrth().rrth_load_image(p.encode(), None, C.byref(w), C.byref(h), buf.ctypes.data, buf.nbytes) == 0, name
        assert (w.value, h.value) == (93, 70)
        ref = np.asarray(Image.open(p).convert("RGB")).astype(int)
        d = np.abs(buf.astype(int) - ref)
        assert d.max() <= tol and d.mean() < (0.2 if tol == 4 else 3.0), (name, d.max(), d.mean())   # subsampled: no fancy upsampling here
    p = str(tmp_path / "grey.jpg")
    Image.fromarray(img[..., 0]).save(p, quality=95)
    buf = np.zeros((70, 93, 3), dtype=np.uint8)
    assert F.rrth().rrth_load_image(p.encode(), None, C.byref(w), C.byref(h), buf.ctypes.data, buf.nbytes) == 0
    assert np.abs(buf[..., 0].astype(int) - np.asarray(Image.open(p)).astype(int)).max() <= 2 and np.array_equal(buf[..., 0], buf[..., 2])
    p = str(tmp_path / "prog.jpg")
    Image.fromarray(img).save(p, progressive=True)
    assert F.rrth().rrth_load_image(p.encode(), None, None, None, None, 0) == -1       # rejected loudly, not mis-decoded


def test_jpeg_decoder_rejects_malformed_files(tmp_path):
    """A corrupt earthmap.jpg must be an error, not undefined behaviour: Huffman selectors outside the
    four tables (SOS), component tables longer than their segment (SOF, SOS), truncation anywhere,
    and random bytes after a valid SOI."""
    from PIL import Image
    rng = np.random.default_rng(11)
    img = rng.integers(0, 255, (24, 40, 3)).astype(np.uint8)
    good = str(tmp_path / "good.jpg")
    Image.fromarray(img).save(good, quality=90, subsampling=0)
    data = bytearray(open(good, "rb").read())
    load = lambda p: F.rrth().rrth_load_image(p.encode(), None, None, None, None, 0)
    assert load(good) == 0

    def variant(name, edit):
        d = bytearray(data)
        edit(d)
        p = str(tmp_path / name)
        open(p, "wb").write(bytes(d))
        return p

    sos = data.index(b"\xff\xda")
    sof = data.index(b"\xff\xc0")
    def bad_selectors(d): d[sos + 6] = 0xF7          # Td = 15, Ta = 7 for the first component
    def short_sos(d): d[sos + 2:sos + 4] = b"\x00\x04"  # length 4 < 6 + 2 * 3
    def short_sof(d): d[sof + 2:sof + 4] = b"\x00\x09"  # length 9 < 8 + 3 * 3
    def many_components(d): d[sof + 9] = 4
    assert load(variant("sel.jpg", bad_selectors)) == -1
    assert load(variant("sos.jpg", short_sos)) == -1
    assert load(variant("sof.jpg", short_sof)) == -1
    assert load(variant("nc.jpg", many_components)) == -1
    for cut in (3, sof + 5, sos + 3, sos + 20, len(data) - 40):      # truncated: error or a decode that stays in bounds
        p = str(tmp_path / f"cut{cut}.jpg")
        open(p, "wb").write(bytes(data[:cut]))
        assert load(p) in (0, -1)
    for k in range(20):                                             # garbage after SOI
        p = str(tmp_path / f"fuzz{k}.jpg")
        d = bytearray(data)
        for pos in rng.integers(2, len(d), 12):
            d[pos] = int(rng.integers(0, 256))
        open(p, "wb").write(bytes(d))
        assert load(p) in (0, -1)
