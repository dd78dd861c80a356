"""CPU tier: the ORACLE against known answers derived by hand from the reference source
(SURVEY.md §4 — the reference ships no tests or golden vectors, so these KATs are the pin)."""
import math

import numpy as np
import pytest

import oracle_ffi as O
import rustraytrace_b200 as R

F = R._ffi


def one_sphere():
    b = R.DescBuilder()
    m = b.lambertian((0.5, 0.5, 0.5))
    sp = b.sphere((0, 0, -1), 0.5, m)
    return b, O.OracleScene(b.desc(b.list([sp]))), sp


def test_sphere_hit_record():
    """sphere.rs:47-53, 69-96: t=0.5, front face, normal (0,0,1), uv=(0.25,0.5)."""
    _, o, sp = one_sphere()
    h = o.hit((0, 0, 0), (0, 0, -1))
    assert h["id"] == sp and h["front"]
    assert h["t"] == pytest.approx(0.5, abs=1e-15)
    assert h["n"] == pytest.approx([0, 0, 1], abs=1e-15)
    assert (h["u"], h["v"]) == (pytest.approx(0.25, abs=1e-12), pytest.approx(0.5, abs=1e-12))
    # from inside: far root, normal flipped against the ray (hittable.rs:49-52)
    h2 = o.hit((0, 0, -1), (0, 0, -1))
    assert h2["t"] == pytest.approx(0.5) and not h2["front"] and h2["n"] == pytest.approx([0, 0, 1])
    # unnormalised direction: t is in units of |d| (ray.rs:31-33)
    assert o.hit((0, 0, 0), (0, 0, -4))["t"] == pytest.approx(0.125)
    assert o.hit((0, 0, 0), (0, 1, 0)) is None


def test_sphere_interval_is_open_quad_interval_is_closed():
    """sphere.rs:84-88 uses surrounds, quad.rs:72 uses contains (interval.rs:32-38)."""
    _, o, _ = one_sphere()
    assert o.hit((0, 0, 0), (0, 0, -1), tmin=0.5)["t"] == pytest.approx(1.5)   # 0.5 not inside (0.5, inf)
    b = R.DescBuilder()
    m = b.lambertian((1, 1, 1))
    q = b.quad((-1, -1, -2), (2, 0, 0), (0, 2, 0), m)
    oq = O.OracleScene(b.desc(b.list([q])))
    assert oq.hit((0, 0, 0), (0, 0, -1), tmin=2.0)["t"] == 2.0                  # 2.0 inside [2.0, inf]
    assert oq.hit((0, 0, 0), (0, 0, -1), tmin=0.001, tmax=2.0)["t"] == 2.0


def test_quad_corners_and_edges_are_inside():
    """quad.rs:51-60: alpha, beta tested on the closed unit interval; uv = (alpha, beta)."""
    b = R.DescBuilder()
    m = b.lambertian((1, 1, 1))
    q = b.quad((0, 0, -1), (1, 0, 0), (0, 1, 0), m)
    o = O.OracleScene(b.desc(b.list([q])))
    for x, y in [(0, 0), (1, 0), (0, 1), (1, 1), (0.5, 0), (0.25, 0.75)]:
        h = o.hit((x, y, 0), (0, 0, -1))
        assert h is not None and (h["u"], h["v"]) == (pytest.approx(x), pytest.approx(y))
    assert o.hit((1.0000001, 0.5, 0), (0, 0, -1)) is None
    assert o.hit((0.5, 0.5, 0), (1, 0, 0)) is None  # parallel: |denom| < 1e-8 (quad.rs:67)


def test_equal_t_tie_rules():
    """hittable_list.rs:46-51 with the interval asymmetry: a later quad replaces an equal-t hit,
    a later sphere does not."""
    b = R.DescBuilder()
    m = b.lambertian((1, 1, 1))
    q1 = b.quad((-1, -1, -2), (2, 0, 0), (0, 2, 0), m)
    q2 = b.quad((-1, -1, -2), (2, 0, 0), (0, 2, 0), m)
    assert O.OracleScene(b.desc(b.list([q1, q2]))).hit((0, 0, 0), (0, 0, -1))["id"] == q2
    s1 = b.sphere((0, 0, -3), 1.0, m)
    s2 = b.sphere((0, 0, -3), 1.0, m)
    assert O.OracleScene(b.desc(b.list([s1, s2]))).hit((0, 0, 0), (0, 0, -1))["id"] == s1
    # quad listed after a sphere touching it at the same t wins; sphere listed after the quad loses
    assert O.OracleScene(b.desc(b.list([s1, q1]))).hit((0, 0, 0), (0, 0, -1))["id"] == q1
    assert O.OracleScene(b.desc(b.list([q1, s1]))).hit((0, 0, 0), (0, 0, -1))["id"] == q1


def test_dielectric_reflectance_and_refract():
    """material.rs:107-111 Schlick; vec3.rs:205-210 refract."""
    assert O.reflectance(1.0, 1 / 1.5) == pytest.approx(0.04, abs=1e-15)
    assert O.reflectance(0.0, 1.5) == pytest.approx(1.0)
    r = O.refract((0, 0, -1), (0, 0, 1), 1 / 1.5)     # normal incidence goes straight
    assert r == pytest.approx([0, 0, -1])
    s = math.sin(math.radians(30))
    r = O.refract((s, 0, -math.cos(math.radians(30))), (0, 0, 1), 1 / 1.5)
    assert r[0] == pytest.approx(s / 1.5) and np.linalg.norm(r) == pytest.approx(1.0)


def test_checker_parity_uses_floor_and_rust_remainder():
    """texture.rs:63-77: floor then % 2 == 0; Rust's -1 % 2 = -1 is odd."""
    b = R.DescBuilder()
    ck = b.checker(1.0, b.solid((1, 0, 0)), b.solid((0, 0, 1)))
    m = b.material(F.M_LAMBERTIAN, ck)
    o = O.OracleScene(b.desc(b.list([b.sphere((0, 0, 0), 1, m)])))
    even, odd = [1, 0, 0], [0, 0, 1]
    assert o.texture(ck, 0, 0, (0.5, 0.5, 0.5)) == even
    assert o.texture(ck, 0, 0, (1.5, 0.5, 0.5)) == odd
    assert o.texture(ck, 0, 0, (-0.5, 0.5, 0.5)) == odd     # floor(-0.5) = -1
    assert o.texture(ck, 0, 0, (-0.5, -0.5, 0.5)) == even   # -2
    assert o.texture(ck, 0, 0, (-0.5, -0.5, -0.5)) == odd   # -3


def test_image_texture_lookup():
    """texture.rs:88-109, rtw_image.rs:70-78: v flipped, u=1 clamps to the last column, /255, no gamma."""
    img = np.zeros((2, 4, 3), dtype=np.uint8)
    img[0, :, 0] = [10, 20, 30, 40]      # top row
    img[1, :, 1] = [50, 60, 70, 80]      # bottom row
    b = R.DescBuilder()
    t = b.image(img)
    m = b.material(F.M_LAMBERTIAN, t)
    o = O.OracleScene(b.desc(b.list([b.sphere((0, 0, 0), 1, m)])))
    assert o.texture(t, 0.0, 1.0, (0, 0, 0)) == pytest.approx([10 / 255, 0, 0])     # v=1 -> top row
    assert o.texture(t, 1.0, 1.0, (0, 0, 0)) == pytest.approx([40 / 255, 0, 0])     # i = width -> clamp
    assert o.texture(t, 0.30, 0.0, (0, 0, 0)) == pytest.approx([0, 60 / 255, 0])    # v=0 -> bottom row (clamped)
    assert o.texture(t, -3.0, 7.0, (0, 0, 0)) == pytest.approx([10 / 255, 0, 0])    # clamp u, v
    b2 = R.DescBuilder()
    t2 = b2.image(None)                   # missing file -> cyan (texture.rs:91-93)
    m2 = b2.material(F.M_LAMBERTIAN, t2)
    o2 = O.OracleScene(b2.desc(b2.list([b2.sphere((0, 0, 0), 1, m2)])))
    assert o2.texture(t2, 0.5, 0.5, (0, 0, 0)) == [0, 1, 1]


def test_write_color_quantisation():
    """color.rs:6-42: 0->0, 1.0->255, 0.25->128, NaN->0 (book 3), negative->0."""
    rgb = np.array([[[0.0, 1.0, 0.25], [np.nan, -1.0, 4.0], [0.999 ** 2, 1e-12, 0.5]]])
    out = O.write_color(rgb, True)
    assert out[0, 0].tolist() == [0, 255, 128]
    assert out[0, 1].tolist() == [0, 0, 255]
    assert out[0, 2].tolist() == [255, 0, int(256 * math.sqrt(0.5))]


def test_resolve_accum_matches_render_io():
    """src/render_io.rs:8-26: scale 1/spp, non-finite -> 0, sqrt, clamp .999, x256, `as u8`."""
    acc = np.array([[[100.0, 25.0, 0.0, 100.0], [np.inf, np.nan, -3.0, 100.0]]], dtype=np.float32)
    out = O.resolve_accum(acc, 100)
    assert out[0, 0].tolist() == [255, 128, 0] and out[0, 1].tolist() == [0, 0, 0]
