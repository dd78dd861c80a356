"""Pins for the code only the book-3 integrator runs and for
the sampling RNG, on both sides of the parity gate.

CPU tier (the oracle and the host-side hooks):
  * light pdfs against their closed forms — quad dist^2 / (cos * area) (quad.rs:93-102), sphere
    1 / (2 pi (1 - cos theta_max)) (sphere.rs:102-115), list = mean of its members
    (hittable_list.rs:60-69);
  * the white furnace (tests/b3_scenes.py) under the book-2 and book-3 integrators: radiance E;
  * Philox4x32 at 7 rounds — the round count the kernel runs — against Random123's kat_vectors.
GPU tier: the same vectors pushed through the DEVICE code (rrt_kat_light, rrt_kat_rng), the furnace on
the GPU, and random book-3 scenes (isotropic media, metal, textures, moving spheres, two lights)
against the oracle with per-pixel standard errors from the oracle's second moments."""
import json
import math
import os

import numpy as np
import pytest

import b3_scenes as B3
from parity_stats import z_stats
import rustraytrace_b200 as R

F = R._ffi
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# Random123 kat_vectors (philox4x32, rounds, counter[4], key[2] -> output[4])
PHILOX_KAT = [
    (7, (0, 0, 0, 0), (0, 0), (0x5f6fb709, 0x0d893f64, 0x4f121f81, 0x4f730a48)),
    (7, (0xffffffff,) * 4, (0xffffffff,) * 2, (0x5207ddc2, 0x45165e59, 0x4d8ee751, 0x8c52f662)),
    (7, (0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0x4dfccaba, 0x190a87f0, 0xc47362ba, 0xb6b5242a)),
    (10, (0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    (10, (0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    (10, (0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
]


def light_scene():
    """A quad light, a sphere light and the list of both, over an otherwise empty world."""
    b = R.DescBuilder()
    m = b.material(F.M_EMPTY)
    q = b.quad((343, 554, 332), (-130, 0, 0), (0, 0, -105), m)      # the_rest_of_your_life/mod.rs:130-135
    s = b.sphere((190, 90, 190), 90, m)                               # mod.rs:136-140
    floor = b.quad((0, 0, 0), (555, 0, 0), (0, 0, 555), b.lambertian((0.7, 0.7, 0.7)))
    return b, q, s, floor


def quad_pdf_closed_form(o, d):
    """dist^2 / (|cos| area) for the quad y = 554, x in [213, 343], z in [227, 332]; 0 on a miss."""
    if abs(d[1]) < 1e-8:
        return 0.0
    t = (554.0 - o[1]) / d[1]
    if t < 0.001:
        return 0.0
    x, z = o[0] + t * d[0], o[2] + t * d[2]
    if not (213.0 <= x <= 343.0 and 227.0 <= z <= 332.0):
        return 0.0
    L = math.sqrt(sum(c * c for c in d))
    return (t * L) ** 2 / (abs(d[1]) / L * 130.0 * 105.0)


def sphere_pdf_closed_form(o, d):
    c, r = np.array([190.0, 90.0, 190.0]), 90.0
    oc = c - np.array(o)
    dd = np.array(d, dtype=np.float64)
    a, h = dd @ dd, dd @ oc
    disc = h * h - a * (oc @ oc - r * r)
    if disc < 0:
        return 0.0
    roots = [(h - math.sqrt(disc)) / a, (h + math.sqrt(disc)) / a]
    if not any(x > 0.001 for x in roots):
        return 0.0
    cos_max = math.sqrt(1.0 - r * r / (oc @ oc))
    return 1.0 / (2.0 * math.pi * (1.0 - cos_max))


def light_queries(n, seed=1):
    """Origins on the floor and in mid-air, directions toward the lights (hits) and random (mostly misses)."""
    rng = np.random.default_rng(seed)
    o = np.column_stack([rng.uniform(20, 535, n), rng.uniform(0, 300, n) * (rng.random(n) < 0.5), rng.uniform(20, 535, n)])
    o[:, 1] = np.where(np.linalg.norm(o - np.array([190, 90, 190.0]), axis=1) < 100, 400.0, o[:, 1])   # stay outside the sphere
    target = np.where((rng.random(n) < 0.5)[:, None],
                      np.column_stack([rng.uniform(200, 356, n), np.full(n, 554.0), rng.uniform(215, 345, n)]),
                      np.array([190, 90, 190.0]) + rng.normal(0, 60, (n, 3)))
    d = np.where((rng.random(n) < 0.85)[:, None], target - o, rng.normal(0, 1, (n, 3)))
    d *= rng.uniform(0.3, 3.0, n)[:, None]       # pdf_value must not depend on |d|
    return o, d


# ---- CPU tier ------------------------------------------------------------------------------------------
def test_philox_known_answers_at_the_round_count_the_kernel_runs():
    assert F.rrt().rrt_philox_rounds() == 7
    out = (F.C.c_uint32 * 4)()
    for rounds, c, k, want in PHILOX_KAT:
        assert F.rrt().rrt_philox_kat_rounds(rounds, *c, *k, out) == 0
        assert tuple(out) == want, (rounds, [hex(x) for x in out])
    assert F.rrt().rrt_philox_kat_rounds(9, 0, 0, 0, 0, 0, 0, out) == F.ERR_INVALID


def test_oracle_light_pdfs_match_their_closed_forms():
    import oracle_ffi as O
    b, q, s, floor = light_scene()
    oq = O.OracleScene(b.desc(b.list([floor]), b.list([q])))
    os_ = O.OracleScene(b.desc(b.list([floor]), b.list([s])))
    ol = O.OracleScene(b.desc(b.list([floor]), b.list([q, s])))
    o, d = light_queries(400)
    hits_q = hits_s = 0
    for i in range(len(o)):
        pq, ps = quad_pdf_closed_form(o[i], d[i]), sphere_pdf_closed_form(o[i], d[i])
        hits_q += pq > 0; hits_s += ps > 0
        assert oq.light_pdf(o[i], d[i]) == pytest.approx(pq, rel=1e-10, abs=1e-300)
        assert os_.light_pdf(o[i], d[i]) == pytest.approx(ps, rel=1e-10, abs=1e-300)
        assert ol.light_pdf(o[i], d[i]) == pytest.approx(0.5 * (pq + ps), rel=1e-10, abs=1e-300)   # hittable_list.rs:60-69
    assert hits_q > 60 and hits_s > 60
    # the book's own numbers: straight below the centre of the light, 554 - 0 away: dist^2 / area
    assert oq.light_pdf((278, 0, 279.5), (0, 2, 0)) == pytest.approx(554.0 ** 2 / (130 * 105), rel=1e-12)
    assert os_.light_pdf((190, 90, 390), (0, 0, -1)) == pytest.approx(1 / (2 * math.pi * (1 - math.sqrt(1 - 0.45 ** 2))), rel=1e-12)


@pytest.mark.parametrize("book3", [False, True])
def test_oracle_white_furnace(book3):
    import oracle_ffi as O
    b, d, cam = B3.furnace(book3, width=32, spp=100, depth=100)
    img, sq = O.OracleScene(d).render_moments(cam, F.INT_BOOK3 if book3 else F.INT_BOOK2, seed=31)
    assert np.isfinite(img).all()
    n = img.shape[0] * img.shape[1] * 100
    for k in range(3):
        se = math.sqrt(max(1e-30, float((sq[..., k] - img[..., k] ** 2).mean()) / n))
        assert abs(img[..., k].mean() - B3.E[k]) <= max(4 * se, 1e-9), (k, img[..., k].mean(), B3.E[k], se)
        assert abs(img[..., k].mean() / B3.E[k] - 1) < 0.01


# ---- GPU tier ------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def ctx():
    c = R.Context(0)
    yield c
    c.close()


def record(section, key, entry):
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    path = os.path.join(out, "parity_r02.json")
    data = json.load(open(path)) if os.path.exists(path) else {}
    data.setdefault(section, {})[key] = entry
    with open(path, "w") as f:
        json.dump(data, f, indent=1, sort_keys=True)


@pytest.mark.gpu
def test_device_philox_is_the_known_answer_function(ctx):
    for rounds, c, k, want in PHILOX_KAT:
        if rounds != F.rrt().rrt_philox_rounds():
            continue
        raw, _ = ctx.kat_rng(k[0] | (k[1] << 32), c[0], c[1], c[2], c[3], 1)
        assert tuple(int(x) for x in raw[0]) == want


@pytest.mark.gpu
def test_device_uniforms_are_uniform_and_the_fifth_is_independent(ctx):
    """Rand4.x..w are the top 24 bits of the four Philox words, .v the low bytes of words 0-2
    (rng.cuh): chi-square on 256 bins each, on 16x16 bins for the pairs (x, v), (y, v), (z, v), and no
    linear correlation; also along the bounce axis, which is how a path walks the counter space."""
    n = 1 << 20
    sets = [ctx.kat_rng(0x1234ABCD, 777, 0, 3, 0, n)[1].astype(np.float64)]
    sets.append(np.concatenate([ctx.kat_rng(99, 5, s0, bnc, 0, 1 << 14)[1] for s0 in (0, 1 << 20) for bnc in range(32)]).astype(np.float64))
    stats = {}
    for name, u in zip(("samples", "bounces"), sets):
        m = len(u)
        assert (u >= 0).all() and (u < 1).all()
        for k, nm in enumerate("xyzwv"):
            cnt = np.bincount((u[:, k] * 256).astype(int), minlength=256)
            chi = float(((cnt - m / 256) ** 2 / (m / 256)).sum())
            stats[f"{name}_chi2_{nm}"] = round(chi, 1)
            assert chi < 255 + 6 * math.sqrt(2 * 255), (name, nm, chi)
            assert abs(u[:, k].mean() - 0.5) < 5 * math.sqrt(1 / 12 / m)
        for k in range(3):
            cell = (u[:, k] * 16).astype(int) * 16 + (u[:, 4] * 16).astype(int)
            cnt = np.bincount(cell, minlength=256)
            chi = float(((cnt - m / 256) ** 2 / (m / 256)).sum())
            stats[f"{name}_chi2_{'xyz'[k]}v"] = round(chi, 1)
            assert chi < 255 + 6 * math.sqrt(2 * 255), (name, k, chi)
            assert abs(np.corrcoef(u[:, k], u[:, 4])[0, 1]) < 5 / math.sqrt(m)
    record("rng", "chi2_255_dof", stats)


@pytest.mark.gpu
def test_device_light_pdf_and_random_match_closed_forms_and_the_oracle(ctx):
    import oracle_ffi as O
    b, q, s, floor = light_scene()
    d_ = b.desc(b.list([floor]), b.list([q, s]))
    gs = ctx.scene(d_)
    ol = O.OracleScene(d_)
    o, d = light_queries(4000, seed=2)
    rng = np.random.default_rng(3)
    u = rng.random((len(o), 2))
    each, lst, rnd = ctx.kat_light(gs, o, d, u)
    want_q = np.array([quad_pdf_closed_form(o[i], d[i]) for i in range(len(o))])
    want_s = np.array([sphere_pdf_closed_form(o[i], d[i]) for i in range(len(o))])
    want_l = np.array([ol.light_pdf(o[i], d[i]) for i in range(len(o))])
    # f32: a direction that grazes the edge of a light can land on the other side of it; everywhere else 1e-4
    for got, want, name in ((each[:, 0], want_q, "quad"), (each[:, 1], want_s, "sphere"), (lst, want_l, "list")):
        flipped = (got > 0) != (want > 0)
        assert flipped.mean() < 2e-3, (name, flipped.mean())
        ok = ~flipped & (want > 0)
        rel = np.abs(got[ok] - want[ok]) / want[ok]
        assert ok.sum() > 500 and rel.max() < 2e-4, (name, rel.max())
        record("light_pdf", name, dict(queries=int(len(o)), hits=int(ok.sum()), edge_flips=int(flipped.sum()), max_rel_err=float(rel.max())))
    # random(): the quad's points are Q + u1 u + u2 v (quad.rs:104-107); the sphere's directions lie inside
    # the cone of half-angle theta_max about the centre direction, uniformly in cos theta (sphere.rs:117-122)
    p = o + rnd[:, 0, :]
    want_p = np.array([343, 554, 332.0]) + u[:, :1] * np.array([-130.0, 0, 0]) + u[:, 1:2] * np.array([0, 0, -105.0])
    assert np.abs(p - want_p).max() < 2e-3
    c = np.array([190, 90, 190.0]) - o
    dist = np.linalg.norm(c, axis=1)
    cos_max = np.sqrt(1 - 90.0 ** 2 / dist ** 2)
    dirs = rnd[:, 1, :] / np.linalg.norm(rnd[:, 1, :], axis=1, keepdims=True)
    cos = (dirs * c).sum(axis=1) / dist
    assert (cos >= cos_max - 1e-5).all()
    z = (1 - cos) / (1 - cos_max)                      # uniform on [0, 1] when cos is uniform on [cos_max, 1]
    cnt = np.bincount(np.clip((z * 16).astype(int), 0, 15), minlength=16)
    assert ((cnt - len(o) / 16) ** 2 / (len(o) / 16)).sum() < 15 + 6 * math.sqrt(30)
    # and a sampled direction has exactly the pdf the closed form gives it
    each2, _, _ = ctx.kat_light(gs, o, rnd[:, 1, :], u)
    assert np.allclose(each2[:, 1], 1 / (2 * math.pi * (1 - cos_max)), rtol=2e-4)
    gs.close()


@pytest.mark.gpu
@pytest.mark.parametrize("book3", [False, True])
def test_gpu_white_furnace(ctx, book3):
    b, d, cam = B3.furnace(book3, width=64, spp=4096, depth=100)
    gs = ctx.scene(d)
    integ = F.INT_BOOK3 if book3 else F.INT_BOOK2
    acc = ctx.render(gs, cam, integ, seed=13).astype(np.float64)
    assert ctx.stats()["kernel_features"] & F.SF_MEDIA
    gs.close()
    lin = acc[..., :3] / acc[..., 3:4]
    assert np.isfinite(lin).all()
    ent = {}
    for k in range(3):
        px = lin[..., k]
        se = px.std() / math.sqrt(px.size)           # every pixel has the same expectation E
        ent["rgb"[k]] = dict(mean=float(px.mean()), expected=B3.E[k], rel_dev=float(px.mean() / B3.E[k] - 1), rel_se=float(se / B3.E[k]))
        assert abs(px.mean() - B3.E[k]) <= max(4 * se, 2e-4 * B3.E[k]), ent
    record("white_furnace", "book3" if book3 else "book2", ent)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(8))
def test_random_book3_scenes_match_the_oracle(ctx, seed):
    """a31 (Isotropic under the book-3 integrator: SpherePdf, 1/4pi), b3 metal, media, textures and moving
    spheres with a two-entry lights list: GPU against oracle, per pixel and in the frame mean."""
    import oracle_ffi as O
    spp_o, spp_g = 1024, 1024
    b, d, cam = B3.random_b3(seed, width=48, spp=spp_o)
    img, sq = O.OracleScene(d).render_moments(cam, F.INT_BOOK3, seed=600 + seed)
    var = np.clip(sq - img * img, 0, None)
    gs = ctx.scene(d)
    feat = R.scene_features(d)
    assert feat & F.SF_MEDIA and feat & F.SF_QUADS and feat & F.SF_TEXTURES
    acc = ctx.render(gs, R.copy_camera(cam, samples_per_pixel=spp_g), F.INT_BOOK3, seed=seed).astype(np.float64)
    gs.close()
    lin = acc[..., :3] / acc[..., 3:4]
    ok = np.isfinite(img).all(axis=2)
    # A NaN sample poisons an oracle pixel as it does a reference pixel (camera.rs:79-83 sums every sample;
    # the GPU drops non-finite samples, include/rrt.h): such pixels — a surface inside the sphere light,
    # where cos_theta_max is NaN, sphere.rs:110 — are left out of the comparison.
    assert ok.mean() > 0.95
    zs = z_stats(np.where(ok[..., None], lin, np.nan), spp_g, img, var, spp_o)
    fm_g, fm_o = lin[ok].mean(axis=0), img[ok].mean(axis=0)
    fse = np.sqrt((var[ok] * (1 / spp_o + 1 / spp_g)).sum(axis=0)) / ok.sum()
    rel = np.abs(fm_g - fm_o) / fm_o
    record("book3_fuzz", str(seed), dict(features=int(feat), frame_mean_rel_diff=[float(x) for x in rel],
                                         frame_mean_rel_se=[float(x) for x in fse / fm_o], **zs))
    assert np.all(rel <= np.maximum(0.005, 4 * fse / fm_o)), (rel, fse / fm_o)
    assert abs(zs["block_mean_z"]) < 0.35 and zs["block_rms_z"] < 1.3 and zs["frac_over_4"] < 2.5e-2 and zs["flat_max_rel"] < 5e-2, zs   # 1 024 spp: one light path in a flat-looking pixel is 2 % of the frame mean


@pytest.mark.gpu
def test_non_finite_samples_are_dropped_not_spread_over_the_pixel(ctx):
    """The documented deviation (include/rrt.h, "non-finite samples"): a floor that runs through a sphere
    light puts scattering points INSIDE the light, where sphere.rs:110 takes the square root of a negative
    number.  The reference sums the NaN into the pixel and color.rs:19-27 writes the whole pixel black; the
    GPU path ends such a path where the pdf stops being a positive number, keeps every finite sample and
    still counts all of them in .w — the oracle (= the reference) shows the poisoned pixels, the GPU none."""
    import oracle_ffi as O
    b = R.DescBuilder()
    grey = b.lambertian((0.6, 0.6, 0.6))
    lightm = b.material(F.M_DIFFUSE_LIGHT, b.solid((4, 4, 4)))
    floor = b.quad((-5, 0, -5), (10, 0, 0), (0, 0, 10), grey)
    lamp = b.sphere((0, 0.5, 0), 1.5, lightm)                      # the floor cuts through it
    world = b.list([floor, lamp])
    d = b.desc(world, b.list([b.sphere((0, 0.5, 0), 1.5, b.material(F.M_EMPTY))]))
    cam = R.make_camera(image_width=64, aspect_ratio=1.0, samples_per_pixel=64, max_depth=10, background=(0.1, 0.1, 0.1),
                        vfov=60.0, lookfrom=(0, 1.0, 0.2), lookat=(0, 0, 0.1))   # camera inside the light, looking at the floor
    gs = ctx.scene(d)
    acc = ctx.render(gs, cam, F.INT_BOOK3, seed=1)
    gs.close()
    assert np.isfinite(acc).all() and np.all(acc[..., 3] == 64) and (acc[..., :3] >= 0).all()
    img, _ = O.OracleScene(d).render(cam, F.INT_BOOK3, seed=1)
    assert np.isnan(img).any(axis=2).mean() > 0.2                   # the reference poisons these pixels
    assert np.array_equal(O.write_color(img, True)[np.isnan(img).any(axis=2)].max(initial=0), 0)
