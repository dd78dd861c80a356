"""The north star's converged-render gate as written, against COMMITTED oracle images
(tests/golden/converged_<scene>.npz, made by scripts/make_golden_converged.py): every book scene,
    PSNR >= 40 dB on the 8-bit gamma image                           (no "or the noise floor"),
    per-channel frame-mean radiance within 0.5 %                      (or 3 standard errors if larger),
    bias: z = (gpu - oracle) / s.e. per 4x4-pixel block has |mean| < 0.3 and rms < 1.3 (tests/parity_stats.py);
    per pixel, fewer than 6e-3 of the |z| exceed 4 (measured 0.3-2.7e-3: path-traced pixels are skewed
    distributions, a Gaussian would give 6e-5; a wrong object shows up as block rms long before this).
The golden images hold the oracle's per-pixel mean and per-sample variance at an spp where two
independent oracle renders agree to >= 42 dB (converged_manifest.json: floor_psnr); the GPU renders
the golden's sample count ("the same high spp"), which puts the comparison 3 dB above that floor.  Under the null
hypothesis (the GPU samples the same per-pixel distribution) the variance of a GPU sample equals the
stored oracle variance, which gives the standard errors without a second-moment buffer on the GPU.

CPU tier: the fixtures are complete and self-consistent, and today's oracle still reproduces them
(a short oracle render must pass the same z test), so an oracle regression cannot hide behind them.
The achieved figures are written to gpurun_out/parity_r02.json (copied to profiles/ per round)."""
import json
import os

import numpy as np
import pytest

import rustraytrace_b200 as R
from parity_stats import z_stats

F = R._ffi
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
MANIFEST = json.load(open(os.path.join(GOLD, "converged_manifest.json")))
SCENES = sorted(MANIFEST)
GPU_SPP_FACTOR = 1


def to8(lin, nan_scrub):
    """write_color (color.rs:14-42): NaN -> 0 (book 3), sqrt gamma, clamp [0, 0.999], truncate 256 x."""
    x = np.array(lin, dtype=np.float64)
    if nan_scrub:
        x = np.where(np.isnan(x), 0.0, x)
    g = np.where(x > 0, np.sqrt(np.maximum(x, 0)), 0.0)
    return (256.0 * np.clip(g, 0.0, 0.999)).astype(np.int32)


def psnr8(a, b):
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return 99.0 if mse == 0 else 10.0 * np.log10(255.0 ** 2 / mse)


def load(key):
    z = np.load(os.path.join(GOLD, f"converged_{key}.npz"))
    return z["mean"].astype(np.float64), z["var"].astype(np.float64)


def build(key, spp):
    m = MANIFEST[key]
    return R.BookScene(m["book"], m["scene"], overrides=dict(image_width=m["width"], samples_per_pixel=spp, max_depth=m["depth"]))


def record(key, entry):
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    path = os.path.join(out, "parity_r02.json")
    data = json.load(open(path)) if os.path.exists(path) else {}
    data.setdefault("converged_golden", {})[key] = entry
    with open(path, "w") as f:
        json.dump(data, f, indent=1, sort_keys=True)


# ---- CPU tier --------------------------------------------------------------------------------------
def test_fixtures_cover_every_book_scene_and_are_converged():
    want = {"in_one_weekend", "in_one_weekend_gpu", "the_rest_of_your_life"} | {f"the_next_week_{i}" for i in range(1, 10)}
    assert set(SCENES) == want
    for key in SCENES:
        m = MANIFEST[key]
        mean, var = load(key)
        assert mean.shape == (m["height"], m["width"], 3) and var.shape == mean.shape
        assert m["floor_psnr"] >= 42.0, f"{key}: two oracle renders agree to {m['floor_psnr']} dB only"
        assert m["nan_pixels"] == 0 and np.isfinite(mean).all() and (var >= 0).all()
        assert max(np.array(m["frame_mean_se"]) / np.array(m["frame_mean"])) < 1e-3   # frame means known to 0.1 %


@pytest.mark.parametrize("key", SCENES)
def test_oracle_still_reproduces_the_converged_fixture(key):
    import oracle_ffi as O
    m = MANIFEST[key]
    spp = 64
    s = build(key, spp)
    img, _ = O.OracleScene(s.desc).render(s.camera, s.integrator, seed=4242)
    gold, var = load(key)
    st = z_stats(img, spp, gold, var, m["n"])
    # 64 samples of a skewed distribution are not Gaussian: the tail bound is loose, the mean is not
    assert abs(st["block_mean_z"]) < 0.3 and st["frac_over_4"] < 2e-2 and 0.5 < st["block_rms_z"] < 1.5, st


# ---- GPU tier --------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def ctx():
    c = R.Context(0)
    yield c
    c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("key", SCENES)
def test_gpu_converged_render_meets_the_north_star_gate(ctx, key):
    m = MANIFEST[key]
    gold, var = load(key)
    b3 = m["book"] == "the_rest_of_your_life"
    n_gpu = GPU_SPP_FACTOR * m["n"]
    s = build(key, n_gpu)
    gs = ctx.scene(s.desc)
    acc = ctx.render(gs, s.camera, s.integrator, seed=77)
    st = ctx.stats()
    n_gpu = st["spp_effective"]
    assert np.all(acc[..., 3] == n_gpu)
    lin = acc[..., :3].astype(np.float64) / n_gpu
    img_gpu = ctx.resolve(acc, n_gpu).astype(np.int32)
    gs.close()
    assert np.array_equal(img_gpu, to8(lin, b3)) or np.abs(img_gpu - to8(lin, b3)).max() <= 1   # f32 resolve vs f64 here
    psnr = psnr8(img_gpu, to8(gold, b3))
    gm, fm = np.array(m["frame_mean"]), lin.mean(axis=(0, 1))
    se = np.sqrt(np.array(m["frame_mean_se"]) ** 2 * (1.0 + m["n"] / n_gpu))
    rel = np.abs(fm - gm) / gm
    zs = z_stats(lin, n_gpu, gold, var, m["n"])
    record(key, dict(width=m["width"], height=m["height"], golden_n=m["n"], gpu_n=int(n_gpu), golden_floor_psnr=m["floor_psnr"],
                     psnr=round(float(psnr), 2), frame_mean_gpu=[float(x) for x in fm], frame_mean_golden=[float(x) for x in gm],
                     frame_mean_rel_diff=[float(x) for x in rel], frame_mean_rel_se=[float(x) for x in se / gm], **zs))
    assert psnr >= 40.0, f"PSNR {psnr:.2f} dB (golden floor {m['floor_psnr']} dB)"
    assert np.all(rel <= np.maximum(0.005, 3.0 * se / gm)), f"frame mean off by {rel} (s.e. {se / gm})"
    assert abs(zs["block_mean_z"]) < 0.3 and zs["block_rms_z"] < 1.3 and zs["flat_max_rel"] < 1e-2, zs
    assert zs["frac_over_4"] < 6e-3, zs
