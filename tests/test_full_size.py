"""GPU tier at BASELINE.json's full sizes (configs C1..C5 of rustraytrace_b200/configs.py).

Probe parity is checked on every pixel of the full frame (the oracle's probe is cheap).  Converged
renders at full size are checked through size-independent properties: every pixel received exactly
its samples, the frame is finite, tile- and sample-sharded shares sum to the unsharded frame, the
frame mean agrees to 0.5 % with the oracle's converged mean for that scene (tests/golden/converged_manifest.json;
the frame mean is the viewport integral, independent of the pixel grid), and — C1, which the
reference CPU renderer can run in full — PSNR against the full oracle render.  Full spp is used
where it costs < 1 s of GPU time, otherwise a stated fraction of it (sampling is per-sample iid)."""
import json
import os

import numpy as np
import pytest

import oracle_ffi as O
import rustraytrace_b200 as R
from rustraytrace_b200 import configs

F = R._ffi
pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLD_NAME = {"C1": "in_one_weekend", "C2": "in_one_weekend", "C3a": "the_next_week_1", "C3b": "the_next_week_3",
             "C3c": "the_next_week_4", "C4": "the_next_week_9", "C5": "the_rest_of_your_life"}
# spp used at full resolution (full where cheap; C4/C5 full spp is seconds to tens of seconds of GPU)
SPP = {"C1": 100, "C2": 500, "C3a": 250, "C3b": 1000, "C3c": 250, "C4": 400, "C5": 64}


def _record(section, key, entry):
    out = os.path.join(os.path.dirname(GOLD), "..", "gpurun_out")
    os.makedirs(out, exist_ok=True)
    path = os.path.join(out, "parity_r02.json")
    data = json.load(open(path)) if os.path.exists(path) else {}
    data.setdefault(section, {})[key] = entry
    with open(path, "w") as f:
        json.dump(data, f, indent=1, sort_keys=True)


@pytest.fixture(scope="module")
def ctx():
    c = R.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("key", list(configs.CONFIGS))
def test_full_size_probe_ids_bit_exact(ctx, key):
    s = configs.build(key)
    gs = ctx.scene(s.desc)
    ids, t = ctx.probe(gs, s.camera, 64)
    oids, ot = O.OracleScene(s.desc).probe(s.camera)
    assert ids.shape == (s.height, s.width)
    assert np.array_equal(ids, oids), f"{int((ids != oids).sum())} of {ids.size} primary-hit ids differ"
    hit = oids >= 0
    assert np.max(np.abs(t[hit] - ot[hit]) / ot[hit]) <= 1e-5
    gs.close()


def _silhouette(ids):
    e = np.zeros(ids.shape, dtype=bool)
    e[:, 1:] |= ids[:, 1:] != ids[:, :-1]; e[:, :-1] |= ids[:, 1:] != ids[:, :-1]
    e[1:, :] |= ids[1:, :] != ids[:-1, :]; e[:-1, :] |= ids[1:, :] != ids[:-1, :]
    return e


@pytest.mark.parametrize("key", list(configs.CONFIGS) + ["C4_no_box_leaves", "C2_no_grid", "C3a_no_grid", "C4_no_grid"])
def test_full_size_f32_hot_path_primary_hits(ctx, key, monkeypatch):
    """The f32 intersection code that traces every sample (sphere_hit32, quad_hit32, box_leaf_hit32, slab2,
    --use_fast_math) against the oracle's f64 probe, every pixel of every config: gated at twice the
    figures measured on a B200 and committed in tests/golden/probe_f32.json (scripts/probe_f32_stats.py).
    Ids may differ on silhouette pixels only; t errors above 1e-4 belong to the r = 1000 ground sphere."""
    cfg = key.split("_")[0]
    golds = json.load(open(os.path.join(GOLD, "probe_f32.json")))
    gold = golds.get(key, golds[cfg])   # (the figures were taken on the tree walk; the grid walk runs the same primitive tests)
    s = configs.build(cfg)
    if key.endswith("_no_box_leaves"):
        monkeypatch.setenv("RRT_NO_BOX_LEAVES", "1")
    if key.endswith("_no_grid"):        # C1, C2, C3a, C4 are walked through grids by default: the tree walk as well
        monkeypatch.setenv("RRT_NO_GRID", "1")
    gs = ctx.scene(s.desc)
    assert (ctx.stats()["n_grid_cells"] > 0) == (key in ("C1", "C2", "C3a", "C4"))   # (without box leaves C4 has no box field)
    ids, t = ctx.probe(gs, s.camera, 32)
    gs.close()
    oids, ot = O.OracleScene(s.desc).probe(s.camera)
    diff = ids != oids
    assert int(diff.sum()) <= 2 * gold["ids_differ"] + 2, (int(diff.sum()), gold["ids_differ"])
    assert not (diff & ~_silhouette(oids)).any(), "an f32 id differs away from a silhouette"
    same = ~diff & (oids >= 0)
    rel = np.abs(t[same] - ot[same]) / ot[same]
    assert rel.max() <= 2 * gold["t_max_rel"] and np.quantile(rel, 0.99) <= 2 * gold["t_p99_rel"], (rel.max(), gold)
    big = np.zeros(ids.shape, dtype=bool)
    big[same] = rel > 1e-4
    if big.any():   # only the outsized spheres (r = 1000 ground) lose more than 1e-4 to f32 cancellation
        assert set(np.unique(oids[big]).tolist()) <= {gold["worst_id"]}, np.unique(oids[big])


@pytest.mark.parametrize("key", list(configs.CONFIGS))
def test_full_size_render_properties(ctx, key):
    s = configs.build(key, samples_per_pixel=SPP[key])
    gs = ctx.scene(s.desc)
    acc = ctx.render(gs, s.camera, s.integrator, seed=21)
    st = ctx.stats()
    spp = st["spp_effective"]
    assert st["paths"] == float(s.width) * s.height * spp
    assert np.all(acc[..., 3] == spp), "every pixel must have received exactly its samples"
    assert np.isfinite(acc).all() and (acc[..., :3] >= 0).all()
    mean = acc[..., :3].astype(np.float64).mean(axis=(0, 1)) / spp
    # The frame mean is the integral of the image over the viewport, whatever the pixel grid: the
    # converged oracle fixtures (64-96 px wide, tests/golden/converged_manifest.json, frame means known
    # to ~0.01-0.03 %) are the target at full size too.  North star: within 0.5 %; widened only by
    # 3 standard errors of the two estimates where that is larger (it is not, at these sample counts).
    g = json.load(open(os.path.join(GOLD, "converged_manifest.json")))[GOLD_NAME[key]]
    want, se_o = np.array(g["frame_mean"]), np.array(g["frame_mean_se"])
    se_g = se_o * np.sqrt(g["n"] * g["width"] * g["height"] / (float(spp) * s.width * s.height))   # same per-sample variance
    rel = np.abs(mean - want) / want
    assert np.all(rel <= np.maximum(0.005, 3 * np.sqrt(se_o ** 2 + se_g ** 2) / want)), (mean, want, rel)
    _record("full_size_frame_mean", key, dict(spp=int(spp), gpu=[float(x) for x in mean], golden=[float(x) for x in want],
                                              rel_diff=[float(x) for x in rel]))
    # same seed, same frame, bit for bit (one float4 store per pixel, or commutative float atomics of
    # per-chunk sums whose order may vary: allow the last ulps there)
    acc2 = ctx.render(gs, s.camera, s.integrator, seed=21)
    assert np.allclose(acc, acc2, rtol=1e-6, atol=1e-4)
    gs.close()


@pytest.mark.parametrize("key,mode", [("C4", F.SHARD_SAMPLE), ("C5", F.SHARD_TILE)])
def test_full_size_shares_sum_to_the_frame(ctx, key, mode):
    """The multi-GPU decomposition of the config that BASELINE.json shards, on one GPU: 8 ranks'
    shares, summed on the host, equal the unsharded frame."""
    s = configs.build(key, samples_per_pixel=16)
    gs = ctx.scene(s.desc)
    full = ctx.render(gs, s.camera, s.integrator, seed=4).astype(np.float64)
    tot = np.zeros_like(full)
    for r in range(8):
        tot += ctx.render(gs, s.camera, s.integrator, seed=4, shard_mode=mode, rank=r, world=8)
    assert np.array_equal(tot[..., 3], full[..., 3])
    assert np.allclose(tot[..., :3], full[..., :3], rtol=2e-4, atol=1e-3)
    gs.close()


@pytest.mark.parametrize("key,row_step", [("C2", 1), ("C3b", 1), ("C3c", 2)])
def test_full_size_full_spp_render_against_the_oracle_render(ctx, key, row_step):
    """North-star config 2 as written: 1200x675, 500 spp, depth 50 on one B200, "converged PSNR check"
    against the CPU render at the same spp — and the two 1920x1080 x 1000 spp book-2 configs the oracle
    finishes in under a minute (C3c on every second row).  PSNR >= 40 dB, no noise-floor escape;
    per-channel mean within 0.5 %."""
    s = configs.build(key)
    gs = ctx.scene(s.desc)
    acc = ctx.render(gs, s.camera, s.integrator, seed=1)
    spp = ctx.stats()["spp_effective"]
    gs.close()
    a, _ = O.OracleScene(s.desc).render(s.camera, s.integrator, seed=301, rows=(0, s.height, row_step))
    rows = slice(0, s.height, row_step)
    img = ctx.resolve(acc, spp)[rows].astype(np.float64)
    ia = O.write_color(a[rows], False).astype(np.float64)
    psnr = 10 * np.log10(255.0 ** 2 / np.mean((img - ia) ** 2))
    m_gpu = acc[rows, :, :3].astype(np.float64).mean(axis=(0, 1)) / spp
    m_ref = a[rows].mean(axis=(0, 1))
    rel = np.abs(m_gpu - m_ref) / m_ref
    _record("full_size_vs_oracle", key, dict(width=s.width, height=s.height, spp=int(spp), rows_compared=len(range(0, s.height, row_step)),
                                             psnr=round(float(psnr), 2), mean_rel_diff=[float(x) for x in rel]))
    assert psnr >= 40.0, psnr
    assert np.all(rel < 0.005), rel


def test_c1_full_render_against_full_oracle_render(ctx):
    """C1 is the config the reference CPU renderer runs in full: same size, same spp, same depth."""
    s = configs.build("C1")
    gs = ctx.scene(s.desc)
    acc = ctx.render(gs, s.camera, s.integrator, seed=1)
    o = O.OracleScene(s.desc)
    a, _ = o.render(s.camera, s.integrator, seed=101)
    b, _ = o.render(s.camera, s.integrator, seed=102)
    img = ctx.resolve(acc, 100)
    ia, ib = O.write_color(a, False), O.write_color(b, False)
    mse = lambda x, y: np.mean((x.astype(np.float64) - y.astype(np.float64)) ** 2)
    psnr = lambda x, y: 10 * np.log10(255.0 ** 2 / mse(x, y))
    floor, got = psnr(ia, ib), psnr(img, ia)
    # 100 spp is the reference's own CPU-runnable setting, not a converged one: the GPU frame must be as
    # close to an oracle frame as a second oracle frame is (0.5 dB = 12 % in MSE); the >= 40 dB gates are
    # test_full_size_full_spp_render_against_the_oracle_render and tests/test_converged_golden.py
    assert got >= floor - 0.5, (got, floor)
    m_gpu = acc[..., :3].astype(np.float64).mean(axis=(0, 1)) / 100
    m_ref = 0.5 * (a.mean(axis=(0, 1)) + b.mean(axis=(0, 1)))
    assert np.all(np.abs(m_gpu - m_ref) / m_ref < 0.005)
    gs.close()
