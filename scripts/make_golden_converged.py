"""Writes tests/golden/converged_<scene>.npz: a CONVERGED oracle image per book scene, with the
per-pixel variance of the per-sample radiance, so that the GPU tests can demand the north star's gate
as written — PSNR >= 40 dB, mean radiance within 0.5 % — and a per-pixel bias test, instead of
"within the oracle's own noise".

For every scene a pilot render measures the per-pixel variance and predicts the samples-per-pixel at
which two independent oracle renders agree to >= 43 dB on the 8-bit gamma image; the scene is then
rendered twice at that spp with different seeds (A, B).  Stored: mean = (A + B) / 2 (linear f64 ->
f32), var = pooled per-sample variance per channel, n = samples behind `mean`, floor_psnr =
PSNR(A, B) (two renders of n/2 samples each; `mean` itself is 3 dB better).  A scene whose floor
ends below 42 dB is reported and must be re-run with a larger --max-spp.

Run here (CPU, ~1 min per scene on 8 threads):   python scripts/make_golden_converged.py [scene ...]
The reference itself cannot be run (no Rust toolchain), so these fixtures pin the ORACLE, and give the
GPU tier an oracle-free target (the oracle is not needed on the GPU box for these tests)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import oracle_ffi as O  # noqa: E402
import rustraytrace_b200 as R  # noqa: E402

F = R._ffi
GOLD = os.path.join(ROOT, "tests", "golden")
# scene -> (book, scene number, image width, max_depth)
SCENES = {
    "in_one_weekend": ("in_one_weekend", None, 96, 50),
    "in_one_weekend_gpu": ("in_one_weekend_gpu", None, 96, 50),
    "the_next_week_1": ("the_next_week", 1, 96, 50),
    "the_next_week_2": ("the_next_week", 2, 96, 50),
    "the_next_week_3": ("the_next_week", 3, 96, 50),
    "the_next_week_4": ("the_next_week", 4, 96, 50),
    "the_next_week_5": ("the_next_week", 5, 64, 50),
    "the_next_week_6": ("the_next_week", 6, 96, 50),
    "the_next_week_7": ("the_next_week", 7, 64, 50),
    "the_next_week_8": ("the_next_week", 8, 64, 50),
    "the_next_week_9": ("the_next_week", 9, 64, 40),
    "the_rest_of_your_life": ("the_rest_of_your_life", None, 64, 50),
}


def psnr8(a, b):
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return 99.0 if mse == 0 else 10.0 * np.log10(255.0 ** 2 / mse)


def predicted_mse(mean, var, n):
    """8-bit MSE between two independent n-sample renders: 2 * E[(256 * d sqrt(x)/dx)^2 var / n], unsaturated pixels."""
    x = np.clip(mean, 1e-4, None)
    g = np.where(mean < 0.998, 256.0 / (2.0 * np.sqrt(x)), 0.0)
    return float(np.mean(2.0 * g * g * var / n) + 1.0 / 6.0)   # + quantisation of both images


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("scenes", nargs="*", default=list(SCENES))
    ap.add_argument("--target-db", type=float, default=43.5)
    ap.add_argument("--max-spp", type=int, default=1 << 18)
    ap.add_argument("--pilot-spp", type=int, default=256)
    args = ap.parse_args()
    os.makedirs(GOLD, exist_ok=True)
    manifest_path = os.path.join(GOLD, "converged_manifest.json")
    manifest = json.load(open(manifest_path)) if os.path.exists(manifest_path) else {}
    for key in args.scenes:
        book, no, width, depth = SCENES[key]
        def scene(spp):
            return R.BookScene(book, no, overrides=dict(image_width=width, samples_per_pixel=spp, max_depth=depth))
        s = scene(args.pilot_spp)
        o = O.OracleScene(s.desc)
        b3 = s.integrator == F.INT_BOOK3
        t0 = time.perf_counter()
        m, sq = o.render_moments(s.camera, s.integrator, seed=90001)
        rate = s.width * s.height * args.pilot_spp / (time.perf_counter() - t0)
        var = np.clip(sq - m * m, 0, None)
        ok = np.isfinite(m).all(axis=2)
        n = 1024
        step = 4 if b3 else 2      # book 3 traces floor(sqrt(spp))^2 samples: keep spp a perfect square
        want_mse = 255.0 ** 2 / 10 ** (args.target_db / 10)
        while n < args.max_spp and predicted_mse(m[ok], var[ok], n) > want_mse:
            n *= step
        print(f"{key}: pilot {rate / 1e6:.2f} Mpaths/s, predicted floor at {n} spp: "
              f"{10 * np.log10(255.0 ** 2 / predicted_mse(m[ok], var[ok], n)):.1f} dB, "
              f"~{2 * s.width * s.height * n / rate:.0f} s for both renders", flush=True)
        s = scene(n)     # same scene (same seed), new sample count; `o` keeps the description it was built from
        nan_scrub = b3
        t0 = time.perf_counter()
        a, asq = o.render_moments(s.camera, s.integrator, seed=90011)
        b, bsq = o.render_moments(s.camera, s.integrator, seed=90012)
        dt = time.perf_counter() - t0
        floor = psnr8(O.write_color(a, nan_scrub), O.write_color(b, nan_scrub))
        mean = 0.5 * (a + b)
        var = np.clip(0.5 * (asq + bsq) - mean * mean, 0, None)
        n_eff = 2 * (int(n ** 0.5) ** 2 if b3 else n)
        np.savez_compressed(os.path.join(GOLD, f"converged_{key}.npz"), mean=mean.astype(np.float32), var=var.astype(np.float32))
        frame_mean = np.nanmean(mean, axis=(0, 1))
        frame_se = np.sqrt(np.nansum(var, axis=(0, 1)) / n_eff) / np.isfinite(mean[..., 0]).sum()
        manifest[key] = dict(book=book, scene=no, width=s.width, height=s.height, depth=depth, spp_per_render=n, n=n_eff,
                             seeds=[90011, 90012], floor_psnr=round(float(floor), 2),
                             frame_mean=[float(x) for x in frame_mean], frame_mean_se=[float(x) for x in frame_se],
                             nan_pixels=int((~np.isfinite(mean).all(axis=2)).sum()), oracle_seconds=round(dt, 1),
                             scene_seed=hex(0x5EED1234))
        flag = "" if floor >= 42.0 else "   <-- BELOW 42 dB, raise --max-spp"
        print(f"{key}: {s.width}x{s.height} n={n_eff} floor {floor:.2f} dB, frame mean {frame_mean} "
              f"(s.e. {frame_se / frame_mean * 100} %), {dt:.0f} s{flag}", flush=True)
        with open(manifest_path, "w") as f:
            json.dump(manifest, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
