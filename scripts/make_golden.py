"""Writes tests/golden/probe_<scene>.npz: primary-hit ids and t of every book scene in probe mode
(pixel-centre rays, no jitter/defocus, time 0, media excluded) from the ORACLE at width 96, with the
scene seed 0x5EED1234.  The reference itself cannot be run here (no Rust toolchain, and it has no
object ids), so these vectors pin the oracle + scene builders against regressions and give the GPU
tests an oracle-free target.  Also writes tests/golden/radiance_means.json: converged mean linear
radiance per scene from the oracle (64x? frame, 4096 spp), the target of the mean-radiance gate.
Run here (CPU):  python scripts/make_golden.py [--radiance]
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import oracle_ffi as O  # noqa: E402
import rustraytrace_b200 as R  # noqa: E402

SCENES = [("in_one_weekend", None), ("in_one_weekend_gpu", None), ("the_rest_of_your_life", None)] + \
         [("the_next_week", i) for i in range(1, 10)]
GOLD = os.path.join(ROOT, "tests", "golden")


def name(book, no):
    return book if no is None else f"{book}_{no}"


def main():
    os.makedirs(GOLD, exist_ok=True)
    for book, no in SCENES:
        s = R.BookScene(book, no, overrides=dict(image_width=96))
        ids, t = O.OracleScene(s.desc).probe(s.camera)
        np.savez_compressed(os.path.join(GOLD, f"probe_{name(book, no)}.npz"), ids=ids.astype(np.int16 if ids.max() < 32767 else np.int32), t=t)
        print(name(book, no), ids.shape, "hit fraction", float((ids >= 0).mean()))
    if "--radiance" in sys.argv:
        out = {}
        for book, no in SCENES:
            s = R.BookScene(book, no, overrides=dict(image_width=48, samples_per_pixel=4096, max_depth=50))
            rgb, _ = O.OracleScene(s.desc).render(s.camera, s.integrator, seed=2024)
            out[name(book, no)] = dict(mean=[float(x) for x in np.nanmean(rgb, axis=(0, 1))], width=48, spp=4096, depth=50, seed=2024)
            print(name(book, no), out[name(book, no)]["mean"], flush=True)
        with open(os.path.join(GOLD, "radiance_means.json"), "w") as f:
            json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
