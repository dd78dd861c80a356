"""Writes tests/golden/probe_<scene>.npz: primary-hit ids and t of every book scene in probe mode
(pixel-centre rays, no jitter/defocus, time 0, media excluded) from the ORACLE at width 96, with the
scene seed 0x5EED1234.  The reference itself cannot be run here (no Rust toolchain, and it has no
object ids), so these vectors pin the oracle + scene builders against regressions and give the GPU
tests an oracle-free target.  (Converged images and frame means: scripts/make_golden_converged.py.)
Run here (CPU):  python scripts/make_golden.py
"""
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


if __name__ == "__main__":
    main()
