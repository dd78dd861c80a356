"""Algorithmic flops per camera path of every BASELINE.json config, from the
ORACLE's op counters on the reference-shaped scene graph (pointer-tree BVH as bvh.rs builds it).
Writes tests/golden/flops_per_path.json.  Run here (CPU only):  python scripts/count_flops.py

Each config is sampled at full resolution on every `row_step`-th row with `spp` samples per pixel
(path statistics do not depend on spp; book 3 uses a square spp).
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import oracle_ffi as O  # noqa: E402
from rustraytrace_b200 import configs  # noqa: E402

SAMPLE = {"C1": (1, 16), "C2": (5, 8), "C3a": (9, 4), "C3b": (9, 16), "C3c": (9, 4), "C4": (8, 4), "C5": (32, 16)}


def main():
    out = {}
    for key in configs.CONFIGS:
        step, spp = SAMPLE[key]
        s = configs.build(key, samples_per_pixel=spp)
        o = O.OracleScene(s.desc)
        t = time.time()
        _, c = o.render(s.camera, s.integrator, seed=1, rows=(0, s.height, step))
        dt = time.time() - t
        out[key] = dict(flops_per_path=round(O.flops_per_path(c), 1), segments_per_path=round(c["segments"] / c["paths"], 3),
                        aabb_per_path=round(c["aabb"] / c["paths"], 2), sample=f"every {step}th row, {spp} spp, {int(c['paths'])} paths",
                        oracle_mpaths_s=round(c["paths"] / dt / 1e6, 4), threads=os.cpu_count())
        print(key, out[key], flush=True)
    with open(os.path.join(ROOT, "tests", "golden", "flops_per_path.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
