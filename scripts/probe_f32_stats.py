"""Primary hits of the f32 intersection code the render kernel traces with (rrt_probe precision 32)
against the oracle's f64 probe, on every BASELINE.json config at full size:
    python scripts/probe_f32_stats.py r02 [C1,C2,...]      -> gpurun_out/probe_f32_<tag>.json
Per config: pixels whose id differs, where they sit (fraction on a silhouette = a pixel with a
neighbour of another oracle id), max / p99 / median relative t error among equal ids, and the same
for the f64 probe kernel.  C4 is run with box leaves on and off (RRT_NO_BOX_LEAVES, read at
rrt_scene_create).  tests/test_full_size.py gates at twice these values (tests/golden/probe_f32.json)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import oracle_ffi as O  # noqa: E402
import rustraytrace_b200 as R  # noqa: E402
from rustraytrace_b200 import configs  # noqa: E402


def silhouette(ids):
    e = np.zeros(ids.shape, dtype=bool)
    e[:, 1:] |= ids[:, 1:] != ids[:, :-1]; e[:, :-1] |= ids[:, 1:] != ids[:, :-1]
    e[1:, :] |= ids[1:, :] != ids[:-1, :]; e[:-1, :] |= ids[1:, :] != ids[:-1, :]
    return e


def stats(ids, t, oids, ot):
    diff = ids != oids
    same = (~diff) & (oids >= 0)
    rel = np.abs(t[same] - ot[same]) / ot[same] if same.any() else np.zeros(1)
    sil = silhouette(oids)
    worst = int(np.argmax(rel)) if same.any() else 0
    ys, xs = np.nonzero(same)
    return dict(ids_differ=int(diff.sum()), pixels=int(ids.size), ids_differ_frac=float(diff.mean()),
                ids_differ_on_silhouette=int((diff & sil).sum()),
                t_max_rel=float(rel.max()), t_p99_rel=float(np.quantile(rel, 0.99)), t_median_rel=float(np.median(rel)),
                t_over_1e5=int((rel > 1e-5).sum()), t_over_1e4=int((rel > 1e-4).sum()),
                worst_pixel=[int(xs[worst]), int(ys[worst])] if same.any() else None,
                worst_id=int(oids[ys[worst], xs[worst]]) if same.any() else None,
                worst_t=float(ot[ys[worst], xs[worst]]) if same.any() else None)


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
    names = sys.argv[2].split(",") if len(sys.argv) > 2 else list(configs.CONFIGS)
    ctx = R.Context(0)
    out = {}
    for key in names:
        s = configs.build(key)
        oids, ot = O.OracleScene(s.desc).probe(s.camera)
        variants = [("", {})]
        if key == "C4":
            variants.append(("_no_box_leaves", {"RRT_NO_BOX_LEAVES": "1"}))
        for suffix, env in variants:
            os.environ.update(env)
            gs = ctx.scene(s.desc)
            for k in env:
                del os.environ[k]
            i32, t32 = ctx.probe(gs, s.camera, 32)
            i64, t64 = ctx.probe(gs, s.camera, 64)
            gs.close()
            out[key + suffix] = dict(width=s.width, height=s.height, f32=stats(i32, t32, oids, ot), f64=stats(i64, t64, oids, ot))
            print(key + suffix, json.dumps(out[key + suffix]), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"probe_f32_{tag}.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
