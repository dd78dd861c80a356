"""Quick GPU diagnostics: throughput of each BASELINE.json config at a reduced spp (device-resident)."""
import json
import sys
import time

import numpy as np
import torch

import rustraytrace_b200 as R

F = R._ffi
CONFIGS = {
    "C1": ("in_one_weekend", None, dict(image_width=400, samples_per_pixel=100, max_depth=50, aspect_ratio=16 / 9)),
    "C2": ("in_one_weekend", None, dict(image_width=1200, samples_per_pixel=500, max_depth=50)),
    "C3a": ("the_next_week", 1, dict(image_width=1920, samples_per_pixel=1000, max_depth=50)),
    "C3b": ("the_next_week", 3, dict(image_width=1920, samples_per_pixel=1000, max_depth=50)),
    "C3c": ("the_next_week", 4, dict(image_width=1920, samples_per_pixel=1000, max_depth=50)),
    "C4": ("the_next_week", 9, dict(image_width=800, samples_per_pixel=10000, max_depth=40)),
    "C5": ("the_rest_of_your_life", None, dict(image_width=3840, samples_per_pixel=4096, max_depth=50)),
}


def main():
    spp_cap = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    names = sys.argv[2].split(",") if len(sys.argv) > 2 else list(CONFIGS)
    ctx = R.Context(0)
    for name in names:
        book, sc, ov = CONFIGS[name]
        ov = dict(ov)
        ov["samples_per_pixel"] = min(ov["samples_per_pixel"], spp_cap)
        s = R.BookScene(book, sc, overrides=ov)
        gs = ctx.scene(s.desc)
        acc = torch.empty((s.height, s.width, 4), dtype=torch.float32, device="cuda")
        for _ in range(2):
            ctx.render_device(gs, s.camera, s.integrator, acc.data_ptr(), seed=1)
        st = ctx.stats()
        a = acc.cpu().numpy()
        print(json.dumps(dict(config=name, w=s.width, h=s.height, spp=st["spp_effective"], paths=st["paths"],
                              kernel_ms=round(st["kernel_ms"], 3), mpaths_s=round(st["paths"] / st["kernel_ms"] / 1e3, 1),
                              mean=[round(float(x), 5) for x in (a[..., :3].sum(axis=(0, 1)) / a[..., 3].sum())],
                              nodes=st["n_bvh_nodes"], prims=st["n_prims"], sm=st["sm_count"])), flush=True)


if __name__ == "__main__":
    main()
