"""Quick GPU diagnostics: throughput of each BASELINE.json config at a reduced spp (device-resident)."""
import json
import sys
import time

import numpy as np
import torch

import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rustraytrace_b200 as R

F = R._ffi
from rustraytrace_b200 import configs
CONFIGS = configs.CONFIGS


def main():
    spp_cap = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    names = sys.argv[2].split(",") if len(sys.argv) > 2 else list(CONFIGS)
    ctx = R.Context(0)
    for name in names:
        s = configs.build(name, samples_per_pixel=min(CONFIGS[name][2]["samples_per_pixel"], spp_cap))
        gs = ctx.scene(s.desc)
        acc = torch.empty((s.height, s.width, 4), dtype=torch.float32, device="cuda")
        for _ in range(2):
            ctx.render_device(gs, s.camera, s.integrator, acc.data_ptr(), seed=1)
        st = ctx.stats()
        a = acc.cpu().numpy()
        print(json.dumps(dict(config=name, w=s.width, h=s.height, spp=st["spp_effective"], paths=st["paths"],
                              kernel_ms=round(st["kernel_ms"], 3), mpaths_s=round(st["paths"] / st["kernel_ms"] / 1e3, 1),
                              mean=[round(float(x), 5) for x in (a[..., :3].sum(axis=(0, 1)) / a[..., 3].sum())],
                              nodes=st["n_bvh_nodes"], grid=st["n_grid_cells"], prims=st["n_prims"], sm=st["sm_count"])), flush=True)


if __name__ == "__main__":
    main()
