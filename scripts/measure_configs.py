"""Measures every BASELINE.json config at FULL size and FULL spp on one B200 and writes
profiles/configs_<tag>.json: device-resident Mpaths/s (kernel CUDA events), FP32-roofline fraction,
CPU-oracle Mpaths/s on a bounded row sample (threads stated), full-frame probe parity, and the
frame mean against the oracle's converged mean.   python scripts/measure_configs.py r01 [C1,C2,...]
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import oracle_ffi as O  # noqa: E402
import rustraytrace_b200 as R  # noqa: E402
from rustraytrace_b200 import configs  # noqa: E402

GOLD_NAME = {"C1": "in_one_weekend", "C2": "in_one_weekend", "C3a": "the_next_week_1", "C3b": "the_next_week_3",
             "C3c": "the_next_week_4", "C4": "the_next_week_9", "C5": "the_rest_of_your_life"}


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
    names = sys.argv[2].split(",") if len(sys.argv) > 2 else list(configs.CONFIGS)
    cpu_seconds = float(os.environ.get("CPU_SECONDS", "6"))
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "converged_manifest.json")))
    threads = os.cpu_count() or 1
    ctx = R.Context(0)
    out = {}
    for key in names:
        s = configs.build(key)
        gs = ctx.scene(s.desc)
        acc = torch.empty((s.height, s.width, 4), dtype=torch.float32, device="cuda")
        warm = configs.build(key, samples_per_pixel=min(16, s.camera.samples_per_pixel))
        ctx.render_device(gs, warm.camera, warm.integrator, acc.data_ptr(), seed=1)
        ctx.render_device(gs, s.camera, s.integrator, acc.data_ptr(), seed=1)
        st = ctx.stats()
        a = acc.double().sum(dim=(0, 1)).cpu().numpy()
        mean = a[:3] / a[3]
        ids, t = ctx.probe(gs, s.camera, 64)
        o = O.OracleScene(s.desc)
        oids, ot = o.probe(s.camera)
        hit = oids >= 0
        # CPU oracle on every k-th row at full width and spp, sized for ~cpu_seconds
        cal = configs.build(key, samples_per_pixel=4)
        t0 = time.perf_counter()
        _, c = O.OracleScene(cal.desc).render(cal.camera, cal.integrator, seed=3, nthreads=threads, rows=(0, s.height, max(1, s.height // 32)))
        rate = c["paths"] / (time.perf_counter() - t0)
        rows = int(max(threads, min(s.height, rate * cpu_seconds / (s.width * st["spp_effective"]))))
        step = max(1, s.height // rows)
        t0 = time.perf_counter()
        _, c = o.render(s.camera, s.integrator, seed=1, nthreads=threads, rows=(0, s.height, step))
        cpu_dt = time.perf_counter() - t0
        mp = st["paths"] / st["kernel_ms"] / 1e3
        fp = configs.F_PATH[key]
        gm = np.array(gold[GOLD_NAME[key]]["frame_mean"])
        out[key] = dict(
            workload=configs.DESCRIPTION[key], width=s.width, height=s.height, spp=st["spp_effective"], paths=st["paths"],
            kernel_ms=round(st["kernel_ms"], 2), mpaths_s=round(mp, 1), flops_per_path=fp,
            tflops=round(mp * 1e6 * fp / 1e12, 2), roofline_frac=round(mp * 1e6 * fp / 1e12 / 74.45, 4),
            cpu_oracle_mpaths_s=round(c["paths"] / cpu_dt / 1e6, 3), cpu_threads=threads,
            cpu_sample=f"every {step}th row at full width/spp, {int(c['paths'])} paths in {cpu_dt:.1f} s",
            speedup_vs_cpu=round(mp / (c["paths"] / cpu_dt / 1e6), 1),
            probe_ids_differ=int((ids != oids).sum()), probe_pixels=int(ids.size),
            probe_t_max_rel=float(np.max(np.abs(t[hit] - ot[hit]) / ot[hit])) if hit.any() else 0.0,
            frame_mean=[round(float(x), 5) for x in mean], oracle_converged_mean=[round(float(x), 5) for x in gm],
            mean_rel_diff=[round(float(x), 5) for x in np.abs(mean - gm) / gm],
            prims=st["n_prims"], bvh_nodes=st["n_bvh_nodes"])
        print(key, json.dumps(out[key]), flush=True)
        gs.close()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"configs_{tag}.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
