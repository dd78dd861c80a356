"""Markdown rows of BASELINE.md section 4 from profiles/configs_<tag>.json:  python scripts/results_table.py r01d"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
d = json.load(open(os.path.join(ROOT, "profiles", f"configs_{tag}.json")))
ref = {"C2": "652 (same scene, 128 spp)"}
for k, v in d.items():
    mean = max(abs(x) for x in v["mean_rel_diff"]) * 100
    print(f"| {k} {v['workload']} | {v['paths']:.3g} | {v['cpu_oracle_mpaths_s']} ({v['cpu_threads']}) | {ref.get(k, '—')} | "
          f"**{v['mpaths_s']}** ({v['kernel_ms']:.2f} ms) | {v['speedup_vs_cpu']}x | {v['flops_per_path']} | "
          f"{v['tflops']} TFLOP/s = {100 * v['roofline_frac']:.1f} % | {v['probe_ids_differ']} of {v['probe_pixels']} / {v['probe_t_max_rel']:.1e} | {mean:.2f} % |")
