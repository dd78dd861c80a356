"""Selected rows of `ncu --page raw` for one kernel launch of a report, as metric,unit,value:
    python scripts/ncu_summary.py gpurun_out/prof.ncu-rep profiles/ncu_rNN_render_kernel_c2_spp64.csv [launch_index]
The metric list is the one of the previous capture in profiles/ (so rounds stay comparable)."""
import csv
import glob
import os
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
launch = int(sys.argv[3]) if len(sys.argv) > 3 else -1
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
prev = sorted(p for p in glob.glob(os.path.join(ROOT, "profiles", "ncu_r*_render_kernel_c2_spp64.csv")) if os.path.abspath(p) != os.path.abspath(out))
want = [ln.split(",")[0] for ln in open(prev[-1]).read().splitlines()[1:]]
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
r = data[launch]
col = {n: i for i, n in enumerate(hdr)}
with open(out, "w") as f:
    f.write("metric,unit,value\n")
    for m in want:
        if m in col:
            f.write(f"{m},{units[col[m]]},{r[col[m]].replace(',', '')}\n")
print(open(out).read())
