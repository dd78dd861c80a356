"""Warp-level structure of the render kernel from an ncu report (needs --import-source on):
runs of consecutive SASS instructions with the same execution count, as executions per
world_hit32 call (one call = one path segment of the warp) and active lanes.
    python scripts/ncu_groups.py gpurun_out/prof.ncu-rep [min_share_percent]
A run executed more than once per segment outside a loop is a group of lanes that did not
re-converge with the others (profiles/README.md, r01j)."""
import csv
import subprocess
import sys

rep = sys.argv[1]
min_share = float(sys.argv[2]) if len(sys.argv) > 2 else 0.4
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hdr = rows[1]
iA, iS, iI, iT = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
data = [r for r in rows[2:] if len(r) > iT]
base = int(data[0][iA], 16)
tot = sum(int(r[iI]) for r in data)
# one segment = one world_hit32 call: the most frequent execution count among the instructions that
# run with (nearly) all lanes of the warp (ray set-up, Philox block, hit record): the mode
from collections import Counter
mode = Counter()
for r in data:
    if int(r[iI]) and int(r[iT]) / int(r[iI]) >= 20:
        mode[int(r[iI])] += int(r[iI])   # weighted by executions, so the once-per-work-item prologue does not win
seg = mode.most_common(1)[0][0]
print(f"warp instructions {tot:.4g}; segment (world_hit32 call) count {seg:.4g}; {tot / seg:.0f} warp instructions per segment")
groups = []
for r in data:
    I, T = int(r[iI]), int(r[iT])
    lanes = T / max(I, 1)
    if groups and abs(groups[-1]["I"] - I) <= 0.04 * max(I, 1) and abs(groups[-1]["l"] - lanes) < 1.0:
        g = groups[-1]; g["n"] += 1; g["sum"] += I; g["last"] = r[iS].strip()[:40]
    else:
        groups.append(dict(I=I, l=lanes, off=int(r[iA], 16) - base, n=1, sum=I, first=r[iS].strip()[:40], last=""))
for g in groups:
    if 100 * g["sum"] / tot >= min_share:
        print(f"{g['off']:05x} n={g['n']:3d} x{g['I'] / seg:5.2f}/seg lanes={g['l']:4.1f} share={100 * g['sum'] / tot:5.1f}%  {g['first']} .. {g['last']}")
