"""Per-source-line instruction counts of an ncu report (needs -lineinfo and --import-source on):
    python scripts/ncu_lines.py gpurun_out/prof.ncu-rep [top_n]"""
import csv
import subprocess
import sys

rep, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur, agg, hdr = None, {}, None
for r in csv.reader(txt.splitlines()):
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split('/')[-1]; continue
    if len(r) > 5 and r[0] == "Line No":
        hdr = r; iI = r.index("Instructions Executed"); iT = r.index("Thread Instructions Executed"); iS = r.index("# Samples"); continue
    if hdr is None or len(r) < len(hdr) - 2:
        continue
    try:
        ln = int(r[0]); inst = int(r[iI] or 0); tinst = int(r[iT] or 0); smp = int(r[iS] or 0)
    except ValueError:
        continue
    a = agg.setdefault((cur, ln), [0, 0, 0, r[1].strip()[:84]]); a[0] += inst; a[1] += tinst; a[2] += smp
tot = sum(a[0] for a in agg.values()); ts = sum(a[2] for a in agg.values()); tt = sum(a[1] for a in agg.values())
print(f"warp inst {tot:.4g}  thread inst {tt:.4g}  lanes/inst {tt / tot:.2f}  samples {ts}")
byf = {}
for (f, l), a in agg.items():
    b = byf.setdefault(f, [0, 0, 0]); b[0] += a[0]; b[1] += a[1]; b[2] += a[2]
for f, b in sorted(byf.items(), key=lambda x: -x[1][0])[:8]:
    print(f"  {f:18s} {100 * b[0] / tot:5.1f}% inst  {b[1] / max(1, b[0]):5.1f} lanes  {100 * b[2] / ts:5.1f}% samples")
for (f, l), a in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    print(f"{f[:12]:12s}:{l:4d} {100 * a[0] / tot:5.1f}% l={a[1] / max(1, a[0]):4.1f} s={100 * a[2] / ts:4.1f}%  {a[3]}")
