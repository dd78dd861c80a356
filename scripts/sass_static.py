"""Static size of one kernel by source region, from `nvdisasm -g -c` of its cubin (needs -lineinfo):
    cuobjdump -xelf all rustraytrace_b200/librrt.so && nvdisasm -g -c kernels_book1.sm_100a.cubin > b1.dis
    python scripts/sass_static.py b1.dis render_kernelILi1ELi64ELb1 [lines_per_bucket]
What is executed every segment has to fit the 32 KB instruction cache level (16 B an instruction): profiles/README.md r02m/n."""
import collections
import re
import sys

dis, name = sys.argv[1], sys.argv[2]
bucket = int(sys.argv[3]) if len(sys.argv) > 3 else 10
lines = open(dis).read().split("\n")
start = end = None
for i, l in enumerate(lines):
    if l.startswith("//---------------------") and name in l and start is None:
        start = i
    elif start is not None and l.startswith("//---------------------") and i > start + 3:
        end = i
        break
cur, by = None, collections.Counter()
for l in lines[start:end]:
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
    elif re.match(r"\s+/\*[0-9a-f]{4,5}\*/", l) and cur:
        by[cur] += 1
tot = sum(by.values())
print(f"{name}: {tot} instructions, {tot * 16 / 1024:.1f} KB")
byfile = collections.Counter()
agg = collections.Counter()
for (f, ln), c in by.items():
    byfile[f] += c
    agg[(f, ln // bucket * bucket)] += c
print("  ".join(f"{f} {c}" for f, c in byfile.most_common(6)))
for (f, ln), c in agg.most_common(25):
    print(f"{f}:{ln}-{ln + bucket - 1}  {c}")
