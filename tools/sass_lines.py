"""Developer tool: join ncu's per-instruction counters (ncu -i rep --page source --csv) with nvdisasm's line info, and print the
source lines that execute most instructions.

    python tools/sass_lines.py <ncu source csv> <cubin> <kernel name substring> [top]
"""
import csv, re, subprocess, sys
from collections import defaultdict

src_csv, cubin, kname = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
rows = list(csv.reader(open(src_csv)))
hdr, data = rows[1], rows[2:]
for i, r in enumerate(data):  # a report with several matching launches: one block per launch, each opened by a 'Kernel Name' row
    if r and r[0] == "Kernel Name":
        data = data[:i]
        break
iE, iT, iW = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
# find the section of the kernel
start = next(i for i, l in enumerate(dis) if l.startswith("\t.section\t.text.") and kname in l)
lines = []
cur = None
for l in dis[start + 1:]:
    if l.startswith("\t.section"):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        lines.append(cur)
assert len(lines) == len(data), (len(lines), len(data))
agg = defaultdict(lambda: [0, 0, 0, 0])
tot = 0
for loc, r in zip(lines, data):
    a = agg[loc]
    a[0] += int(r[iE]); a[1] += int(r[iT]); a[2] += int(r[iW]); a[3] += 1
    tot += int(r[iE])
print(f"total warp instructions {tot}")
for loc, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{a[0] / tot * 100:5.1f}%  lanes {a[1] / max(a[0], 1):5.1f}  samples {a[2]:7d}  sass {a[3]:4d}  {loc}")
