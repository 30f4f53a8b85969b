"""Developer tool: which source lines collect which warp-stall samples (ncu source CSV + nvdisasm line info).

    python tools/sass_stalls.py <ncu source csv> <cubin> <kernel name substring> <stall column, e.g. stall_long_sb> [top]
"""
import csv, re, subprocess, sys
from collections import defaultdict

src_csv, cubin, kname, col = sys.argv[1:5]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 15
rows = list(csv.reader(open(src_csv)))
hdr, data = rows[1], rows[2:]
ic, iS = hdr.index(col), hdr.index("Source")
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(dis) if l.startswith("\t.section\t.text.") and kname in l)
lines, cur = [], None
for l in dis[start + 1:]:
    if l.startswith("\t.section"):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        lines.append(cur)
agg = defaultdict(int)
ops = defaultdict(lambda: defaultdict(int))
tot = 0
for loc, r in zip(lines, data):
    v = int(r[ic] or 0)
    agg[loc] += v
    tot += v
    if v:
        ops[loc][r[iS].split()[0] if not r[iS].strip().startswith("@") else r[iS].split()[1]] += v
print(f"{col}: {tot} samples")
for loc, v in sorted(agg.items(), key=lambda kv: -kv[1])[:top]:
    o = sorted(ops[loc].items(), key=lambda kv: -kv[1])[:3]
    print(f"{v / max(tot, 1) * 100:5.1f}%  {loc}  {o}")
