"""Developer tool: per-kernel shares of an `ncu --metrics gpu__time_duration.sum --csv` launch list.

    python tools/launch_summary.py gpurun_out/x_launches.csv "command that was profiled" > profiles/x_launches_summary.txt
"""
import csv, sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
by = {}
for r in rows:
    name = r[4].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
    t = by.setdefault(name, [0, 0.0])
    t[0] += 1
    t[1] += float(r[14].replace(",", "")) / 1e6
total = sum(v[1] for v in by.values())
print(f"ncu --metrics gpu__time_duration.sum --clock-control none: {sys.argv[2] if len(sys.argv) > 2 else ''}")
print("launch times under ncu are serialised and cold-cache: compare SHARES, not absolute times")
print(f"{'kernel':66s} {'n':>4s} {'total ms':>10s} {'avg ms':>9s}  share")
for name, (n, ms) in sorted(by.items(), key=lambda kv: -kv[1][1]):
    print(f"{name[:66]:66s} {n:4d} {ms:10.3f} {ms / n:9.3f} {100 * ms / total:5.1f}%")
