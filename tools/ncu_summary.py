"""Developer tool: .ncu-rep (ncu --set full) -> compact CSV of the metrics DESIGN.md and bench.py quote, one row per launch.

    python tools/ncu_summary.py gpurun_out/x.ncu-rep profiles/x.csv [profiles/traffic.json]
The optional JSON maps kernel name -> DRAM bytes per launch (read + written, mean over the captured launches): what
bench.py reports as roofline.traffic.
"""
import csv, io, json, subprocess, sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"]
UNIT_SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    cols = [k for k in KEYS if k in hdr]
    traffic = {}
    with open(out, "w", newline="") as fh:
        w = csv.writer(fh)
        w.writerow(["kernel"] + [f"{k} [{units[hdr.index(k)]}]" for k in cols])
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            name = d["Kernel Name"].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
            w.writerow([name] + [d[k] for k in cols])
            total = 0.0
            for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                total += float(d[k].replace(",", "")) * UNIT_SCALE[units[hdr.index(k)]]
            traffic.setdefault(name, []).append(total)
    if len(sys.argv) > 3:
        json.dump({k: sum(v) / len(v) for k, v in traffic.items()}, open(sys.argv[3], "w"), indent=1)
    print(f"{len(rows) - 2} launches -> {out}")


main()
