#!/usr/bin/env python
"""Condense ncu output into the small tracked files under profiles/.

    python tools/ncu_summary.py launches <launches.csv>            -> per-kernel time shares
    python tools/ncu_summary.py full <report.ncu-rep> <out.csv> <traffic.json>
"""
import collections
import csv
import json
import subprocess
import sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "sm__warps_active.avg.pct_of_peak_sustained_active",
           "l1tex__throughput.avg.pct_of_peak_sustained_active",
           "lts__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
           "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size",
           "launch__block_size"]


def short(name):
    name = name.replace("<unnamed>::", "").replace("void ", "")
    return name.split("(")[0]


def launches(path):
    rows = list(csv.reader(open(path)))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    d = collections.OrderedDict()
    for r in rows[hdr + 1:]:
        if len(r) >= 15:
            d.setdefault(short(r[4]), []).append(float(r[-1]))
    tot = sum(sum(v) for v in d.values())
    print("kernel,launches,avg_us,total_us,share")
    for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
        print(f"{k},{len(v)},{sum(v) / len(v) / 1e3:.2f},{sum(v) / 1e3:.1f},{sum(v) / tot:.4f}")


def full(rep, out_csv, out_json):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--metrics", ",".join(METRICS)],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    cols = [i for i, h in enumerate(hdr) if h in METRICS]
    ki = hdr.index("Kernel Name")
    traffic = {}
    with open(out_csv, "w", newline="") as fh:
        w = csv.writer(fh)
        w.writerow(["kernel"] + [f"{hdr[i]} [{units[i]}]" for i in cols])
        for r in rows[2:]:
            name = short(r[ki])
            w.writerow([name] + [r[i] for i in cols])
            rd = float(r[hdr.index("dram__bytes_read.sum")] or 0)
            wr = float(r[hdr.index("dram__bytes_write.sum")] or 0)
            mult = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}
            rd *= mult.get(units[hdr.index("dram__bytes_read.sum")], 1.0)
            wr *= mult.get(units[hdr.index("dram__bytes_write.sum")], 1.0)
            t = traffic.setdefault(name, {"launches": 0, "dram_bytes": 0.0, "time_us": 0.0})
            t["launches"] += 1
            t["dram_bytes"] += rd + wr
            t["time_us"] += float(r[hdr.index("gpu__time_duration.sum")])
    for t in traffic.values():
        t["dram_bytes_per_launch"] = t["dram_bytes"] / t["launches"]
        t["time_us_per_launch"] = t["time_us"] / t["launches"]
    json.dump(traffic, open(out_json, "w"), indent=1)


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    else:
        full(*sys.argv[2:5])
