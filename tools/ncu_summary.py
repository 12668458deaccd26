#!/usr/bin/env python
"""Summarise an `ncu --set full` report into one CSV row per launch (the table kept under profiles/).

    python tools/ncu_summary.py gpurun_out/v5_full.ncu-rep "comment line; build=<id from gpurun_out/<tag>_build.txt>" > profiles/r1_ncu_full_v5_summary.csv
"""
import csv
import io
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum",
]


def main():
    rep = sys.argv[1]
    note = sys.argv[2] if len(sys.argv) > 2 else ""
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    ci = {h: i for i, h in enumerate(hdr)}
    keep = ["Kernel Name", "Grid Size", "Block Size"] + [m for m in METRICS if m in ci]
    w = csv.writer(sys.stdout)
    w.writerow(["# ncu --set full --clock-control none; %s; from %s" % (note, rep)])
    w.writerow(keep)
    w.writerow([units[ci[k]] for k in keep])
    for r in body:
        if len(r) != len(hdr):
            continue
        out = []
        for k in keep:
            v = r[ci[k]]
            if k == "Kernel Name":
                v = v.split("(")[0].replace("hf::<unnamed>::", "")[-60:]
            out.append(v)
        w.writerow(out)


if __name__ == "__main__":
    main()
