#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` export: top SASS lines by stall samples.

    ncu -i prof.ncu-rep --page source --csv --kernel-name regex:<k> --launch-count 1 > src.csv
    python profiles/srcpage.py src.csv [top_n]
"""
import csv
import sys


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    rows = list(csv.reader(open(path)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    body = []
    for r in rows[hdr_i + 1:]:
        if r and r[0] == "Kernel Name":
            break  # next kernel of a multi-kernel export
        if len(r) == len(hdr) and r[0].startswith("0x"):
            body.append(r)
    ci = {h: i for i, h in enumerate(hdr)}
    samp = ci["# Samples"]
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    total = sum(int(r[samp] or 0) for r in body)
    print("kernel:", rows[0][1] if len(rows[0]) > 1 else "?", "| SASS lines:", len(body), "| samples:", total)
    tot_stall = {}
    for r in body:
        for i in stall_cols:
            tot_stall[hdr[i]] = tot_stall.get(hdr[i], 0) + int(r[i] or 0)
    print("stall totals:", ", ".join("%s=%d" % kv for kv in sorted(tot_stall.items(), key=lambda kv: -kv[1]) if kv[1]))
    order = sorted(range(len(body)), key=lambda k: -int(body[k][samp] or 0))[:top]
    for k in sorted(order):
        r = body[k]
        st = sorted(((int(r[i] or 0), hdr[i][6:]) for i in stall_cols), reverse=True)[:3]
        print("%5d %6s  %-70s %s" % (k, r[samp], r[ci["Source"]].strip()[:70],
                                     " ".join("%s:%d" % (n, v) for v, n in st if v)))


if __name__ == "__main__":
    main()
