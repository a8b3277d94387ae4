#!/usr/bin/env python3
"""Per-source-line hot spots of one kernel from an .ncu-rep (needs -lineinfo + --import-source on).

    python scripts/ncu_hotspots.py report.ncu-rep [top_n] > profiles/xxx_hotspots.csv

Columns: file, line, source, % of executed warp instructions, % of stall samples, average
active threads per executed instruction, dominant stall reasons.
"""
import csv
import io
import os
import subprocess
import sys
from collections import defaultdict


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 60
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    agg = defaultdict(lambda: defaultdict(float))
    src = {}
    cur_file = None
    hdr = None
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            cur_file = os.path.basename(r[1])
            continue
        if r and r[0] == "Line No":
            hdr = r
            continue
        if hdr is None or len(r) != len(hdr) or not r[0].isdigit():
            continue
        d = dict(zip(hdr, r))
        # the first 'Source' column is the CUDA line, rows with an empty Address are line summaries
        if r[2] != "-":
            continue
        key = (cur_file, int(r[0]))
        src[key] = r[1].strip()
        a = agg[key]
        for k in ("# Samples", "Instructions Executed", "Thread Instructions Executed"):
            try:
                a[k] += float(d[k] or 0)
            except ValueError:
                pass
        for k in hdr:
            if k.startswith("stall_") and "Not Issued" not in k:
                try:
                    a[k] += float(d[k] or 0)
                except ValueError:
                    pass
    ti = sum(a["Instructions Executed"] for a in agg.values()) or 1.0
    ts = sum(a["# Samples"] for a in agg.values()) or 1.0
    w = csv.writer(sys.stdout)
    w.writerow(["file", "line", "source", "pct_warp_inst", "pct_stall_samples", "avg_active_threads", "top_stalls"])
    for key, a in sorted(agg.items(), key=lambda kv: -kv[1]["# Samples"])[:top]:
        st = sorted(((v, k) for k, v in a.items() if k.startswith("stall_") and v > 0), reverse=True)[:3]
        act = a["Thread Instructions Executed"] / a["Instructions Executed"] if a["Instructions Executed"] else 0
        w.writerow([key[0], key[1], src[key][:110], "%.2f" % (100 * a["Instructions Executed"] / ti),
                    "%.2f" % (100 * a["# Samples"] / ts), "%.1f" % act,
                    " ".join("%s:%.0f%%" % (k[6:], 100 * v / max(a["# Samples"], 1)) for v, k in st)])
    # totals by stall reason
    tot = defaultdict(float)
    for a in agg.values():
        for k, v in a.items():
            if k.startswith("stall_"):
                tot[k] += v
    s = sum(tot.values()) or 1.0
    sys.stderr.write("stall mix: " + ", ".join("%s %.1f%%" % (k[6:], 100 * v / s)
                                               for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:8]) + "\n")


if __name__ == "__main__":
    main()
