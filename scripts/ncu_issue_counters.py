#!/usr/bin/env python3
"""Extracts from an `ncu --set full` report the counters bench.py's roofline line is built from
and writes them as JSON (committed under profiles/, read back by bench.py):

    python scripts/ncu_issue_counters.py report.ncu-rep [kernel-substring] [--scene NAME]
        [--shade-report shade.ncu-rep --rays-per-launch N] > profiles/rN_issue_counters.json

  issue_slot_busy_pct   sm__inst_issued.avg.pct_of_peak_sustained_active ("Issue Slots Busy")
  active_threads        smsp__thread_inst_executed_per_inst_executed.ratio
  nonpred_threads       smsp__thread_inst_executed_pred_on_per_inst_executed.ratio
  lane_issue_frac       issue_slot_busy_pct / 100 * nonpred_threads / 32  — the fraction of the
                        machine's lane-issue capacity that does useful work: the roofline a
                        divergent traversal kernel lives under (neither HBM nor the FP32 pipe
                        is anywhere near its peak, see the dram_* and fp32 fields)
plus occupancy, cache hit rates, DRAM bytes and duration of the profiled launches (averaged over
the launches in the report that match the kernel substring), and a hash of the kernel sources
the report belongs to.  With --shade-report and --rays-per-launch (captures of full waves of a
known size, FS2_TUNE pool) it also writes dram_bytes_per_segment = (DRAM bytes of one k_wf_trace
+ one k_wf_shade launch) / rays: the measured HBM traffic per ray segment bench.py scales.
"""
import csv
import hashlib
import io
import json
import os
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(REPO, "final-stage-path-tracer-2-0_b200", "csrc")

FIELDS = {
    "issue_slot_busy_pct": "sm__inst_issued.avg.pct_of_peak_sustained_active",
    "active_threads": "smsp__thread_inst_executed_per_inst_executed.ratio",
    "nonpred_threads": "smsp__thread_inst_executed_pred_on_per_inst_executed.ratio",
    "achieved_occupancy_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
    "ipc_active": "sm__inst_executed.avg.per_cycle_active",
    "l1_hit_pct": "l1tex__t_sector_hit_rate.pct",
    "l2_hit_pct": "lts__t_sector_hit_rate.pct",
    "dram_read_mb": "dram__bytes_read.sum",
    "dram_write_mb": "dram__bytes_write.sum",
    "dram_pct_of_peak": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "duration_ms": "gpu__time_duration.sum",
    "registers": "launch__registers_per_thread",
    "warp_instructions": "smsp__inst_executed.sum",
}


def csrc_hash():
    """sha256 over the kernel sources: ties a counters file to the build it was captured from."""
    h = hashlib.sha256()
    for name in sorted(os.listdir(CSRC)):
        if name.endswith((".cu", ".cuh", ".h", ".inc")):
            h.update(name.encode())
            h.update(open(os.path.join(CSRC, name), "rb").read())
    return h.hexdigest()[:16]


def to_unit(value, unit, want):
    v = float(value.replace(",", ""))
    scale = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3, "ns": 1e-6, "us": 1e-3, "ms": 1.0,
             "s": 1e3, "usecond": 1e-3, "msecond": 1.0, "nsecond": 1e-6, "second": 1e3}
    if want in ("mb", "ms") and unit in scale:
        return v * scale[unit]
    return v


def load(rep, pattern):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    name_col = col.get("Kernel Name")
    picked = [r for r in data if name_col is None or pattern in r[name_col]]
    if not picked:
        raise SystemExit("no launch of %r in %s" % (pattern, rep))
    return col, units, picked, name_col


def dram_mb(rep, pattern):
    col, units, picked, _ = load(rep, pattern)
    tot = []
    for r in picked:
        tot.append(sum(to_unit(r[col[m]], units[col[m]], "mb") for m in ("dram__bytes_read.sum", "dram__bytes_write.sum")))
    return sum(tot) / len(tot)


def main():
    argv = sys.argv[1:]
    opts = {}
    for flag in ("--scene", "--shade-report", "--rays-per-launch"):
        if flag in argv:
            i = argv.index(flag)
            opts[flag] = argv[i + 1]
            del argv[i:i + 2]
    rep = argv[0]
    pattern = argv[1] if len(argv) > 1 else "k_wf_trace"
    col, units, picked, name_col = load(rep, pattern)
    out = {"kernel": picked[0][name_col] if name_col is not None else pattern, "launches": len(picked),
           "report": os.path.basename(rep), "csrc_sha16": csrc_hash(), "scene": opts.get("--scene", "mesh100k")}
    for key, metric in FIELDS.items():
        if metric not in col:
            out[key] = None
            continue
        want = "mb" if key.endswith("_mb") else "ms" if key.endswith("_ms") else ""
        vals = []
        for r in picked:
            try:
                vals.append(to_unit(r[col[metric]], units[col[metric]], want))
            except ValueError:
                pass
        out[key] = sum(vals) / len(vals) if vals else None
    if out["issue_slot_busy_pct"] is not None and out["nonpred_threads"] is not None:
        out["lane_issue_frac"] = out["issue_slot_busy_pct"] / 100.0 * out["nonpred_threads"] / 32.0
    if "--shade-report" in opts and "--rays-per-launch" in opts:
        rays = float(opts["--rays-per-launch"])
        shade = dram_mb(opts["--shade-report"], "k_wf_shade")
        trace = (out["dram_read_mb"] or 0.0) + (out["dram_write_mb"] or 0.0)
        out["shade_dram_mb"] = shade
        out["rays_per_launch"] = rays
        out["dram_bytes_per_segment"] = (shade + trace) * 1e6 / rays
    json.dump(out, sys.stdout, indent=1, sort_keys=True)
    sys.stdout.write("\n")


if __name__ == "__main__":
    main()
