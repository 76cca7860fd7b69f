#!/usr/bin/env python
"""Summarise ncu output into profiles/ (tracked):  launches CSV -> per-kernel time shares; .ncu-rep -> key counters.

    python scripts/ncu_summary.py launches gpurun_out/launches.csv  > profiles/rNN_launches.txt
    python scripts/ncu_summary.py full gpurun_out/prof.ncu-rep      > profiles/rNN_full.txt
"""
import collections
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
        "smsp__inst_executed.sum", "sm__cycles_elapsed.max"]


def launches(path):
    hdr, agg = None, collections.OrderedDict()
    for r in csv.reader(open(path)):
        if r and r[0] == "ID":
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            d = dict(zip(hdr, r))
            if d.get("Metric Name") != "gpu__time_duration.sum":
                continue
            v = float(d["Metric Value"].replace(",", ""))
            v *= {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "nsecond": 1e-3}.get(d["Metric Unit"], 1.0)
            a = agg.setdefault(d["Kernel Name"].split("(")[0][:70], [0, 0.0])
            a[0] += 1
            a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"# {path}: per-kernel device time (ncu gpu__time_duration.sum, --clock-control none; cold-cache, serialised: compare SHARES)")
    print(f"{'kernel':72s} {'n':>5s} {'total_us':>12s} {'avg_us':>10s} {'share':>6s}")
    for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{k:72s} {n:5d} {t:12.1f} {t / n:10.1f} {t / tot:6.3f}")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, units = rows[0], rows[1]
    print(f"# {path}: ncu --set full --clock-control none (one row per captured launch)")
    for row in rows[2:]:
        d = dict(zip(h, row))
        print("kernel:", d.get("Kernel Name", "?")[:100])
        for k in KEYS:
            if k in d:
                print(f"    {k:70s} {d[k]:>16s} {units[h.index(k)]}")
        rd = float(d.get("dram__bytes_read.sum", "0").replace(",", "") or 0)
        wr = float(d.get("dram__bytes_write.sum", "0").replace(",", "") or 0)
        ur, uw = units[h.index("dram__bytes_read.sum")], units[h.index("dram__bytes_write.sum")]
        print(f"    traffic = dram read {rd} {ur} + write {wr} {uw}")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
