#!/usr/bin/env python
"""Per-source-line hot spots of one kernel from an .ncu-rep captured with --import-source on (needs -lineinfo).

    python scripts/ncu_lines.py gpurun_out/prof.ncu-rep k_validate [top_n [launch_index]]

Aggregates the 'cuda,sass' source page: warp-stall samples and executed warp instructions per (file, line), plus the
dominant stall reasons of each line.
"""
import collections
import csv
import subprocess
import sys


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    which = int(sys.argv[4]) + 1 if len(sys.argv) > 4 else 1    # n-th captured launch that matches (0-based on the command line)
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-name", "regex:" + kern],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    fname, hdr = "?", None
    agg = collections.OrderedDict()
    kernels_seen = 0
    for r in rows:
        if not r:
            continue
        if r[0] == "Kernel Name":
            kernels_seen += 1
            if kernels_seen > which:
                break           # one captured launch only
            continue
        if kernels_seen and kernels_seen != which:   # reports without 'Kernel Name' rows aggregate every captured launch
            continue
        if r[0] == "File Name":
            fname = (r[1] or "?").split("/")[-1]
            continue
        if r[0] == "Line No":
            hdr = r
            continue
        if hdr is None or len(r) < len(hdr) or not r[0].isdigit():
            continue
        d = dict(zip(hdr[4:], r[4:]))
        key = (fname, int(r[0]))
        a = agg.setdefault(key, {"src": r[1], "samples": 0, "inst": 0, "stalls": collections.Counter()})
        try:
            a["samples"] += int(d["# Samples"] or 0)
            a["inst"] += int(d["Instructions Executed"] or 0)
        except ValueError:
            continue
        for k, v in d.items():
            if k.startswith("stall_") and "Not Issued" not in k and v not in ("", "0"):
                a["stalls"][k[6:]] += int(v)
    ts = sum(a["samples"] for a in agg.values()) or 1
    ti = sum(a["inst"] for a in agg.values()) or 1
    print(f"# {rep} kernel~{kern}: {ts} samples, {ti} warp instructions (captured launch {which - 1} of the match)")
    print(f"{'file:line':28s} {'samp%':>6s} {'inst%':>6s}  top stalls | source")
    for (f, ln), a in sorted(agg.items(), key=lambda x: -x[1]["samples"])[:top]:
        st = " ".join(f"{k}:{v}" for k, v in a["stalls"].most_common(3))
        print(f"{f + ':' + str(ln):28s} {100 * a['samples'] / ts:6.2f} {100 * a['inst'] / ti:6.2f}  {st:40s} | {a['src'].strip()[:90]}")


if __name__ == "__main__":
    main()
