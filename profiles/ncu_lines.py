#!/usr/bin/env python
"""Per-CUDA-source-line view of an .ncu-rep captured with --import-source on (read on the CPU box):
stall samples and executed warp instructions per line, plus totals for line ranges.

    python profiles/ncu_lines.py gpurun_out/x.ncu-rep [--top 30] [--file sparse_kernel.cuh] [--regions name:lo-hi,...]
"""
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 30
    fname = sys.argv[sys.argv.index("--file") + 1] if "--file" in sys.argv else None
    regions = []
    if "--regions" in sys.argv:
        for item in sys.argv[sys.argv.index("--regions") + 1].split(","):
            n, r = item.split(":")
            lo, hi = r.split("-")
            regions.append((n, int(lo), int(hi)))
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source=cuda,sass"],
                         capture_output=True, text=True).stdout
    cur, hdr, rows = None, None, []
    for r in csv.reader(io.StringIO(out)):
        if r and r[0] in ("File Path", "File Name"):
            cur = r[1].split("/")[-1]
        elif r and r[0] == "Line No":
            hdr = r
        elif hdr and cur and len(r) == len(hdr) and r[0].isdigit():
            g = lambda name: int(r[hdr.index(name)] or 0)
            rows.append((cur, int(r[0]), r[1].strip(), g("# Samples"), g("Instructions Executed"),
                         {h[6:]: int(r[i] or 0) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h and "(" not in h}))
    ts = sum(x[3] for x in rows) or 1
    ti = sum(x[4] for x in rows) or 1
    print(f"{ts} samples, {ti} warp instructions")
    agg = {}
    for x in rows:
        for k, v in x[5].items():
            agg[k] = agg.get(k, 0) + v
    print("stalls:", ", ".join(f"{k}={100 * v / ts:.1f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
    for f, l, src, sm, ie, st in sorted(rows, key=lambda x: -x[3])[:top]:
        why = ",".join(k for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:2] if v)
        print(f"{f}:{l:4d} smp {100 * sm / ts:5.1f}% inst {100 * ie / ti:5.1f}% [{why}]  {src[:90]}")
    for n, lo, hi in regions:
        sel = [x for x in rows if (fname is None or x[0] == fname) and lo <= x[1] <= hi]
        print(f"region {n:14s} lines {lo}-{hi}: smp {100 * sum(x[3] for x in sel) / ts:5.1f}% inst {100 * sum(x[4] for x in sel) / ti:5.1f}%")


if __name__ == "__main__":
    main()
