#!/usr/bin/env python
"""Aggregate an `ncu --page source --print-source cuda,sass --csv` dump by CUDA source line:
python tools/ncu_by_line.py x.csv [top] -> warp instructions, thread instructions, lanes/inst and stall samples per line,
plus totals per line range given as name=lo-hi arguments (e.g. march=160-235)."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
top = 40
ranges = []
for a in sys.argv[2:]:
    if "=" in a:
        n, r = a.split("="); f, r = r.split(":") if ":" in r else ("vt_march.cuh", r); lo, hi = r.split("-"); ranges.append((n, f, int(lo), int(hi)))
    else:
        top = int(a)
cur, hdr = None, None
agg = collections.defaultdict(lambda: [0, 0, 0, ""])
for r in rows:
    if len(r) == 2 and r[0] in ("File Path", "File Name"):
        cur = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No":
        hdr = {h: i for i, h in enumerate(r)}; continue
    if hdr is None or len(r) < 10 or not r[0]:
        continue
    try:
        wi = int(r[hdr["Instructions Executed"]]); ti = int(r[hdr["Thread Instructions Executed"]]); ss = int(r[hdr["# Samples"]])
    except (ValueError, KeyError):
        continue
    a = agg[(cur, int(r[0]))]
    a[0] += wi; a[1] += ti; a[2] += ss; a[3] = a[3] or r[1].strip()[:90]
tw = sum(a[0] for a in agg.values()); tt = sum(a[1] for a in agg.values()); ts = sum(a[2] for a in agg.values())
print("total warp inst %d, thread inst %d, lanes/inst %.2f, samples %d" % (tw, tt, tt / max(1, tw), ts))
for n, f, lo, hi in ranges:
    w = sum(a[0] for (ff, l), a in agg.items() if ff == f and lo <= l <= hi)
    t = sum(a[1] for (ff, l), a in agg.items() if ff == f and lo <= l <= hi)
    s = sum(a[2] for (ff, l), a in agg.items() if ff == f and lo <= l <= hi)
    print("  %-14s %s:%d-%d  %5.1f%% inst %5.1f%% samp lanes %4.1f" % (n, f, lo, hi, 100 * w / tw, 100 * s / max(1, ts), t / max(1, w)))
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5.1f%% inst %5.1f%% samp lanes %4.1f  %s:%d  %s" % (100 * a[0] / tw, 100 * a[2] / max(1, ts), a[1] / max(1, a[0]), f, l, a[3]))
