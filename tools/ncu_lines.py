#!/usr/bin/env python
"""Per-source-line summary of an ncu report:  ncu -i rep --page source --print-source cuda,sass --csv > x.csv ;
python tools/ncu_lines.py x.csv [top]   -> share of warp instructions, of stall samples, active lanes per line."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur, hdr, lines = None, None, []
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No":
        hdr = {h: i for i, h in enumerate(r)}; continue
    if hdr is None or len(r) < 10 or not r[0]:
        continue
    try:
        wi = int(r[hdr["Instructions Executed"]]); ti = int(r[hdr["Thread Instructions Executed"]]); ss = int(r[hdr["# Samples"]])
    except (ValueError, KeyError):
        continue
    lines.append((wi, ti, ss, cur, r[0], r[1].strip()[:100]))
tw = sum(l[0] for l in lines); tt = sum(l[1] for l in lines); ts = sum(l[2] for l in lines)
print("total warp inst %d, lanes/inst %.2f, samples %d" % (tw, tt / max(1, tw), ts))
byfile = collections.Counter()
for l in lines:
    byfile[l[3]] += l[0]
print("by file:", {k: "%.1f%%" % (100 * v / tw) for k, v in byfile.most_common(8)})
for wi, ti, ss, f, ln, src in sorted(lines, reverse=True)[:top]:
    print("%5.1f%% inst %5.1f%% samp lanes %4.1f  %s:%s  %s" % (100 * wi / tw, 100 * ss / max(1, ts), ti / max(1, wi), f, ln, src))
