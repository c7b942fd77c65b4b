#!/usr/bin/env python3
"""Per-CUDA-source-line instruction counts and stall samples from `ncu --page source --csv
--print-source cuda,sass` output.  usage: ncu_lines.py file.csv [min_percent]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
hdr = None
out, tot, tots = [], 0, 0
fname = ""
for r in rows:
    if r and r[0] == "File Path":
        fname = r[1].split("/")[-1]
    if r and r[0] == "Line No":
        hdr = r
        ix = hdr.index("Instructions Executed"); isamp = hdr.index("# Samples")
        continue
    if hdr is None or len(r) <= ix or not r[0]:
        continue
    try:
        n = int(r[ix]); s = int(r[isamp])
    except ValueError:
        continue
    tot += n; tots += s
    out.append((fname, int(r[0]), r[1].strip()[:100], n, s))
print(f"total warp instructions {tot}, samples {tots}")
for f, ln, src, n, s in out:
    if n > tot * thr / 100 or s > tots * thr / 100:
        print(f"{n / tot * 100:5.1f}% inst {s / max(tots, 1) * 100:5.1f}% smp  {f}:{ln:<4} {src}")
