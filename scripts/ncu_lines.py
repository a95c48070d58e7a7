#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv --print-source sass,cuda` export per source line:
warp instructions, thread instructions, SIMD efficiency, stall samples.  Usage: ncu_lines.py file.csv [top_n]"""
import csv, sys, collections
path = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
cur_file = None; hdr = None
agg = collections.defaultdict(lambda: [0, 0, 0, ""])
files = collections.defaultdict(lambda: [0, 0, 0])
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; hdr = None; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or cur_file is None: continue
    d = dict(zip(hdr, r))
    try:
        ln = int(d["Line No"]); inst = int(d["Instructions Executed"] or 0); th = int(d["Thread Instructions Executed"] or 0); smp = int(d["# Samples"] or 0)
    except (ValueError, KeyError):
        continue
    a = agg[(cur_file, ln)]; a[0] += inst; a[1] += th; a[2] += smp
    src = r[1] if len(r) > 1 else ""
    if src and not a[3]: a[3] = src.strip()[:90]
    f = files[cur_file]; f[0] += inst; f[1] += th; f[2] += smp
tot = sum(a[0] for a in agg.values()) or 1; tots = sum(a[2] for a in agg.values()) or 1
print(f"total warp inst {tot:,}  thread inst {sum(a[1] for a in agg.values()):,}  samples {tots:,}")
print("--- per file"); 
for f, (i, t, s) in sorted(files.items(), key=lambda kv: -kv[1][0]):
    print(f"{f:16s} inst {100 * i / tot:5.1f}%  eff {t / max(i, 1):5.1f}/32  samples {100 * s / tots:5.1f}%")
print("--- top lines by warp instructions")
for (f, ln), (i, t, s, src) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{f:14s}:{ln:4d} inst {100 * i / tot:5.2f}%  eff {t / max(i, 1):5.1f}  smp {100 * s / tots:5.2f}%  {src}")
