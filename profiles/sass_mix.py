#!/usr/bin/env python
"""Executed-instruction mix of one kernel from an `ncu --page source --csv` export:
   python profiles/sass_mix.py <source.csv> [n_units]   (n_units: rays / samples per launch, for per-unit counts)"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
n_units = float(sys.argv[2]) if len(sys.argv) > 2 else None
hdr = rows[1]
isrc, iex, ith = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
ist = {h: hdr.index(h) for h in hdr if h.startswith("stall_") and "Not Issued" not in h}
mix = collections.Counter(); tot = 0; thr = 0; stalls = collections.Counter()
for r in rows[2:]:
    if r and r[0] == "Kernel Name": break            # first kernel of the export only
    if len(r) <= ith or not r[iex].isdigit(): continue
    op = r[isrc].split()
    if not op: continue
    o = op[1] if op[0].startswith("@") else op[0]
    o = o.split(".")[0]
    n = int(r[iex]); mix[o] += n; tot += n; thr += int(r[ith])
    for h, i in ist.items():
        stalls[h] += int(r[i] or 0)
print(f"warp instructions {tot}, thread instructions {thr}" + (f", per unit {thr / n_units:.1f}" if n_units else ""))
for o, n in mix.most_common(28):
    print(f"  {o:10s} {n:12d} {100 * n / tot:5.1f}%" + (f"  {32 * n / n_units:7.1f}/unit" if n_units else ""))
s = sum(stalls.values())
print("stall samples:", ", ".join(f"{h[6:]} {100 * n / s:.1f}%" for h, n in stalls.most_common(8)))
