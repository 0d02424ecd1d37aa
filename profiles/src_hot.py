#!/usr/bin/env python
"""Top stalled SASS instructions of an `ncu --page source --csv` dump: python profiles/src_hot.py <file.csv> [n]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]; data = []
for r in rows[hi + 1:]:                      # first kernel section only
    if not r or r[0] in ("Address", "Kernel Name"):
        break
    if len(r) == len(h):
        data.append(r)
I = lambda r, c: int(r[h.index(c)] or 0)
S, N, E = "Warp Stall Sampling (All Samples)", "Warp Stall Sampling (Not-issued Samples)", "Instructions Executed"
tot, totn, tote = (sum(I(r, c) for r in data) for c in (S, N, E))
print("samples", tot, "not-issued", totn, "warp inst", tote, "sass lines", len(data))
names = [c for c in h if c.startswith("stall_")]
agg = {c: sum(I(r, c) for r in data) for c in names}
print("stall reasons:", sorted(((v, k) for k, v in agg.items() if v), reverse=True)[:8])
mx = max(I(r, E) for r in data)
hot = [r for r in data if I(r, E) > 0.5 * mx]
print("hot-loop lines", len(hot), "inst share %.3f" % (sum(I(r, E) for r in hot) / tote), "sample share %.3f" % (sum(I(r, S) for r in hot) / tot))
for r in sorted(data, key=lambda r: -I(r, N))[:int(sys.argv[2]) if len(sys.argv) > 2 else 25]:
    rs = sorted(((I(r, c), c[6:]) for c in names if I(r, c)), reverse=True)[:3]
    print(f"{I(r, N):6d} {I(r, E):9d}  {r[1].strip()[:64]:64s} {rs}")
