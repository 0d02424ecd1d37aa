#!/usr/bin/env python
"""Static opcode histogram of one kernel of libsunsky_b200.so (whole function, cold paths included):
   python profiles/sass_static.py <substring of the mangled name> [top-n]"""
import collections, os, re, subprocess, sys
so = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "mitsuba3-sunsky_b200", "libsunsky_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
on = False; c = collections.Counter()
for line in txt.splitlines():
    if "Function :" in line:
        on = sys.argv[1] in line
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if on and m:
        c[m.group(2)] += 1
print("total", sum(c.values()))
for o, n in c.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 30):
    print(f"{n:6d} {o}")
