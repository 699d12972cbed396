#!/usr/bin/env python
"""Top source lines of one kernel by warp-stall samples.
usage: ncu -i X.ncu-rep --page source --csv --kernel-name K --print-source cuda,sass > k.csv
       python profiles/hotlines.py k.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr, cur, out = None, None, []
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif len(r) > 10 and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) == len(hdr) and r[0]:
        try:
            s = int(r[4])
        except ValueError:
            continue
        ie = hdr.index("Instructions Executed")
        n = int(r[ie]) if r[ie].isdigit() else 0
        out.append((s, n, cur, r[0], r[1].strip()[:120]))
tot = sum(o[0] for o in out)
print(f"total samples {tot}, warp instructions {sum(o[1] for o in out)}")
for o in sorted(out, reverse=True)[:top]:
    print(f"{o[0]:6d} {100 * o[0] / tot:5.1f}% inst={o[1]:9d} {o[2]}:{o[3]}  {o[4]}")
