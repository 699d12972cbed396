#!/usr/bin/env python
"""Warp-state samples and executed instructions of k_encode_frame per device function.
usage: python profiles/r2/by_function.py X.ncu-rep [git-rev of the source the capture was built from, default HEAD]"""
import bisect, collections, csv, io, re, subprocess, sys
rep, rev = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "HEAD")
rows = list(csv.reader(io.StringIO(subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout)))
cur, hd, out = None, None, []
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif len(r) > 10 and r[0] == "Line No":
        hd = r
    elif hd and len(r) == len(hd) and r[0]:
        try:
            ie = hd.index("Instructions Executed")
            out.append((cur, int(r[0]), int(r[4]), int(r[ie]) if r[ie].isdigit() else 0, r[1].strip()[:100]))
        except ValueError:
            pass
funcs = {}
for name in ["enc_common.h", "enc_state.h", "enc_pred.h", "enc_xform.h", "enc_me.h", "enc_core.h"]:
    txt = subprocess.run(["git", "show", f"{rev}:thor_b200/csrc/enc/{name}"], capture_output=True, text=True).stdout
    funcs[name] = [(i, m.group(1)) for i, ln in enumerate(txt.splitlines(), 1)
                   for m in [re.match(r"\s*(?:template.*>\s*)?(?:TDN|TD|__device__|__global__)\s+.*?(\w+)\s*\(", ln)] if m and not ln.strip().startswith("//")]
agg, aggi = collections.Counter(), collections.Counter()
for f, l, s, n, _ in out:
    L = funcs.get(f, [])
    idx = bisect.bisect_right([x[0] for x in L], l) - 1
    fn = L[idx][1] if idx >= 0 else "?"
    agg[(f, fn)] += s
    aggi[(f, fn)] += n
tot, toti = sum(agg.values()) or 1, sum(aggi.values()) or 1
print(f"samples {tot}  warp instructions {toti}  (source at {rev})")
for k, v in agg.most_common(40):
    print(f"{100 * v / tot:5.1f}% samples  {100 * aggi[k] / toti:5.1f}% inst  {k[0]}:{k[1]}")
print("\nhottest lines:")
for o in sorted(out, key=lambda o: -o[2])[:40]:
    print(f"{100 * o[2] / tot:5.1f}%  inst {100 * o[3] / toti:4.1f}%  {o[0]}:{o[1]}  {o[4]}")
