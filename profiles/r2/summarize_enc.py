#!/usr/bin/env python
"""Condense an `ncu --set full --import-source on` capture of k_encode_frame into the text kept under profiles/.
usage: python profiles/r2/summarize_enc.py X.ncu-rep > profiles/r2_ncu_encode_frame_<what>.txt
Prints: launch geometry and the raw counters the roofline uses, warp-state samples by stall reason, the opcode mix by
executed warp instructions and by samples, and the hottest source lines (hotlines.py's view)."""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]


def page(*args):
    return list(csv.reader(io.StringIO(subprocess.run(["ncu", "-i", rep, "--csv"] + list(args), capture_output=True, text=True).stdout)))


raw = page("--page", "raw")
hdr, units = raw[0], raw[1]
KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum", "sm__inst_executed.sum.per_cycle_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed"]
for r in raw[2:]:
    d = dict(zip(hdr, r))
    print(f"kernel {d.get('Kernel Name')}  grid {d.get('Grid Size')}  block {d.get('Block Size')}")
    for k in KEYS:
        if k in d:
            print(f"  {k:70s} {d[k]} {units[hdr.index(k)]}")
src = page("--page", "source", "--print-source", "sass")
h = src[1]
ix = {n: i for i, n in enumerate(h)}
data = [r for r in src[2:] if len(r) >= len(h)]
stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
total = sum(int(r[ix["# Samples"]]) for r in data)
inst = sum(int(r[ix["Instructions Executed"]]) for r in data)
tinst = sum(int(r[ix["Thread Instructions Executed"]]) for r in data)
print(f"\nSASS instructions {len(data)}; executed warp instructions {inst}; active threads per instruction {tinst / max(inst, 1):.1f}; samples {total}")
print("warp-state samples by reason:")
for s, v in sorted(((s, sum(int(r[ix[s]]) for r in data)) for s in stalls), key=lambda x: -x[1])[:10]:
    print(f"  {s:26s} {100 * v / total:5.1f} %")
op, ops = collections.Counter(), collections.Counter()
for r in data:
    w = [x for x in r[ix["Source"]].split() if not x.startswith("@")]
    if w:
        op[w[0].split(".")[0]] += int(r[ix["Instructions Executed"]])
        ops[w[0].split(".")[0]] += int(r[ix["# Samples"]])
print("opcode mix (share of executed warp instructions / share of samples):")
for k, v in op.most_common(16):
    print(f"  {k:10s} {100 * v / inst:5.1f} % / {100 * ops[k] / total:5.1f} %")
cu = page("--page", "source", "--print-source", "cuda,sass")
cur, hd, out = None, None, []
for r in cu:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif len(r) > 10 and r[0] == "Line No":
        hd = r
    elif hd and len(r) == len(hd) and r[0]:
        try:
            out.append((int(r[4]), cur, r[0], r[1].strip()[:110]))
        except ValueError:
            pass
tot = sum(o[0] for o in out) or 1
print("\nhottest source lines (samples, share, file:line):")
for o in sorted(out, reverse=True)[:30]:
    print(f"  {o[0]:6d} {100 * o[0] / tot:5.1f} %  {o[1]}:{o[2]}  {o[3]}")
byfile = collections.Counter()
for o in out:
    byfile[o[1]] += o[0]
print("samples by source file:")
for k, v in byfile.most_common(10):
    print(f"  {k:20s} {100 * v / tot:5.1f} %")
