#!/usr/bin/env python
"""SASS instructions of k_encode_frame per source function (nvdisasm -gi line info; an inlined call counts for the function whose
source lines it carries -- i.e. where the code came from, which is what tells which helper is replicated most).
usage: python profiles/r2/code_size.py [libthorcuda.so]"""
import bisect, collections, glob, os, re, subprocess, sys, tempfile
lib = sys.argv[1] if len(sys.argv) > 1 else "thor_b200/libthorcuda.so"
root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
with tempfile.TemporaryDirectory() as d:
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=d, capture_output=True)
    cub = [f for f in glob.glob(d + "/*.cubin") if "encoder" in f][0]
    txt = subprocess.run(["nvdisasm", "-gi", "-c", cub], capture_output=True, text=True).stdout
funcs = {}
for f in glob.glob(root + "/thor_b200/csrc/enc/*.h") + [root + "/thor_b200/csrc/encoder.cu"]:
    funcs[os.path.basename(f)] = [(i, m.group(1)) for i, ln in enumerate(open(f), 1)
                                   for m in [re.match(r"\s*(?:template.*>\s*)?(?:TDN|TD|__device__|__global__)\s+.*?(\w+)\s*\(", ln)] if m and not ln.strip().startswith("//")]
cnt, cur, inkernel = collections.Counter(), None, False
for ln in txt.splitlines():
    if ln.startswith(".text."):
        inkernel = "k_encode_frame" in ln
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if inkernel and re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln) and cur:
        L = funcs.get(cur[0], [])
        i = bisect.bisect_right([x[0] for x in L], cur[1]) - 1
        cnt[(cur[0], L[i][1] if i >= 0 else "?")] += 1
tot = sum(cnt.values())
print(f"k_encode_frame: {tot} SASS instructions = {tot * 16 // 1024} KB")
for k, v in cnt.most_common(45):
    print(f"{v:6d} {100 * v / tot:5.1f} %  {k[0]}:{k[1]}")
