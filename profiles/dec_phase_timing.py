#!/usr/bin/env python
"""Host-phase table of integration/_build/Thordec_gpu (THORCUDA_HOOK_TIMING=1) on a 1080p stream encoded here by the
reference's Thorenc, pipelined (default) and serial (THORCUDA_HOOK_SERIAL=1).  Measurement tooling.
    python profiles/dec_phase_timing.py [LDB_low_complexity|RA_low_complexity] [frames] [WxH] [hard]"""
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

cfg = sys.argv[1] if len(sys.argv) > 1 else "LDB_low_complexity"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 105
W, H = (int(v) for v in (sys.argv[3] if len(sys.argv) > 3 else "1920x1080").split("x"))
enc, dec = os.path.join(ROOT, "oracle/_ref/Thorenc"), os.path.join(ROOT, "oracle/_ref/Thordec")
gpu = os.path.join(ROOT, "integration/_build/Thordec_gpu")
with tempfile.TemporaryDirectory(dir="/dev/shm" if os.path.isdir("/dev/shm") else None) as d:
    yuv, bit = os.path.join(d, "c.yuv"), os.path.join(d, "c.bit")
    hard = len(sys.argv) > 4 and sys.argv[4] == "hard"
    bench.write_clip(yuv, W, H, frames, noise=5.0 if hard else 2.0, obj=hard)
    subprocess.run([enc, "-if", yuv, "-of", bit, "-rf", os.path.join(d, "r.yuv"), "-stat", os.path.join(d, "s.txt"), "-width", str(W),
                    "-height", str(H), "-n", str(frames), "-f", "60", "-qp", "32"] + bench.DEC_CONFIGS[cfg], check=True, capture_output=True)
    outs = {}
    for name, exe, env in (("stock", dec, {}), ("pipelined", gpu, {"THORCUDA_HOOK_TIMING": "1"}),
                           ("serial", gpu, {"THORCUDA_HOOK_TIMING": "1", "THORCUDA_HOOK_SERIAL": "1"})):
        dt = 1e9
        for rep in range(4):
            out = os.path.join(d, name + ".yuv")
            t0 = time.perf_counter()
            r = subprocess.run([exe, bit, out], capture_output=True, text=True, env=dict(os.environ, **env))
            dt = min(dt, time.perf_counter() - t0)
        print(f"== {name}: {dt:.3f} s for {frames} frames ({cfg})")
        print("\n".join(l for l in r.stderr.splitlines() if "thorcuda" in l))
        outs[name] = open(out, "rb").read()
    print("identical:", outs["stock"] == outs["pipelined"] == outs["serial"])
