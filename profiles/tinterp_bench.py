#!/usr/bin/env python
"""Timing of the temporal interpolation (interpolate_frames + pad_yuv_frame): libthorcuda on cuda:0 with
the two padded references resident in HBM (CUDA events on the library's stream, warm) beside the UNMODIFIED reference's
interpolate_frames (oracle/_ref/libthorref.so, one host thread as in the codec), outputs compared byte for byte.

    python profiles/tinterp_bench.py [--size 1920x1080] [--reps 20] [--no-cpu]       -> one JSON line

Measurement tooling, not product code: it imports tests/thorref.py for the picture helpers and the reference build."""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", default="1920x1080")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--ratio", type=int, default=8)
    ap.add_argument("--pos", type=int, default=3)
    ap.add_argument("--motion", default="4,-9")
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    w, h = (int(x) for x in a.size.split("x"))
    import thor_b200 as tb
    from thorref import Picture, interp_pair, load_ref, set_ref_simd

    tb.init(0)
    lib = tb.cdll()
    p0, p1 = interp_pair(w, h, seed=w + 11, motion=tuple(int(x) for x in a.motion.split(",")))
    dev = []
    for p in (p0, p1):
        d = tb.Picture(w, h, p.pad_y, p.pad_c)
        d.upload(p.plane(0), p.plane(1), p.plane(2))
        tb.ops.pad_yuv_frame(d)
        dev.append(d)
    out = tb.Picture(w, h, 96, 48)
    n0 = lib.thorcuda_launch_count()
    for _ in range(3):
        tb.ops.interpolate_frames(out, dev[0], dev[1], a.ratio, a.pos)
        tb.ops.pad_yuv_frame(out)
    tb.sync()
    per_call = (lib.thorcuda_launch_count() - n0) // 3
    t = lib.thorcuda_timer_begin()
    for _ in range(a.reps):
        tb.ops.interpolate_frames(out, dev[0], dev[1], a.ratio, a.pos)
        tb.ops.pad_yuv_frame(out)
    lib.thorcuda_timer_end(t)
    tb.sync()
    gpu_ms = lib.thorcuda_timer_elapsed_ms(t) / a.reps
    got = out.download()
    res = {"op": "interpolate_frames + pad_yuv_frame", "size": a.size, "ratio": a.ratio, "pos": a.pos, "gpu_ms": round(gpu_ms, 4),
           "gpu_launches_per_call": per_call, "reps": a.reps,
           # both references read once at full resolution + the pyramid (1/4 + 1/16 + ...), the new picture written padded
           "algorithmic_bytes": int(2 * 1.5 * w * h * 4 / 3 + p0.stride_y * (h + 2 * p0.pad_y) * 1.5)}
    res["algorithmic_GBps"] = round(res["algorithmic_bytes"] / gpu_ms / 1e6, 1)
    if not a.no_cpu:
        ref = load_ref()
        set_ref_simd(ref, 1)          # the reference's SSE4 kernels, as init_use_simd() selects on this host
        want = Picture(w, h, 96, 48)
        s0, s1, sw = p0.as_ref_struct(), p1.as_ref_struct(), want.as_ref_struct()
        t0 = time.perf_counter()
        ref.interpolate_frames(C.byref(sw), C.byref(s0), C.byref(s1), a.ratio, a.pos)
        res["cpu_reference_ms"] = round(1e3 * (time.perf_counter() - t0), 2)
        res["cpu_threads"] = 1
        res["identical"] = all(np.array_equal(got[i], want.plane(i)) for i in range(3))
        res["speedup"] = round(res["cpu_reference_ms"] / gpu_ms, 1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
