"""Diagnostic: user CPU seconds of the STOCK Thorenc on the encoder leg's clip, alone / with a CUDA context preloaded into
the process / and of Thorenc_tail.  python profiles/diag/host_slowdown.py (GPU box, repo root)"""
import os
import resource
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

d = "/dev/shm/diag"
os.makedirs(d, exist_ok=True)
n = 8
bench.write_clip(d + "/c.yuv", 1920, 1080, n, noise=5.0, obj=True)
args = ["-if", d + "/c.yuv", "-of", d + "/o.bit", "-rf", d + "/r.yuv", "-stat", d + "/s.txt", "-width", "1920", "-height", "1080", "-n", str(n),
        "-f", "60", "-qp", "32"] + bench.LDB_MEDIUM
pre = os.path.join(ROOT, "profiles", "diag", "cuda_init_preload.so")


import time


def run(exe, env, frames):
    a = list(args)
    a[a.index("-n") + 1] = str(frames)
    r0 = resource.getrusage(resource.RUSAGE_CHILDREN)
    t0 = time.perf_counter()
    r = subprocess.run([os.path.join(ROOT, exe)] + a, capture_output=True, text=True, env=dict(os.environ, THORCUDA_HOOK_TIMING="1", **env), check=True)
    wall = time.perf_counter() - t0
    r1 = resource.getrusage(resource.RUSAGE_CHILDREN)
    return {"frames": frames, "wall": round(wall, 2), "user": round(r1.ru_utime - r0.ru_utime, 2), "sys": round(r1.ru_stime - r0.ru_stime, 2),
            "hook": r.stderr.strip()[-420:]}


for rep in range(2):
    for exe, env in (("oracle/_ref/Thorenc", {}), ("oracle/_ref/Thorenc", {"LD_PRELOAD": pre}), ("integration/_build/Thorenc_tail", {}),
                     ("integration/_build/Thorenc_tail", {"THORCUDA_HOOK_NOPIN": "1"})):
        for frames in (2, n):
            print(exe.split("/")[-1], env, run(exe, env, frames), flush=True)
