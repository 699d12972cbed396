"""Where the hooked encoder's time goes: stock Thorenc vs integration/_build/Thorenc_tail on the bench's 1080p LDB-medium
clip, 2 and N frames, with the hook's own per-phase host clock (THORCUDA_HOOK_TIMING=1).  Run on a GPU box from the repo root:
python profiles/enc_phase_timing.py [frames]"""
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 12
d = "/dev/shm/enc_t" if os.path.isdir("/dev/shm") else "/tmp/enc_t"
os.makedirs(d, exist_ok=True)
yuv = d + "/clip.yuv"
bench.write_clip(yuv, 1920, 1080, frames)
for exe, extra in (("oracle/_ref/Thorenc", {}), ("integration/_build/Thorenc_tail", {}), ("integration/_build/Thorenc_tail", {"THORCUDA_HOOK_NOPIN": "1"})):
    for n in (2, frames):
        t0 = time.perf_counter()
        r = subprocess.run([exe, "-if", yuv, "-of", d + "/o.bit", "-rf", d + "/rec.yuv", "-stat", d + "/s.txt", "-width", "1920",
                            "-height", "1080", "-n", str(n), "-f", "60", "-qp", "32"] + bench.LDB_MEDIUM,
                           capture_output=True, text=True, env=dict(os.environ, THORCUDA_HOOK_TIMING="1", **extra))
        print(exe, extra, n, "frames", round(time.perf_counter() - t0, 2), "s", r.stderr[-700:].strip())
