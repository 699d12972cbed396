"""Round-2 driver script (run under gpurun): Thorenc_gpu against the stock Thorenc on seeded hard clips -- byte identity of
.bit and reconstruction, and the frame rate of each.  Prints one JSON line per case; appends them to gpurun_out/."""
import json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import thorref as T

GPU = os.path.join(ROOT, "integration", "_build", "Thorenc_gpu")
cases = [tuple(c.split(":")) for c in sys.argv[1:]] or [("LDB_low_complexity", "416", "240", "10", "32")]
shm = "/dev/shm" if os.path.isdir("/dev/shm") else "/tmp"
out = []
for cfg, w, h, n, qp in cases:
    w, h, n, qp = int(w), int(h), int(n), int(qp)
    yuv = os.path.join(shm, f"clip_{w}x{h}_{n}.yuv")
    if not os.path.exists(yuv):
        T.write_hard_clip(yuv, w, h, n, 77)
    res = {"config": cfg, "size": f"{w}x{h}", "frames": n, "qp": qp}
    try:
        os.environ["THORCUDA_HOOK_TIMING"] = "1"
        cmd = [GPU, "-if", yuv, "-of", shm + "/g.bit", "-rf", shm + "/g_rec.yuv", "-stat", shm + "/g_stat.txt", "-width", str(w), "-height", str(h),
               "-n", str(n), "-f", "30", "-qp", str(qp)] + T.THOR_CONFIGS[cfg]
        t0 = time.perf_counter()
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=1500)
        res["gpu_seconds"] = round(time.perf_counter() - t0, 3)
        if r.returncode != 0:
            res["gpu_error"] = (r.stdout[-600:] + r.stderr[-1200:])
        else:
            res["gpu_timing"] = [ln for ln in r.stderr.splitlines() if "thorcuda" in ln]
            b1, r1 = open(shm + "/g.bit", "rb").read(), open(shm + "/g_rec.yuv", "rb").read()
            if os.environ.get("SKIP_REF") != "1":
                b0, r0, t = T.run_thorenc("Thorenc", cfg, yuv, w, h, n, qp, shm + "/ref", timeout=3000)
                res.update(ref_seconds=round(t, 3), identical_bit=b0 == b1, identical_rec=r0 == r1, bytes=len(b0))
    except Exception as e:                                    # noqa: BLE001
        res["exception"] = repr(e)[:800]
    print(json.dumps(res), flush=True)
    out.append(res)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", "enc_gpu_check.jsonl"), "a") as f:
    for r in out:
        f.write(json.dumps(r) + "\n")
