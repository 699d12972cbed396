"""Which coded frame of Thorenc_gpu's stream first differs from the stock encoder's (frames are 4-byte-length-prefixed chunks,
enc/putbits.c:57-95).  usage: first_diff.py <config> <w> <h> <frames> <qp>"""
import os, struct, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import thorref as T
cfg, w, h, n, qp = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
yuv = f"/dev/shm/clip_{w}x{h}_{n}.yuv"
if not os.path.exists(yuv):
    T.write_hard_clip(yuv, w, h, n, 77)
b0, r0, _ = T.run_thorenc("Thorenc", cfg, yuv, w, h, n, qp, "/dev/shm/fd_ref")
b1, r1, _ = T.run_thorenc(os.path.join(ROOT, "integration", "_build", "Thorenc_gpu"), cfg, yuv, w, h, n, qp, "/dev/shm/fd_gpu")
def chunks(b):
    out, i = [], 0
    while i + 4 <= len(b):
        ln = struct.unpack(">I", b[i:i + 4])[0]
        out.append(b[i + 4:i + 4 + ln]); i += 4 + ln
    return out
c0, c1 = chunks(b0), chunks(b1)
fs = w * h * 3 // 2
bad_rec = [k for k in range(len(r0) // fs) if r0[k * fs:(k + 1) * fs] != r1[k * fs:(k + 1) * fs]]
bad = [k for k in range(min(len(c0), len(c1))) if c0[k] != c1[k]]
print(cfg, f"{w}x{h}", "chunks", len(c0), len(c1), "differing coded frames", bad[:8], "differing display frames", bad_rec[:8])
