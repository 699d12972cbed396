"""Prints the Thorenc_gpu command line for a seeded hard clip (writes the clip to /dev/shm first): used to wrap the encoder
in ncu under gpurun.  usage: enc_cmd.py <config> <w> <h> <frames> <qp>"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import thorref as T
cfg, w, h, n, qp = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
yuv = f"/dev/shm/clip_{w}x{h}_{n}.yuv"
if not os.path.exists(yuv):
    T.write_hard_clip(yuv, w, h, n, 77)
print(" ".join([os.path.join(ROOT, "integration", "_build", "Thorenc_gpu"), "-if", yuv, "-of", "/dev/shm/p.bit", "-rf", "/dev/shm/p_rec.yuv", "-stat",
                "/dev/shm/p_stat.txt", "-width", str(w), "-height", str(h), "-n", str(n), "-f", "30", "-qp", str(qp)] + T.THOR_CONFIGS[cfg]))
