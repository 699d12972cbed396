"""The reference DECODER with libthorcuda's fused frame tail hooked in (integration/):
integration/_build/Thordec_tail = the reference's decoder objects, its decode_frame.c with the deblock / CLPF /
create_reference_frame call sites substituted at build time, and integration/dec_tail_hook.c.  Every frame's
deblocking, CLPF (bits parsed from the stream by the hook) and padded reference come from the B200; the decoded
.yuv must be byte-identical to the stock decoder's for P-frame, B-frame and interpolated-reference streams."""
import os
import subprocess

import pytest

from thorref import REF_DIR, ROOT, run_thorenc, write_clip

pytestmark = pytest.mark.gpu
BUILD = os.path.join(ROOT, "integration", "_build")


_STREAMS = {}


def _decode(exe, bit, out):
    r = subprocess.run([exe, bit, out], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    return open(out, "rb").read()


@pytest.mark.parametrize("cfg,w,h,frames,qp", [
    ("LDB_medium_complexity", 416, 240, 5, 30),     # P frames, bi-pred blocks, CLPF bits per superblock; 240 = 3*64 + 48
    ("LDB_low_complexity", 200, 136, 4, 36),        # partial superblocks right and bottom
    ("RA_high_efficiency", 128, 128, 9, 32),        # B frames, interpolated references, tb/pb split, delta qp
    ("RA_high_efficiency", 192, 136, 17, 38),       # two sub-GOPs of the B hierarchy, partial superblock row, coarse qp
    ("RA_low_complexity", 416, 240, 17, 32),        # 4-level interpolation pyramid (device-side in Thordec_gpu), 240 = 3*64 + 48
    ("HDB16_low_complexity", 256, 192, 33, 34),     # two 16-frame sub-GOPs: interpolation ratios 2..16, four B levels
    ("RA16_medium_complexity", 192, 128, 17, 30),
    ("LDB_medium_complexity", 136, 72, 14, 22),     # fine qp (dense coefficients), HQ-period frame, 33-slot ring not yet wrapped
    # the rest of the 15 shipped configurations ("all 15 configs"), small pictures
    ("LDB_high_efficiency", 128, 72, 6, 30),        # four references, tb/pb split, delta qp, full search
    ("HDB_low_complexity", 128, 128, 17, 34),
    ("HDB_medium_complexity", 128, 72, 9, 30),
    ("HDB_high_efficiency", 128, 128, 9, 33),
    ("RA_medium_complexity", 200, 136, 9, 31),
    ("HDB16_medium_complexity", 128, 72, 17, 32),
    ("HDB16_high_efficiency", 128, 72, 17, 35),
    ("RA16_low_complexity", 136, 128, 33, 36),
    ("RA16_high_efficiency", 128, 128, 17, 29),
])
@pytest.mark.parametrize("exe", ["Thordec_tail", "Thordec_gpu"])
def test_thordec_with_gpu_frame_tail(tmp_path, cfg, w, h, frames, qp, exe):
    """Thordec_tail: deblock + CLPF + reference creation on the device (2.2).  Thordec_gpu: additionally every
    prediction, inverse transform and reconstruction -- inter blocks in two launches, intra blocks in wavefront
    levels -- with the reference pictures resident on the device (2.3); the host only parses the stream."""
    TAIL = os.path.join(BUILD, exe)
    for need in (os.path.join(REF_DIR, "Thorenc"), os.path.join(REF_DIR, "Thordec"), TAIL):
        if not os.path.exists(need):
            pytest.skip(f"{need} not built (needs /root/reference at build time)")
    key = (cfg, w, h, frames, qp)
    if key not in _STREAMS:                      # one encode + stock decode per stream, shared by the two decoders
        yuv = str(tmp_path / "clip.yuv")
        write_clip(yuv, w, h, frames, seed=77)
        bit, rec, _ = run_thorenc("Thorenc", cfg, yuv, w, h, frames, qp, str(tmp_path / "enc"))
        _STREAMS[key] = (bit, rec, _decode(os.path.join(REF_DIR, "Thordec"), str(tmp_path / "enc.bit"), str(tmp_path / "cpu.yuv")))
    bit, rec, cpu = _STREAMS[key]
    with open(str(tmp_path / "in.bit"), "wb") as f:
        f.write(bit)
    gpu = _decode(TAIL, str(tmp_path / "in.bit"), str(tmp_path / "gpu.yuv"))
    assert cpu == rec, "stock decoder does not reproduce the encoder's reconstruction"
    assert gpu == cpu, "decoder output with the GPU frame tail differs"


@pytest.mark.parametrize("cfg,w,h,frames,qp", [
    ("LDB_medium_complexity", 416, 240, 4, 30),
    ("LDB_low_complexity", 200, 136, 4, 36),
    ("RA_high_efficiency", 128, 128, 9, 32),
    ("LDB_high_efficiency", 128, 72, 6, 30),
    ("HDB16_medium_complexity", 128, 72, 17, 32),
    ("RA16_low_complexity", 136, 128, 33, 36),
    ("HDB_low_complexity", 200, 136, 9, 34),
])
def test_thorenc_with_gpu_frame_tail(tmp_path, cfg, w, h, frames, qp):
    """integration/_build/Thorenc_tail: the reference ENCODER with deblocking, the CLPF decision (against the original
    frame) and reference creation on the device; bitstream and reconstruction must match the stock encoder's."""
    exe = os.path.join(BUILD, "Thorenc_tail")
    for need in (os.path.join(REF_DIR, "Thorenc"), exe):
        if not os.path.exists(need):
            pytest.skip(f"{need} not built (needs /root/reference at build time)")
    from thorref import THOR_CONFIGS
    yuv = str(tmp_path / "clip.yuv")
    write_clip(yuv, w, h, frames, seed=91)
    bit_cpu, rec_cpu, _ = run_thorenc("Thorenc", cfg, yuv, w, h, frames, qp, str(tmp_path / "cpu"))
    pre = str(tmp_path / "gpu")
    cmd = [exe, "-if", yuv, "-of", pre + ".bit", "-rf", pre + "_rec.yuv", "-stat", pre + "_stat.txt", "-width", str(w), "-height", str(h),
           "-n", str(frames), "-f", "30", "-qp", str(qp)] + THOR_CONFIGS[cfg]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    assert open(pre + ".bit", "rb").read() == bit_cpu, "bitstream differs"
    assert open(pre + "_rec.yuv", "rb").read() == rec_cpu, "reconstruction differs"


@pytest.mark.parametrize("env", [{"THORCUDA_HOOK_SERIAL": "1"}, {"THORCUDA_HOOK_SYNC_OUTPUT": "1"},
                                 {"THORCUDA_HOOK_SERIAL": "1", "THORCUDA_HOOK_SYNC_OUTPUT": "1"}, {"THORCUDA_HOOK_TIMING": "1"}])
def test_thordec_gpu_pipeline_switches(tmp_path, env):
    """Thordec_gpu keeps one frame in flight and writes the output on its own thread; the switches that turn either off
    (and the timing report) must not change a byte.  RA stream: pictures leave the decoder out of coding order."""
    exe = os.path.join(BUILD, "Thordec_gpu")
    for need in (os.path.join(REF_DIR, "Thorenc"), os.path.join(REF_DIR, "Thordec"), exe):
        if not os.path.exists(need):
            pytest.skip(f"{need} not built (needs /root/reference at build time)")
    w, h, frames = 192, 136, 25
    yuv = str(tmp_path / "clip.yuv")
    write_clip(yuv, w, h, frames, seed=5)
    run_thorenc("Thorenc", "RA_low_complexity", yuv, w, h, frames, 33, str(tmp_path / "enc"))
    cpu = _decode(os.path.join(REF_DIR, "Thordec"), str(tmp_path / "enc.bit"), str(tmp_path / "cpu.yuv"))
    r = subprocess.run([exe, str(tmp_path / "enc.bit"), str(tmp_path / "gpu.yuv")], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, **env))
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    assert open(str(tmp_path / "gpu.yuv"), "rb").read() == cpu
    assert len(cpu) == frames * w * h * 3 // 2
