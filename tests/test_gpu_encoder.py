"""The reference ENCODER with every block decision on the B200: integration/_build/Thorenc_gpu is the
reference's encoder objects with the superblock loop, the frame tail and reference creation of encode_frame() going through
integration/enc_gpu_hook.c -> thorcuda_enc_frame / thorcuda_loopfilter_frame.  Motion search, trial encodes, rate-distortion
decisions and the quad tree run in k_encode_frame (thor_b200/csrc/encoder.cu); the host only entropy-codes the decided trees
with the reference's own write_block.  The .bit file and the reconstruction must be byte-identical to the stock encoder's
(BASELINE.json metric 1: "bit-identical .bit"), including BASELINE configs #1 (416x240 LDB low) and the 1080p geometry
(1080 = 16*64 + 56: a partial superblock row)."""
import os
import subprocess

import pytest

from thorref import REF_DIR, ROOT, THOR_CONFIGS, run_thorenc, write_hard_clip

pytestmark = pytest.mark.gpu
GPU = os.path.join(ROOT, "integration", "_build", "Thorenc_gpu")

CASES = [
    ("LDB_low_complexity", 416, 240, 10, 32),       # BASELINE.json configs[0]
    ("LDB_low_complexity", 1920, 1080, 3, 32),      # the 1080p geometry: 30 x 17 superblocks, last row 56 samples
    ("LDB_medium_complexity", 1920, 1080, 12, 32),  # configs[1] at its own size (12 of its 60 frames: 15 s of device time)
    ("RA_medium_complexity", 1920, 1080, 9, 32),    # north_star's operating point at 1080p: one sub-GOP, references interpolated on the device
    ("LDB_low_complexity", 3840, 2160, 2, 32),      # the 2160p geometry of configs[3] / [4]: 60 x 34 superblocks, last row 48 samples
    ("LDB_medium_complexity", 416, 240, 6, 32),     # configs[1]'s operating point (bi-prediction, intra RDO, telescope search)
    ("LDB_medium_complexity", 200, 136, 5, 30),     # partial superblocks on both sides
    ("LDB_medium_complexity", 128, 64, 3, 44),
    ("RA_medium_complexity", 200, 136, 17, 31),     # north_star's operating point: B hierarchy, interpolated references
    ("RA_low_complexity", 128, 128, 9, 32),
    ("HDB16_medium_complexity", 128, 72, 17, 32),
    ("LDB_high_efficiency", 128, 72, 4, 30),        # tb / pb split, delta qp, encoder_speed 0
    ("RA_high_efficiency", 64, 64, 9, 32),          # + simultaneous bi-prediction search on B frames
]


@pytest.fixture(scope="module")
def binaries():
    for need in (os.path.join(REF_DIR, "Thorenc"), GPU):
        if not os.path.exists(need):
            pytest.skip(f"{need} not built (needs /root/reference at build time)")


@pytest.mark.parametrize("cfg,w,h,frames,qp", CASES, ids=[f"{c[0]}-{c[1]}x{c[2]}x{c[3]}-qp{c[4]}" for c in CASES])
def test_gpu_encoder_is_byte_identical(tmp_path, binaries, cfg, w, h, frames, qp):
    yuv = str(tmp_path / "clip.yuv")
    write_hard_clip(yuv, w, h, frames, seed=77)
    bit0, rec0, _ = run_thorenc("Thorenc", cfg, yuv, w, h, frames, qp, str(tmp_path / "ref"))
    bit1, rec1, _ = run_thorenc(GPU, cfg, yuv, w, h, frames, qp, str(tmp_path / "gpu"))
    assert bit1 == bit0, f"{cfg}: bitstream differs ({len(bit1)} vs {len(bit0)} bytes)"
    assert rec1 == rec0, f"{cfg}: reconstruction differs"


def test_gpu_encoder_runs_its_kernels(tmp_path, binaries):
    """no silent host path: the hooked encoder reports the device time of its frames, and refuses to start without a device"""
    yuv = str(tmp_path / "clip.yuv")
    write_hard_clip(yuv, 128, 64, 3, seed=5)
    cmd = [GPU, "-if", yuv, "-of", str(tmp_path / "a.bit"), "-width", "128", "-height", "64", "-n", "3", "-f", "30", "-qp", "32"] + THOR_CONFIGS["LDB_low_complexity"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=dict(os.environ, THORCUDA_HOOK_TIMING="1"))
    assert r.returncode == 0, r.stderr[-1000:]
    assert "thorcuda frame encoder: 3 frames" in r.stderr
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=dict(os.environ, CUDA_VISIBLE_DEVICES="-1"))
    assert r.returncode != 0, "Thorenc_gpu ran without a CUDA device"


def test_enc_frame_rejects_bad_arguments():
    import ctypes as C
    import thor_b200
    thor_b200.init(0)
    lib = thor_b200.cdll()
    try:
        assert lib.thorcuda_enc_frame(None, None, None) != 0
        buf = (C.c_int32 * 64)()                     # a zeroed thorcuda_enc_params: width 0
        out = (C.c_uint8 * 64)()
        assert lib.thorcuda_enc_frame(buf, out, out) != 0
        lib.thorcuda_last_error.restype = C.c_char_p
        assert b"enc_frame" in lib.thorcuda_last_error()
    finally:
        thor_b200.shutdown()


def test_gpu_encoder_segments_join_into_the_serial_stream(tmp_path, binaries):
    """closed-GOP segments with the device encoder: three Thorenc_gpu processes share the GPU, each encodes one
    range of an RA stream (-seg_start / -seg_end, integration/enc_segment_hook.c); thor_b200/segments.py joins their chunks
    into the stock encoder's serial stream, byte for byte"""
    import sys
    sys.path.insert(0, ROOT)
    from thor_b200 import segments as sg
    w, h, frames, period = 200, 136, 37, 16
    yuv = str(tmp_path / "clip.yuv")
    write_hard_clip(yuv, w, h, frames, seed=9)
    args = ["-if", yuv, "-width", str(w), "-height", str(h), "-n", str(frames), "-f", "30", "-qp", "33"] + THOR_CONFIGS["RA_medium_complexity"] + ["-intra_period", str(period)]
    subprocess.run([os.path.join(REF_DIR, "Thorenc")] + args + ["-of", str(tmp_path / "s.bit"), "-rf", str(tmp_path / "s.yuv")], check=True, capture_output=True, timeout=600)
    ranges = sg.segment_ranges(frames, period, 3)
    assert ranges == [(0, 16), (16, 32), (32, None)]
    files = sg.encode_segments(GPU, args, ranges, str(tmp_path / "seg"), timeout=600)
    bit = sg.join_segment_streams(ranges, [open(f[0], "rb").read() for f in files], frames, 7, period)
    assert bit == (tmp_path / "s.bit").read_bytes(), "joined bitstream differs from the serial one"
    assert b"".join(open(f[1], "rb").read() for f in files) == (tmp_path / "s.yuv").read_bytes(), "joined reconstruction differs"
