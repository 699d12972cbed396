"""The reference DECODER with libthorcuda's fused frame tail hooked in (integration/):
integration/_build/Thordec_tail = the reference's decoder objects, its decode_frame.c with the deblock / CLPF /
create_reference_frame call sites substituted at build time, and integration/dec_tail_hook.c.  Every frame's
deblocking, CLPF (bits parsed from the stream by the hook) and padded reference come from the B200; the decoded
.yuv must be byte-identical to the stock decoder's for P-frame, B-frame and interpolated-reference streams."""
import os
import subprocess

import pytest

import hashlib
import json

from thorref import DECODER_STREAMS, REF_DIR, ROOT, run_thorenc, write_clip

pytestmark = pytest.mark.gpu
BUILD = os.path.join(ROOT, "integration", "_build")


_STREAMS = {}


def _decode(exe, bit, out):
    r = subprocess.run([exe, bit, out], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    return open(out, "rb").read()


@pytest.mark.parametrize("cfg,w,h,frames,qp", DECODER_STREAMS)
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
        _STREAMS[key] = (bit, rec, _decode(os.path.join(REF_DIR, "Thordec"), str(tmp_path / "enc.bit"), str(tmp_path / "cpu.yuv")),
                         hashlib.md5(open(yuv, "rb").read()).hexdigest())
    bit, rec, cpu, clip_md5 = _STREAMS[key]
    with open(str(tmp_path / "in.bit"), "wb") as f:
        f.write(bit)
    gpu = _decode(TAIL, str(tmp_path / "in.bit"), str(tmp_path / "gpu.yuv"))
    # the committed hashes of the same stream (tests/golden/streams.json, written where the reference tree is mounted),
    # when this numpy build reproduces the clip: the stream the box's Thorenc made is the golden one, and so is our output
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "streams.json"))).get(f"{cfg}_{w}x{h}x{frames}_qp{qp}_seed77")
    assert gold is not None, "stream missing from tests/golden/streams.json: rerun tests/golden/make_golden.py"
    if clip_md5 == gold["clip_md5"]:
        assert hashlib.md5(bit).hexdigest() == gold["bit_md5"], "the box's Thorenc does not reproduce the golden stream"
        assert hashlib.md5(gpu).hexdigest() == gold["rec_md5"], "decoder output differs from the golden reconstruction"
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
