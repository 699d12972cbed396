"""CPU checks of the frame encoder's DECISION LOGIC (process_block and everything below it, re-hosted for the
device in thor_b200/csrc/enc/).  integration/_build/Thorenc_emu is the reference encoder with its superblock loop replaced
by that very source compiled for the host (tests/hostemu: one thread stands in for the CTA, barriers are no-ops) and the
decided trees written by the reference's own write_block.  Its .bit and reconstruction must be byte-identical to the stock
encoder's.  What this cannot see -- races between the threads of a CTA, the wavefront hand-over -- is covered by
tests/test_gpu_encoder.py on the B200."""
import os

import pytest

from thorref import REF_DIR, ROOT, THOR_CONFIGS, run_thorenc, write_hard_clip

EMU = os.path.join(ROOT, "integration", "_build", "Thorenc_emu")

# (config, width, height, frames, qp): sized so that the whole file runs in about a minute on one core
CASES = [
    ("LDB_low_complexity", 128, 64, 4, 32),         # P frames, early skip, 2 references, no bi-prediction
    ("LDB_low_complexity", 200, 136, 3, 26),        # partial superblocks right and bottom: rectangular skip blocks
    ("LDB_medium_complexity", 136, 72, 5, 30),      # telescope + widesad search, bi-prediction search, intra RDO
    ("LDB_medium_complexity", 128, 64, 3, 44),      # coarse qp: large early-skip regions, 64x64 blocks
    ("RA_low_complexity", 128, 128, 9, 32),         # B frames, interpolated reference, signs of future references
    ("HDB_medium_complexity", 128, 72, 9, 30),
    ("HDB16_low_complexity", 128, 72, 17, 34),      # 16-picture hierarchy
    ("LDB_high_efficiency", 64, 64, 3, 30),         # tb / pb split, delta qp search, encoder_speed 0, 4 references
    ("RA_high_efficiency", 64, 64, 9, 32),          # + motion_estimate_bi on B frames
    # the remaining shipped operating points, so that all 15 configuration files pass through the device decision source
    ("RA_medium_complexity", 128, 72, 9, 33),
    ("HDB_low_complexity", 128, 64, 5, 32),
    ("HDB_high_efficiency", 64, 64, 5, 30),
    ("HDB16_medium_complexity", 64, 64, 17, 32),
    ("HDB16_high_efficiency", 64, 64, 17, 34),
    ("RA16_low_complexity", 128, 64, 17, 32),
    ("RA16_medium_complexity", 64, 64, 17, 30),
    ("RA16_high_efficiency", 64, 64, 17, 33),
]


@pytest.fixture(scope="module")
def binaries():
    for need in (os.path.join(REF_DIR, "Thorenc"), EMU):
        if not os.path.exists(need):
            pytest.skip(f"{need} not built (needs /root/reference at build time)")


@pytest.mark.parametrize("cfg,w,h,frames,qp", CASES, ids=[f"{c[0]}-{c[1]}x{c[2]}x{c[3]}-qp{c[4]}" for c in CASES])
def test_emulated_encoder_is_byte_identical(tmp_path, binaries, cfg, w, h, frames, qp):
    yuv = str(tmp_path / "clip.yuv")
    write_hard_clip(yuv, w, h, frames, seed=77)
    bit0, rec0, _ = run_thorenc("Thorenc", cfg, yuv, w, h, frames, qp, str(tmp_path / "ref"))
    bit1, rec1, _ = run_thorenc(EMU, cfg, yuv, w, h, frames, qp, str(tmp_path / "emu"))
    assert bit1 == bit0, f"{cfg}: bitstream differs ({len(bit1)} vs {len(bit0)} bytes)"
    assert rec1 == rec0, f"{cfg}: reconstruction differs"


def test_every_shipped_configuration_is_accepted():
    """the device code has no unsupported shipped operating point: the binding only refuses -sync 1 and -rdoq 1"""
    src = open(os.path.join(ROOT, "integration", "enc_gpu_hook.c")).read()
    body = src[src.index("static void check_config"):src.index("#ifdef THOR_EMU")]
    assert body.count("fatalerror(") == 3 and "sync" in body and "rdoq" in body
    for name, opts in THOR_CONFIGS.items():
        assert "-sync" not in opts and "-rdoq" not in opts, name
