"""integration/apply_hooks.py names the reference's call sites it replaces and must fail loudly if they move.  Where the
reference tree is mounted (the build container), every mode is applied to the real sources and the result is checked:
each hook is present exactly where expected and nothing else changed.  No GPU, nothing is compiled."""
import difflib
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.exists(os.path.join(REF, "Makefile")), reason="reference tree not mounted")

MODES = {
    "tail": ("dec/decode_frame.c", ["thorcuda_dec_tail(decoder_info, stream, qp);", "thorcuda_dec_make_reference(decoder_info->ref[0],decoder_info->rec);"]),
    "recon": ("dec/decode_frame.c", ["thorcuda_dec_frame_begin(decoder_info);", "thorcuda_dec_flush(decoder_info, stream, qp);",
                                     "thorcuda_dec_fetch_reference(decoder_info->ref[0],decoder_info->rec);",
                                     "thorcuda_dec_interpolate(decoder_info->interp_frames[0], ref1, ref2, off1+off2, off2);"]),
    "block": ("dec/decode_block.c", ["hk_luma(", "hk_chroma(", "hk_recon_inter(rec_", "hk_recon_intra(rec_"]),
    "main": ("dec/maindec.c", ["thorcuda_dec_write_frame(&rec[op_rec_buffer_idx],width,height,outfile);", "thorcuda_dec_finish();"]),
    "enc": ("enc/encode_frame.c", ["thorcuda_enc_tail(encoder_info, stream, qp);", "thorcuda_enc_make_reference(encoder_info->ref[0],encoder_info->rec);"]),
}


@pytest.mark.parametrize("mode", sorted(MODES))
def test_apply_hooks(tmp_path, mode):
    src, needles = MODES[mode]
    out = str(tmp_path / "out.c")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "integration", "apply_hooks.py"), mode, os.path.join(REF, src), out],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    before, after = open(os.path.join(REF, src)).read(), open(out).read()
    for n in needles:
        assert n in after, n
    # the substitution is local: every changed line mentions a hook or one of the calls it replaces
    words = ("thorcuda_", "hk_", "deblock_frame", "clpf_frame", "clpf_true", "clpf_bit", "create_reference_frame", "interpolate_frames",
             "pad_yuv_frame", "get_inter_prediction", "decode_and_reconstruct_block", "write_yuv_frame", "qpc", "putbits", "sb_signal",
             "getbits", "detect_clpf", "{", "}", "stream,", "downleft_available", "bit_count_t")
    changed = [l[1:] for l in difflib.unified_diff(before.splitlines(), after.splitlines(), lineterm="", n=0)
               if l[:1] in "+-" and l[:3] not in ("+++", "---") and l[1:].strip()]
    stray = [l for l in changed if not any(w in l for w in words)]
    assert not stray, stray[:5]
    assert len(changed) < 140


def test_apply_hooks_rejects_a_changed_reference(tmp_path):
    src = open(os.path.join(REF, "dec/decode_frame.c")).read().replace("create_reference_frame(decoder_info->ref[0],decoder_info->rec);",
                                                                      "create_reference_frame(decoder_info->ref[1],decoder_info->rec);")
    p = str(tmp_path / "decode_frame.c")
    open(p, "w").write(src)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "integration", "apply_hooks.py"), "recon", p, str(tmp_path / "o.c")],
                       capture_output=True, text=True)
    assert r.returncode != 0 and "apply_hooks" in (r.stdout + r.stderr)


def test_thor_configs_match_reference():
    """tests/thorref.THOR_CONFIGS restates the reference's 15 config_*.txt files as options (the files do not travel to the
    GPU box): same options, same values, same order, for every file."""
    import glob
    from thorref import THOR_CONFIGS
    files = sorted(glob.glob("/root/reference/config_*.txt"))
    if not files:
        pytest.skip("reference tree not mounted")
    per_run = {"-if", "-of", "-rf", "-stat", "-width", "-height", "-n", "-f", "-qp"}
    assert len(files) == len(THOR_CONFIGS) == 15
    for f in files:
        opts = []
        for line in open(f, errors="replace"):
            parts = line.split(";")[0].split()
            if parts and parts[0] not in per_run:
                opts += parts[:2]
        name = os.path.basename(f)[len("config_"):-len(".txt")]
        assert THOR_CONFIGS[name] == opts, name
