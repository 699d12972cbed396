"""Link-time drop-in test (SURVEY 8b/8c "End-to-end encoder / decoder").

oracle/_ref/Thorenc_cuda and Thordec_cuda are the reference's UNMODIFIED objects linked against
libthorcuda.so instead of common_kernels.o / enc_kernels.o / transform.o, so every kernel call of the
reference (SAD family, sub-pel prediction, transforms, CLPF blocks, block average) runs on the B200 through
the per-block API of include/thorcuda_compat.h.  The bitstream, the encoder's reconstruction and the
decoder's output must be byte-identical to the stock CPU binaries'.

The per-block API is synchronous (one PCIe round trip per call), so the clips are tiny; the point is
bit-exactness of the whole encode/decode with the CUDA kernels in the loop, not speed."""
import hashlib
import json
import os

import pytest

from thorref import REF_DIR, run_thordec, run_thorenc, write_clip

pytestmark = pytest.mark.gpu


def _need(*names):
    for n in names:
        if not os.path.exists(os.path.join(REF_DIR, n)):
            pytest.skip(f"oracle/_ref/{n} not built (needs /root/reference at build time)")


@pytest.mark.parametrize("cfg,w,h,frames,qp", [
    ("LDB_low_complexity", 64, 64, 3, 32),
    ("LDB_medium_complexity", 72, 64, 3, 28),       # 72: a partial superblock column
    ("RA_high_efficiency", 64, 64, 3, 34),          # bi-pred, tb/pb split, interpolated reference, delta qp
])
def test_thorenc_thordec_linked_against_libthorcuda(tmp_path, cfg, w, h, frames, qp):
    _need("Thorenc", "Thordec", "Thorenc_cuda", "Thordec_cuda")
    yuv = str(tmp_path / "clip.yuv")
    write_clip(yuv, w, h, frames, seed=1234)
    bit_cpu, rec_cpu, t_cpu = run_thorenc("Thorenc", cfg, yuv, w, h, frames, qp, str(tmp_path / "cpu"))
    bit_gpu, rec_gpu, t_gpu = run_thorenc("Thorenc_cuda", cfg, yuv, w, h, frames, qp, str(tmp_path / "gpu"), timeout=900)
    print(f"{cfg}: {len(bit_cpu)} bytes; stock {t_cpu:.2f} s, linked against libthorcuda {t_gpu:.2f} s")
    # the committed golden hashes of the same clip (tests/golden/streams.json), when numpy reproduces the clip
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "streams.json")))
    e = gold.get(f"{cfg}_{w}x{h}x{frames}_qp{qp}_seed1234")
    if e and hashlib.md5(open(yuv, "rb").read()).hexdigest() == e["clip_md5"]:
        assert (hashlib.md5(bit_gpu).hexdigest(), hashlib.md5(rec_gpu).hexdigest()) == (e["bit_md5"], e["rec_md5"]), "golden hash differs"
    assert bit_gpu == bit_cpu, "bitstream differs"
    assert rec_gpu == rec_cpu, "encoder reconstruction differs"
    dec_cpu = run_thordec("Thordec", str(tmp_path / "cpu.bit"), str(tmp_path / "dec_cpu.yuv"))
    dec_gpu = run_thordec("Thordec_cuda", str(tmp_path / "cpu.bit"), str(tmp_path / "dec_gpu.yuv"), timeout=900)
    assert dec_gpu == dec_cpu == rec_cpu, "decoder output differs"
