"""GPU parity of the per-block compatibility API (include/thorcuda_compat.h): every symbol the reference
declares in common/common_kernels.h, enc/enc_kernels.h and common/transform.h is called in libthorcuda.so
with HOST pointers and compared bit for bit with the same symbol of the unmodified reference
(oracle/_ref/libthorref.so) -- or with the oracle port when the reference build did not travel."""
import ctypes as C

import numpy as np
import pytest

from thorref import load_oracle, load_ref, ptr, set_ref_simd

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cu():
    import thor_b200 as tb
    lib = tb.cdll()
    tb.init(0)
    for n in ("sad_calc_simd", "sad_calc_simd_unaligned", "ssd_calc_simd"):
        getattr(lib, n).restype = C.c_int
    for n in ("sad_calc_fasthalf_simd", "sad_calc_fastquarter_simd", "widesad_calc_simd"):
        getattr(lib, n).restype = C.c_uint
    yield lib
    tb.shutdown()


@pytest.fixture(scope="module")
def ref():
    r = load_ref()
    if r is None:
        pytest.skip("oracle/_ref/libthorref.so not present")
    set_ref_simd(r, 1)
    for n in ("sad_calc_fasthalf_simd", "sad_calc_fastquarter_simd", "widesad_calc_simd"):
        getattr(r, n).restype = C.c_uint
    return r


def _aligned(a):
    """a 64-byte aligned C-contiguous copy (the reference's SIMD kernels use aligned loads on compact blocks)"""
    raw = np.zeros(a.nbytes + 64, np.uint8)
    off = (-raw.ctypes.data) % 64
    out = raw[off:off + a.nbytes].view(a.dtype).reshape(a.shape)
    out[...] = a
    return out


def _frame(rng, h=200, w=256):
    """a smooth-ish random plane (so sub-pel costs have distinct minima)"""
    a = rng.integers(0, 256, (h, w), dtype=np.uint8)
    a = ((a.astype(np.uint16) + np.roll(a, 1, 0) + np.roll(a, 1, 1) + np.roll(a, 2, 1)) // 4).astype(np.uint8)
    return _aligned(a)


def _at(a, y, x):
    return C.c_void_p(a.ctypes.data + y * a.strides[0] + x)


SIZES = [(8, 8), (16, 16), (16, 8), (8, 16), (32, 32), (64, 64), (64, 32), (32, 64)]


@pytest.mark.parametrize("wh", SIZES)
def test_sad_ssd(cu, ref, wh):
    w, h = wh
    rng = np.random.default_rng(w * 100 + h)
    fr = _frame(rng)
    org = _aligned(rng.integers(0, 256, (h, w), dtype=np.uint8))
    for (y, x) in ((40, 37), (64, 64), (21, 90)):
        # sad_calc_simd_unaligned: the reference's loop advances its row pointers inside the column loop
        # (common/common_kernels.c:111-119), i.e. it is only meaningful -- and only ever called,
        # common/temporal_interp.c:471-609 -- for widths up to 16; libthorcuda rejects wider blocks
        for name in ("sad_calc_simd", "sad_calc_simd_unaligned") if w <= 16 else ("sad_calc_simd",):
            got = getattr(cu, name)(ptr(org), _at(fr, y, x), w, fr.strides[0], w, h)
            want = getattr(ref, name)(ptr(org), _at(fr, y, x), w, fr.strides[0], w, h)
            assert got == want, (name, wh, y, x)
        if w == h and x % 16 == 0:      # ssd_calc_simd is only ever called on aligned compact blocks
            assert cu.ssd_calc_simd(ptr(org), _at(fr, y, x), w, fr.strides[0], w) == ref.ssd_calc_simd(ptr(org), _at(fr, y, x), w, fr.strides[0], w)


def test_widesad(cu, ref):
    rng = np.random.default_rng(5)
    fr = _frame(rng)
    for k in range(12):
        y, x = 30 + 7 * k, 20 + 11 * k
        org = fr[y + (k % 3) - 1:y + (k % 3) - 1 + 16, x + (k % 7) - 3:x + (k % 7) - 3 + 16]
        org = _aligned(np.clip(org.astype(np.int16) + rng.integers(-3, 4, org.shape), 0, 255).astype(np.uint8))
        gx, wx = C.c_int(99), C.c_int(99)
        got = cu.widesad_calc_simd(ptr(org), _at(fr, y, x), 16, fr.strides[0], 16, 16, C.byref(gx))
        want = ref.widesad_calc_simd(ptr(org), _at(fr, y, x), 16, fr.strides[0], 16, 16, C.byref(wx))
        assert (got, gx.value) == (want, wx.value), k


@pytest.mark.parametrize("wh", SIZES)
def test_fasthalf_fastquarter(cu, ref, wh):
    w, h = wh
    rng = np.random.default_rng(w * 7 + h)
    fr = _frame(rng)
    for k in range(6):
        y, x = 24 + 9 * k, 16 + 13 * k
        org = _aligned(np.clip(fr[y:y + h, x:x + w].astype(np.int16) + rng.integers(-4, 5, (h, w)), 0, 255).astype(np.uint8))
        gx, gy, wx, wy = C.c_int(0), C.c_int(0), C.c_int(0), C.c_int(0)
        got = cu.sad_calc_fasthalf_simd(ptr(org), _at(fr, y, x), w, fr.strides[0], w, h, C.byref(gx), C.byref(gy))
        want = ref.sad_calc_fasthalf_simd(ptr(org), _at(fr, y, x), w, fr.strides[0], w, h, C.byref(wx), C.byref(wy), None)
        assert (got, gx.value, gy.value) == (want, wx.value, wy.value), ("half", wh, k)
        for hx, hy in ((0, 0), (2, 0), (0, 2), (2, 2), (-2, -2), (-2, 0), (0, -2), (2, -2)):
            gx, gy, wx, wy = C.c_int(hx), C.c_int(hy), C.c_int(hx), C.c_int(hy)
            got = cu.sad_calc_fastquarter_simd(ptr(org), _at(fr, y, x), w, fr.strides[0], w, h, C.byref(gx), C.byref(gy))
            want = ref.sad_calc_fastquarter_simd(ptr(org), _at(fr, y, x), w, fr.strides[0], w, h, C.byref(wx), C.byref(wy))
            assert (got, gx.value, gy.value) == (want, wx.value, wy.value), ("quarter", wh, k, hx, hy)


@pytest.mark.parametrize("bipred", [0, 1])
def test_inter_prediction_luma(cu, ref, bipred):
    rng = np.random.default_rng(11 + bipred)
    fr = np.ascontiguousarray(rng.integers(0, 256, (160, 192), dtype=np.uint8))
    for (w, h) in ((8, 8), (16, 16), (32, 16), (64, 64), (8, 4), (4, 8)):
        for yoff in range(4):
            for xoff in range(4):
                if not xoff and not yoff:
                    continue
                got = _aligned(np.zeros((h, 64 + 16), np.uint8))
                want = _aligned(np.zeros((h, 64 + 16), np.uint8))
                cu.get_inter_prediction_luma_simd(w, h, xoff, yoff, ptr(got), got.strides[0], _at(fr, 40, 50), fr.strides[0], bipred)
                ref.get_inter_prediction_luma_simd(w, h, xoff, yoff, ptr(want), want.strides[0], _at(fr, 40, 50), fr.strides[0], bipred)
                assert np.array_equal(got[:, :w], want[:, :w]), (w, h, xoff, yoff)
                assert not got[:, w:].any()          # nothing written beyond the block


def test_inter_prediction_chroma(cu, ref):
    rng = np.random.default_rng(13)
    fr = np.ascontiguousarray(rng.integers(0, 256, (96, 128), dtype=np.uint8))
    for (w, h) in ((4, 4), (8, 8), (16, 8), (32, 32), (4, 8)):
        for yoff in range(8):
            for xoff in range(8):
                got = _aligned(np.zeros((h, 48), np.uint8))
                want = _aligned(np.zeros((h, 48), np.uint8))
                cu.get_inter_prediction_chroma_simd(w, h, xoff, yoff, ptr(got), got.strides[0], _at(fr, 30, 40), fr.strides[0])
                ref.get_inter_prediction_chroma_simd(w, h, xoff, yoff, ptr(want), want.strides[0], _at(fr, 30, 40), fr.strides[0])
                assert np.array_equal(got, want), (w, h, xoff, yoff)


@pytest.mark.parametrize("n", [4, 8, 16, 32, 64])
def test_transforms(cu, ref, n):
    rng = np.random.default_rng(n)
    q = min(n, 16)
    cases = [rng.integers(-255, 256, (n, n)), np.full((n, n), 255), np.full((n, n), -255),
             255 * (1 - 2 * ((np.add.outer(np.arange(n), np.arange(n))) & 1)), rng.integers(-20, 21, (n, n))]
    for blk in cases:
        blk = _aligned(np.ascontiguousarray(blk, np.int16))
        for fast in ((0, 1) if n > 16 else (0,)):
            for fwd in ("transform_simd", "transform"):
                got = _aligned(np.full((n, n), 7777, np.int16))
                want = _aligned(np.full((n, n), 7777, np.int16))
                getattr(cu, fwd)(ptr(blk), ptr(got), n, fast)
                getattr(ref, fwd)(ptr(blk), ptr(want), n, fast)
                assert np.array_equal(got[:q, :q], want[:q, :q]), (fwd, n, fast)
        # inverse on quantiser-like sparse coefficients and on dense ones (low q x q corner)
        for dens in (0.1, 1.0):
            co = _aligned(np.zeros((n, n), np.int16))
            co[:q, :q] = rng.integers(-600, 601, (q, q)) * (rng.random((q, q)) < dens)
            # inverse_transform_simd handles sizes up to 32; 64 exists only in the dispatcher (common/transform.c:496-516)
            for inv in ("inverse_transform_simd", "inverse_transform") if n < 64 else ("inverse_transform",):
                got = _aligned(np.zeros((n, n), np.int16))
                want = _aligned(np.zeros((n, n), np.int16))
                getattr(cu, inv)(ptr(co), ptr(got), n)
                getattr(ref, inv)(ptr(co), ptr(want), n)
                assert np.array_equal(got, want), (inv, n, dens)


def test_clpf_blocks_and_detect(cu, ref):
    rng = np.random.default_rng(17)
    # clpf_frame only visits superblocks that lie completely inside the frame (common/common_frame.c:496-497);
    # the reference's SIMD and C versions disagree outside them, so blocks of partial SBs are not part of the contract
    W, H = 136, 72
    rec = _aligned((rng.integers(0, 256, (H, W)) // 8 * 8 + rng.integers(0, 4, (H, W))).astype(np.uint8))
    org = _aligned(rng.integers(0, 256, (H, W), dtype=np.uint8))
    for (fn, size, sb) in (("clpf_block8", 8, 64), ("clpf_block4", 4, 32)):
        for y0 in range(0, (H // sb) * sb, size):
            for x0 in range(0, (W // sb) * sb, size * 3):
                got = _aligned(np.zeros((sb, sb), np.uint8))
                want = _aligned(np.zeros((sb, sb), np.uint8))
                getattr(cu, fn)(ptr(rec), ptr(got), W, sb, x0, y0, W, H)
                getattr(ref, fn)(ptr(rec), ptr(want), W, sb, x0, y0, W, H)
                assert np.array_equal(got, want), (fn, x0, y0)
    for y0 in range(0, (H // 64) * 64, 8):
        for x0 in range(0, (W // 64) * 64, 24):
            g0, g1, w0, w1 = C.c_int(5), C.c_int(9), C.c_int(5), C.c_int(9)      # the sums ACCUMULATE
            cu.detect_clpf_simd(ptr(rec), ptr(org), x0, y0, W, H, W, W, C.byref(g0), C.byref(g1))
            ref.detect_clpf_simd(ptr(rec), ptr(org), x0, y0, W, H, W, W, C.byref(w0), C.byref(w1))
            assert (g0.value, g1.value) == (w0.value, w1.value), (x0, y0)


def test_block_avg(cu, ref):
    rng = np.random.default_rng(19)
    a = np.ascontiguousarray(rng.integers(0, 256, (64, 80), dtype=np.uint8))
    b = np.ascontiguousarray(rng.integers(0, 256, (64, 96), dtype=np.uint8))
    for (w, h) in ((4, 4), (8, 8), (4, 8), (8, 4), (16, 16)):
        got = _aligned(np.zeros((h, 16), np.uint8))
        want = _aligned(np.zeros((h, 16), np.uint8))
        cu.block_avg_simd(ptr(got), _at(a, 3, 5), _at(b, 7, 9), 16, a.strides[0], b.strides[0], w, h)
        ref.block_avg_simd(ptr(want), _at(a, 3, 5), _at(b, 7, 9), 16, a.strides[0], b.strides[0], w, h)
        assert np.array_equal(got, want), (w, h)


def test_port_agrees_when_reference_absent(cu):
    """The same checks against the oracle port for one representative of each family, so the file still
    tests something on a box where oracle/_ref did not travel."""
    orc = load_oracle()
    orc.orc_sad.restype = C.c_uint
    rng = np.random.default_rng(23)
    fr = _frame(rng)
    org = np.ascontiguousarray(rng.integers(0, 256, (16, 16), dtype=np.uint8))
    assert cu.sad_calc_simd(ptr(org), _at(fr, 33, 41), 16, fr.strides[0], 16, 16) == orc.orc_sad(ptr(org), _at(fr, 33, 41), 16, fr.strides[0], 16, 16)
    blk = np.ascontiguousarray(rng.integers(-255, 256, (16, 16)), np.int16)
    got, want = np.zeros((16, 16), np.int16), np.zeros((16, 16), np.int16)
    cu.transform(ptr(blk), ptr(got), 16, 0)
    orc.orc_transform(ptr(blk), ptr(want), 16, 0)
    assert np.array_equal(got, want)
