"""Pins oracle/orc_*.c (our C restatement) to the unmodified reference compiled into
oracle/_ref/libthorref.so, on both of the reference's code paths (use_simd = 1 and 0).
Skipped when the reference build is absent (it cannot be rebuilt without /root/reference)."""
import ctypes as C

import numpy as np
import pytest

from thorref import (BLOCKINFO, COEFF_KINDS, Picture, blocky_frame, YuvFrame, coeff_block, interp_pair, load_oracle, load_ref, ptr,
                     random_blockinfo, ref_bitpos, ref_stream, set_ref_simd, smooth_frame)

ref = load_ref()
pytestmark = pytest.mark.skipif(ref is None, reason="oracle/_ref/libthorref.so not built")
orc = load_oracle()
orc.orc_dct_matrix.restype = C.c_void_p
orc.orc_zigzag.restype = C.c_void_p


@pytest.fixture(params=[1, 0], ids=["simd", "plainc"])
def simd(request):
    set_ref_simd(ref, request.param)
    return request.param


def test_tables():
    for n, name in ((4, "g1mat_hevc"), (8, "g2mat_hevc"), (16, "g3mat_hevc"), (32, "g4mat_hevc")):
        want = np.ctypeslib.as_array((C.c_int16 * (n * n)).in_dll(ref, name))
        got = np.ctypeslib.as_array((C.c_int16 * (n * n)).from_address(orc.orc_dct_matrix(n)))
        assert np.array_equal(want, got), name
    for q, name in ((4, "zigzag16"), (8, "zigzag64"), (16, "zigzag256")):
        want = np.ctypeslib.as_array((C.c_int * (q * q)).in_dll(ref, name))
        got = np.ctypeslib.as_array((C.c_int * (q * q)).from_address(orc.orc_zigzag(q)))
        assert np.array_equal(want, got), name
    cq = np.ctypeslib.as_array((C.c_int * 52).in_dll(ref, "chroma_qp"))
    assert [orc.orc_chroma_qp(q) for q in range(52)] == list(cq)


@pytest.mark.parametrize("wh", [(416, 240), (192, 136)])
@pytest.mark.parametrize("qp", [0, 20, 32, 51])
def test_deblock(simd, wh, qp):
    w, h = wh
    rng = np.random.default_rng(qp * 7 + w)
    y, u, v = blocky_frame(w, h, seed=qp + 1)
    a = Picture(w, h)
    a.set_visible(y, u, v)
    b = a.copy()
    bi = random_blockinfo(w, h, rng)
    fa = a.as_ref_struct()
    qpc = orc.orc_chroma_qp(qp)
    ref.deblock_frame_y(C.byref(fa), ptr(bi), w, h, C.c_uint8(qp))
    ref.deblock_frame_uv(C.byref(fa), ptr(bi), w, h, C.c_uint8(qpc))
    orc.orc_deblock_frame_y(C.c_void_p(b.origin_ptr(0)), b.stride_y, ptr(bi), w, h, qp)
    orc.orc_deblock_frame_uv(C.c_void_p(b.origin_ptr(1)), C.c_void_p(b.origin_ptr(2)), b.stride_c, ptr(bi), w, h, qpc)
    for i in range(3):
        assert np.array_equal(a.plane(i), b.plane(i)), f"plane {i}"
    if qp >= 20:
        assert not np.array_equal(a.plane(0), y), "filter never fired: weak test"


DECISION = C.CFUNCTYPE(C.c_int, C.c_int, C.c_int, C.POINTER(YuvFrame), C.POINTER(YuvFrame), C.c_void_p, C.c_int, C.c_void_p)


def _ref_clpf(pic, org, bi, w, h, bits_in, simd):
    """Drive the reference clpf_frame with a Python decision callback equal to enc/encode_frame.c:50-63
    (org given) or dec/decode_frame.c:130-133 (bits given).  Returns the per-SB bits."""
    nsbx = w // 64
    out = np.full((h // 64) * nsbx, 0xFF, np.uint8)
    bia = bi.reshape(h // 4, w // 4)
    detect = ref.detect_clpf_simd if simd else ref.detect_clpf

    def decide(k, l, rec, o, dd, bs, stream):
        if org is None:
            on = int(bits_in[k * nsbx + l])
        else:
            s0, s1 = C.c_int(0), C.c_int(0)
            for m in range(8):
                for n in range(8):
                    e = bia[k * 16 + m * 2, l * 16 + n * 2]
                    if e["cbp_y"] and e["mode"] != 3:
                        detect(C.c_void_p(pic.origin_ptr(0)), C.c_void_p(org.origin_ptr(0)), l * 64 + n * 8, k * 64 + m * 8,
                               w, h, org.stride_y, pic.stride_y, C.byref(s0), C.byref(s1))
            on = int(s1.value < s0.value)
        out[k * nsbx + l] = on
        return on

    f = pic.as_ref_struct()
    fo = (org or pic).as_ref_struct()
    ref.clpf_frame(C.byref(f), C.byref(fo), ptr(bi), None, DECISION(decide))
    return out


@pytest.mark.parametrize("wh", [(416, 240), (256, 192)])
@pytest.mark.parametrize("enc", [True, False])
def test_clpf_frame(simd, wh, enc):
    w, h = wh
    rng = np.random.default_rng(w + enc)
    y, u, v = blocky_frame(w, h, seed=3)
    org = Picture(w, h)
    org.set_visible(y, u, v)
    a = Picture(w, h)
    noisy = [np.clip(p.astype(int) + rng.integers(-3, 4, p.shape), 0, 255).astype(np.uint8) for p in (y, u, v)]
    a.set_visible(*noisy)
    b = a.copy()
    bi = random_blockinfo(w, h, rng)
    nsb = (w // 64) * (h // 64)
    bits_in = rng.integers(0, 2, nsb).astype(np.uint8)
    want_bits = _ref_clpf(a, org if enc else None, bi, w, h, bits_in, simd)
    got_bits = np.zeros(nsb, np.uint8)
    orc.orc_clpf_frame(C.c_void_p(b.origin_ptr(0)), C.c_void_p(b.origin_ptr(1)), C.c_void_p(b.origin_ptr(2)),
                       b.stride_y, b.stride_c, C.c_void_p(org.origin_ptr(0)) if enc else None, org.stride_y,
                       ptr(bi), w, h, ptr(bits_in), ptr(got_bits))
    assert np.array_equal(want_bits, got_bits)
    for i in range(3):
        assert np.array_equal(a.plane(i), b.plane(i)), f"plane {i}"
    assert not np.array_equal(a.plane(0), noisy[0])


def test_pad_and_reference():
    w, h = 416, 240
    rng = np.random.default_rng(5)
    rec = Picture(w, h, rng=rng)
    a = Picture(w, h, 96, 48, rng=rng)
    b = a.copy()
    fr, fa = rec.as_ref_struct(), a.as_ref_struct()
    ref.create_reference_frame(C.byref(fa), C.byref(fr))
    for i in range(3):
        b.plane(i)[...] = rec.plane(i)
    orc.orc_pad_plane(C.c_void_p(b.origin_ptr(0)), b.stride_y, w, h, 96, 96)
    orc.orc_pad_plane(C.c_void_p(b.origin_ptr(1)), b.stride_c, w // 2, h // 2, 48, 48)
    orc.orc_pad_plane(C.c_void_p(b.origin_ptr(2)), b.stride_c, w // 2, h // 2, 48, 48)
    assert np.array_equal(a.buf_y, b.buf_y) and np.array_equal(a.buf_u, b.buf_u) and np.array_equal(a.buf_v, b.buf_v)


class Mv(C.Structure):
    _fields_ = [("x", C.c_int16), ("y", C.c_int16)]


@pytest.mark.parametrize("bipred", [0, 1])
def test_inter_luma_all_phases(simd, bipred):
    rng = np.random.default_rng(11)
    refp = Picture(128, 128, 96, 48, rng=rng)
    refp.buf_y[::7, ::5] = 255
    refp.buf_y[3::11, 1::3] = 0
    for (w, h) in ((4, 4), (8, 8), (16, 16), (32, 32), (64, 64), (16, 8), (8, 16), (4, 8), (64, 24)):
        for fy in range(4):
            for fx in range(4):
                for sign in (0, 1):
                    mv = Mv(int(rng.integers(-6, 6)) * 4 + fx, int(rng.integers(-6, 6)) * 4 + fy)
                    base = refp.origin_ptr(0) + 32 * refp.stride_y + 40
                    a = np.zeros((h, w), np.uint8)
                    b = np.zeros((h, w), np.uint8)
                    ref.get_inter_prediction_luma(ptr(a), C.c_void_p(base), w, h, refp.stride_y, w, C.byref(mv), sign, bipred)
                    orc.orc_inter_luma(ptr(b), C.c_void_p(base), w, h, refp.stride_y, w, mv.x, mv.y, sign, bipred)
                    assert np.array_equal(a, b), (w, h, fx, fy, sign)


def test_inter_chroma_all_phases(simd):
    rng = np.random.default_rng(12)
    refp = Picture(128, 128, 96, 48, rng=rng)
    refp.buf_u[::3, ::4] = 255
    refp.buf_u[1::5, 2::7] = 0
    for (w, h) in ((2, 2), (4, 4), (8, 8), (16, 16), (32, 32), (8, 4), (4, 8), (2, 4), (32, 12)):
        for fy in range(8):
            for fx in range(8):
                sign = (fx + fy) & 1
                mv = Mv(int(rng.integers(-4, 4)) * 8 + fx, int(rng.integers(-4, 4)) * 8 + fy)
                base = refp.origin_ptr(1) + 16 * refp.stride_c + 24
                a = np.zeros((h, w), np.uint8)
                b = np.zeros((h, w), np.uint8)
                ref.get_inter_prediction_chroma(ptr(a), C.c_void_p(base), w, h, refp.stride_c, w, C.byref(mv), sign)
                orc.orc_inter_chroma(ptr(b), C.c_void_p(base), w, h, refp.stride_c, w, mv.x, mv.y, sign)
                assert np.array_equal(a, b), (w, h, fx, fy, sign)


def _residuals(rng, n):
    yield np.zeros((n, n), np.int16)
    yield np.full((n, n), 255, np.int16)
    yield np.full((n, n), -255, np.int16)
    cb = ((np.add.outer(np.arange(n), np.arange(n)) & 1) * 510 - 255).astype(np.int16)
    yield cb
    for s in (2, 20, 120):
        yield np.clip(rng.normal(0, s, (n, n)).round(), -255, 255).astype(np.int16)
    yield rng.integers(-255, 256, (n, n)).astype(np.int16)


@pytest.mark.parametrize("n", [4, 8, 16, 32, 64])
def test_transform_quant_chain(simd, n):
    rng = np.random.default_rng(n)
    for blk in _residuals(rng, n):
        for fast in ((0, 1) if n > 16 else (0,)):
            ca = np.zeros((n, n), np.int16)
            cb = np.zeros((n, n), np.int16)
            ref.transform(ptr(blk), ptr(ca), n, fast)
            orc.orc_transform(ptr(blk), ptr(cb), n, fast)
            q = min(n, 16)
            assert np.array_equal(ca[:q, :q], cb[:q, :q]), ("fwd", n, fast)
            for qp in (0, 13, 26, 32, 45, 51):
                for cbt in range(4):
                    qa = np.zeros((n, n), np.int16)
                    qb = np.zeros((n, n), np.int16)
                    ra = ref.quantize(ptr(ca), ptr(qa), qp, n, cbt, 0)
                    rb = orc.orc_quantize(ptr(cb), ptr(qb), qp, n, cbt)
                    assert ra == rb and np.array_equal(qa, qb), ("quant", n, qp, cbt)
                da = np.zeros((n, n), np.int16)
                db = np.zeros((n, n), np.int16)
                ref.dequantize(ptr(qa), ptr(da), qp, n)
                orc.orc_dequantize(ptr(qb), ptr(db), qp, n)
                assert np.array_equal(da, db), ("dequant", n, qp)
                ia = np.zeros((n, n), np.int16)
                ib = np.zeros((n, n), np.int16)
                ref.inverse_transform(ptr(da), ptr(ia), n)
                orc.orc_inverse_transform(ptr(db), ptr(ib), n)
                assert np.array_equal(ia, ib), ("inv", n, qp)


@pytest.mark.parametrize("n", [4, 8, 16, 32, 64])
def test_inverse_transform_dense(simd, n):
    """Dense / large coefficients (exercises the int16 clips and defeats the sparse shortcuts)."""
    rng = np.random.default_rng(100 + n)
    q = min(n, 16)
    for amp in (3, 300, 6000, 32767):
        for fill in ("dense", "dc", "4x4"):
            c = np.zeros((n, n), np.int16)
            if fill == "dense":
                c[:q, :q] = rng.integers(-amp, amp + 1, (q, q))
            elif fill == "dc":
                c[0, 0] = amp
            else:
                c[:4, :4] = rng.integers(-amp, amp + 1, (4, 4))
            ia = np.zeros((n, n), np.int16)
            ib = np.zeros((n, n), np.int16)
            ref.inverse_transform(ptr(c), ptr(ia), n)
            orc.orc_inverse_transform(ptr(c), ptr(ib), n)
            assert np.array_equal(ia, ib), (n, amp, fill)


def test_residual_reconstruct():
    rng = np.random.default_rng(9)
    for n in (4, 8, 16, 32, 64):
        org = rng.integers(0, 256, (n, 80), dtype=np.uint8)
        pred = rng.integers(0, 256, (n, n), dtype=np.uint8)
        ra = np.zeros((n, n), np.int16)
        rb = np.zeros((n, n), np.int16)
        ref.get_residual(ptr(ra), ptr(pred), ptr(org), n, 80)
        orc.orc_get_residual(ptr(rb), ptr(pred), ptr(org), n, 80)
        assert np.array_equal(ra, rb)
        blk = rng.integers(-600, 600, (n, n)).astype(np.int16)
        oa = np.zeros((n, 96), np.uint8)
        ob = np.zeros((n, 96), np.uint8)
        ref.reconstruct_block(ptr(blk), ptr(pred), ptr(oa), n, 96)
        orc.orc_reconstruct_block(ptr(blk), ptr(pred), ptr(ob), n, 96)
        assert np.array_equal(oa, ob)


def _aligned(shape, align=16):
    n = int(np.prod(shape))
    raw = np.zeros(n + align, np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + n].reshape(shape)


def test_sad_family(simd):
    rng = np.random.default_rng(21)
    refp = Picture(160, 160, 96, 48, rng=rng)
    y, _, _ = smooth_frame(160 + 192, 160 + 192, seed=4)
    refp.buf_y[...] = y[:refp.rows_y, :refp.stride_y]
    for r in (ref, orc):
        for f in ("sad_calc", "widesad_calc", "sad_calc_fasthalf", "sad_calc_fastquarter", "orc_sad", "orc_widesad",
                  "orc_sad_fasthalf", "orc_sad_fastquarter"):
            if hasattr(r, f) or f in ("sad_calc",):
                try:
                    getattr(r, f).restype = C.c_uint
                except AttributeError:
                    pass
    for (w, h, cb) in ((8, 8, 8), (16, 16, 16), (32, 32, 32), (64, 64, 64), (16, 8, 16), (8, 16, 16), (4, 8, 8), (4, 4, 8), (32, 16, 32)):
        for trial in range(6):
            org = _aligned((cb, cb))
            oy, ox = int(rng.integers(8, 90)), int(rng.integers(8, 90))
            src = refp.plane(0)[oy:oy + cb, ox:ox + cb]
            org[...] = np.clip(src.astype(int) + rng.integers(-9, 10, src.shape), 0, 255)
            dy, dx = int(rng.integers(-3, 4)), int(rng.integers(-3, 4))
            bptr = C.c_void_p(refp.origin_ptr(0) + (oy + dy) * refp.stride_y + ox + dx)
            s = refp.stride_y
            assert ref.sad_calc(ptr(org), bptr, cb, s, w, h) == orc.orc_sad(ptr(org), bptr, cb, s, w, h)
            # ssd_calc's SIMD twin needs both operands 16-byte aligned compact blocks (enc/enc_kernels.c:100-122)
            recb = _aligned((cb, cb))
            recb[...] = refp.plane(0)[oy + dy:oy + dy + cb, ox + dx:ox + dx + cb]
            assert ref.ssd_calc(ptr(org), ptr(recb), cb, cb, w, h) == orc.orc_ssd(ptr(org), ptr(recb), cb, cb, w, h)
            xa, xb = C.c_int(9), C.c_int(9)
            assert ref.widesad_calc(ptr(org), bptr, cb, s, w, h, C.byref(xa)) == orc.orc_widesad(ptr(org), bptr, cb, s, w, h, C.byref(xb))
            assert xa.value == xb.value
            hx, hy, gx, gy = C.c_int(0), C.c_int(0), C.c_int(0), C.c_int(0)
            fh = ref.sad_calc_fasthalf_simd if (simd and w > 4) else ref.sad_calc_fasthalf
            fh.restype = C.c_uint
            va = fh(ptr(org), bptr, cb, s, w, h, C.byref(hx), C.byref(hy), None)
            vb = orc.orc_sad_fasthalf(ptr(org), bptr, cb, s, w, h, C.byref(gx), C.byref(gy))
            assert (va, hx.value, hy.value) == (vb, gx.value, gy.value), ("half", w, h)
            fq = ref.sad_calc_fastquarter_simd if (simd and w > 4) else ref.sad_calc_fastquarter
            fq.restype = C.c_uint
            for (px, py) in ((0, 0), (2, 0), (-2, 0), (0, 2), (0, -2), (2, 2), (-2, 2), (2, -2), (-2, -2)):
                qx, qy, rx, ry = C.c_int(px), C.c_int(py), C.c_int(px), C.c_int(py)
                va = fq(ptr(org), bptr, cb, s, w, h, C.byref(qx), C.byref(qy))
                vb = orc.orc_sad_fastquarter(ptr(org), bptr, cb, s, w, h, C.byref(rx), C.byref(ry))
                assert (va, qx.value, qy.value) == (vb, rx.value, ry.value), ("quarter", w, h, px, py)


def test_clip_mv():
    rng = np.random.default_rng(2)
    for _ in range(4000):
        fw, fh = (416, 240) if rng.random() < 0.5 else (1920, 1080)
        size = int(rng.choice([8, 16, 32, 64]))
        ypos, xpos = int(rng.integers(0, fh // 8)) * 8, int(rng.integers(0, fw // 8)) * 8
        sign = int(rng.integers(0, 2))
        mx, my = int(rng.integers(-900, 900)), int(rng.integers(-900, 900))
        if rng.random() < 0.3:
            mx *= 8
            my *= 8
        m = Mv(mx, my)
        ref.clip_mv(C.byref(m), ypos, xpos, fw, fh, size, sign)
        ax, ay = C.c_int16(mx), C.c_int16(my)
        orc.orc_clip_mv(C.byref(ax), C.byref(ay), ypos, xpos, fw, fh, size, sign)
        assert (m.x, m.y) == (ax.value, ay.value)


def test_block_avg(simd):
    rng = np.random.default_rng(1)
    for n in (4, 8):
        a = _aligned((n, 16))
        b = _aligned((n, 16))
        a[...] = rng.integers(0, 256, a.shape)
        b[...] = rng.integers(0, 256, b.shape)
        pa = _aligned((n, 16))
        pb = _aligned((n, 16))
        ref.block_avg_simd(ptr(pa), ptr(a), ptr(b), 16, 16, 16, n, n)
        orc.orc_block_avg(ptr(pb), ptr(a), ptr(b), 16, 16, 16, n, n)
        assert np.array_equal(pa[:, :n], pb[:, :n])


def test_early_skip(simd):
    """a19: check_early_skip_sub_block / _sub_blockC (enc/encode_block.c:2507-2610), both outcomes exercised."""
    rng = np.random.default_rng(77)
    for f in (ref.check_early_skip_sub_block, ref.check_early_skip_sub_blockC):
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_float]
    for f in (orc.orc_early_skip_luma, orc.orc_early_skip_chroma):
        f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_float]
    seen = set()
    for trial in range(400):
        chroma = trial & 1
        size = int(rng.choice([4, 8])) if chroma else int(rng.choice([8, 16]))
        qp = int(rng.integers(18, 46))
        thr = float(rng.choice([0.3, 0.8, 1.0, 1.3]))
        orig = rng.integers(0, 256, (size, 64), dtype=np.uint8)
        amp = int(rng.choice([1, 2, 4, 8, 16]))
        pb = _aligned((size, size))
        pb[...] = np.clip(orig[:, :size].astype(int) + rng.integers(-amp, amp + 1, (size, size)) + int(rng.integers(-amp, amp + 1)), 0, 255)
        if chroma:
            a = ref.check_early_skip_sub_blockC(None, ptr(orig), 64, size, qp, ptr(pb), thr)
            b = orc.orc_early_skip_chroma(ptr(orig), 64, size, qp, ptr(pb), thr)
        else:
            a = ref.check_early_skip_sub_block(None, ptr(orig), 64, size, qp, ptr(pb), thr)
            b = orc.orc_early_skip_luma(ptr(orig), 64, size, qp, ptr(pb), thr)
        assert a == b, (chroma, size, qp, thr, amp)
        seen.add((chroma, a))
    assert seen == {(0, 0), (0, 1), (1, 0), (1, 1)}, "both outcomes must occur for luma and chroma"


def _intra_case(rng, size, stride=96):
    """a random picture patch with the block's top-left at (x, y) of a `stride`-wide plane"""
    plane = rng.integers(0, 256, (2 * size + 40, stride), dtype=np.uint8)
    return plane


@pytest.mark.parametrize("size", [4, 8, 16, 32, 64])
def test_intra_neighbours(size):
    """make_top_and_left (common/intra_prediction.c:57-143), both branches: whole blocks and the four sub-blocks
    under tb_split, at picture edges and inside, every availability combination."""
    rng = np.random.default_rng(size)
    stride = 200
    plane = rng.integers(0, 256, (200, stride), dtype=np.uint8)
    n = 0
    for ypos in (0, 64):
        for xpos in (0, 64):
            for up in (0, 1):
                for dl in (0, 1):
                    # whole block
                    cases = [(0, 0, size, 0)]
                    if size <= 32:
                        cases += [(i, j, size, 1) for i in (0, size) for j in (0, size)]     # sub-blocks of a 2*size CB
                    for (i, j, s, split) in cases:
                        rec = plane[ypos:, xpos:]
                        L = [np.full(2 * s + 2, 7, np.uint8) for _ in range(2)]
                        T = [np.full(2 * s + 2, 9, np.uint8) for _ in range(2)]
                        tl_ref, tl_orc = C.c_uint8(1), C.c_uint8(2)
                        rb = plane[ypos + i:, xpos + j:]
                        ref.make_top_and_left(ptr(L[0]), ptr(T[0]), C.byref(tl_ref), ptr(rec), stride, ptr(rb), stride, i, j, ypos, xpos, s, up, dl, split)
                        if split:
                            upr = 1 if (j == 0 or (i == 0 and up)) else 0
                            dwl = 1 if (j == 0 and (i == 0 or dl)) else 0
                        else:
                            upr, dwl = up, dl
                        orc.orc_intra_neighbours(ptr(rb), stride, xpos + j, ypos + i, xpos, s, upr, dwl, ptr(L[1]), ptr(T[1]), C.byref(tl_orc))
                        assert np.array_equal(L[0][:2 * s], L[1][:2 * s]) and np.array_equal(T[0][:2 * s], T[1][:2 * s]), (ypos, xpos, i, j, up, dl, split)
                        assert tl_ref.value == tl_orc.value, (ypos, xpos, i, j, up, dl, split)
                        n += 1
    assert n >= 16


@pytest.mark.parametrize("size", [4, 8, 16, 32, 64])
def test_intra_modes(size):
    """get_intra_prediction (common/intra_prediction.c:363-389): all ten modes + an unknown one, smooth and extreme edges."""
    rng = np.random.default_rng(100 + size)
    for trial in range(12):
        if trial < 8:
            left = rng.integers(0, 256, 2 * size, dtype=np.uint8)
            top = rng.integers(0, 256, 2 * size, dtype=np.uint8)
        else:
            left = np.full(2 * size, 255 if trial & 1 else 0, np.uint8)
            top = np.full(2 * size, 0 if trial & 1 else 255, np.uint8)
        tl = int(rng.integers(0, 256))
        for mode in range(11):
            for (yp, xp) in ((0, 0), (0, 8), (8, 0), (8, 8)):
                a = np.zeros((size, size), np.uint8)
                b = np.ones((size, size), np.uint8)
                ref.get_intra_prediction(ptr(left), ptr(top), tl, yp, xp, size, ptr(a), mode)
                orc.orc_intra_predict(ptr(left), ptr(top), tl, yp, xp, size, ptr(b), mode)
                assert np.array_equal(a, b), (size, mode, yp, xp, trial)


def test_intra_availability():
    for size in (8, 16, 32, 64):
        for ypos in range(0, 192, size):
            for xpos in range(0, 192, size):
                assert ref.get_upright_available(ypos, xpos, size, 192) == orc.orc_upright_available(ypos, xpos, size, 192)
                assert ref.get_downleft_available(ypos, xpos, size, 192) == orc.orc_downleft_available(ypos, xpos, size, 192)


# ------------------------------------------------------------------ temporal interpolation (8f #2) ----
def oracle_interpolate(p0, p1, ratio, pos, want_field=False):
    """oracle/orc_interp.c on two thorref.Picture references -> Picture (visible part) [, mv0, mv1]"""
    w, h = p0.width, p0.height
    out = Picture(w, h, 96, 48)
    bw, bh = C.c_int(), C.c_int()
    orc.orc_interp_field_size(w, h, C.byref(bw), C.byref(bh))
    mv0 = np.zeros((bh.value, bw.value, 2), np.int16)
    mv1 = np.zeros((bh.value, bw.value, 2), np.int16)
    orc.orc_interpolate_frames(C.c_void_p(out.origin_ptr(0)), C.c_void_p(out.origin_ptr(1)), C.c_void_p(out.origin_ptr(2)), out.stride_y, out.stride_c,
                               C.c_void_p(p0.origin_ptr(0)), C.c_void_p(p0.origin_ptr(1)), C.c_void_p(p0.origin_ptr(2)),
                               C.c_void_p(p1.origin_ptr(0)), C.c_void_p(p1.origin_ptr(1)), C.c_void_p(p1.origin_ptr(2)),
                               w, h, p0.stride_y, p0.stride_c, p0.pad_y, ratio, pos, ptr(mv0), ptr(mv1))
    return (out, mv0, mv1) if want_field else out


@pytest.mark.parametrize("case", [(416, 240, "scene", (3, 5)), (192, 136, "scene", (-7, 12)), (128, 128, "noise", None),
                                  (320, 192, "scene", (40, -70)), (64, 48, "scene", (1, 2)), (200, 72, "noise", None),
                                  (416, 240, "scene", (6, 200))])
@pytest.mark.parametrize("rp", [(2, 1), (4, 1), (4, 3), (8, 3), (3, 1)])
def test_interpolate_frames(simd, case, rp):
    w, h, kind, motion = case
    ratio, pos = rp
    p0, p1 = interp_pair(w, h, seed=w + ratio, motion=motion or (0, 0), kind=kind)
    s0, s1 = p0.as_ref_struct(), p1.as_ref_struct()
    want = Picture(w, h, 96, 48)
    sw = want.as_ref_struct()
    ref.interpolate_frames(C.byref(sw), C.byref(s0), C.byref(s1), ratio, pos)
    got = oracle_interpolate(p0, p1, ratio, pos)
    for i in range(3):
        assert np.array_equal(want.plane(i), got.plane(i)), f"plane {i}"
    assert not np.array_equal(got.plane(0), p0.plane(0)) and not np.array_equal(got.plane(0), p1.plane(0))


def test_interpolate_coverage():
    """the cases above reach every branch: clamped SADs, skipped blocks, all three blend forms, refinement rounds"""
    tot = np.zeros(6, np.int64)
    for (w, h, kind, motion) in [(416, 240, "scene", (3, 5)), (416, 240, "scene", (6, 200)), (128, 128, "noise", None)]:
        for (ratio, pos) in [(2, 1), (8, 3)]:
            p0, p1 = interp_pair(w, h, seed=w + ratio, motion=motion or (0, 0), kind=kind)
            orc.orc_interp_counters(None, 1)
            oracle_interpolate(p0, p1, ratio, pos)
            c = np.zeros(6, np.int64)
            orc.orc_interp_counters(ptr(c), 0)
            tot += c
    assert (tot > 0).all(), tot


# ---------------------------------------------------------------- bit-cost model ----
def test_vlc_tables():
    """orc_vlc_len == quote_vlc for every table and the code numbers each can see"""
    ref.quote_vlc.restype = C.c_int
    for n in range(14):
        hi = 3 if n == 8 else 5000
        for cn in list(range(min(hi, 700))) + ([] if n == 8 else [1023, 1024, 4095, 4096, 65535, 65536, 131071]):
            assert orc.orc_vlc_len(n, cn) == ref.quote_vlc(n, cn), (n, cn)


def test_find_code():
    for chroma in (0, 1):
        for maxrun in (0, 1, 3, 4, 5, 14, 62, 254):
            for run in range(0, maxrun + 1):
                for level in (1, 2, 9):
                    assert orc.orc_find_code(run, level, maxrun, chroma, 0) == ref.find_code(run, level, maxrun, chroma, 0)
        assert orc.orc_find_code(0, 0, 0, chroma, 1) == ref.find_code(0, 0, 0, chroma, 1)


@pytest.mark.parametrize("size", [4, 8, 16, 32, 64])
def test_coeff_bits(size):
    """orc_coeff_bits == bits write_coeff puts into a real stream, all four block types, from single +-1 to dense and huge levels"""
    rng = np.random.default_rng(size)
    for kind in COEFF_KINDS:
        for rep in range(12 if size <= 16 else 5):
            c = coeff_block(rng, size, kind)
            for typ in range(4):
                s = ref_stream()
                ref.write_coeff(C.byref(s), ptr(c), size, typ)
                want = ref_bitpos(s)
                got = orc.orc_coeff_bits(ptr(c), size, size, typ)
                assert got == want, (size, kind, rep, typ, got, want)
    assert orc.orc_coeff_bits(ptr(np.zeros((size, size), np.int16)), size, size, 0) == 0


def test_mv_bits():
    class Mv(C.Structure):
        _fields_ = [("x", C.c_int16), ("y", C.c_int16)]        # common/types.h:105-109
    rng = np.random.default_rng(3)
    for _ in range(300):
        mv, mvp = rng.integers(-600, 601, 2), rng.integers(-600, 601, 2)
        s = ref_stream()
        a, b = Mv(int(mv[0]), int(mv[1])), Mv(int(mvp[0]), int(mvp[1]))
        ref.write_mv(C.byref(s), C.byref(a), C.byref(b))
        assert orc.orc_mv_bits(int(mv[0]), int(mv[1]), int(mvp[0]), int(mvp[1])) == ref_bitpos(s)


def test_block_bits():
    """orc_block_bits / orc_block_syntax_bits against the reference's write_block (through oracle/bits_shim.c): 4000
    random blocks over every frame type, mode, partition, context, transform split and cbp pattern."""
    from thorref import BS, load_ref_bits, random_block_syntax
    shim = load_ref_bits()
    rng = np.random.default_rng(11)
    seen = set()
    for _ in range(4000):
        d, cy, cu, cv = random_block_syntax(rng)
        want = shim.ref_write_block_bits(ptr(d), ptr(cy), ptr(cu), ptr(cv))
        got = orc.orc_block_bits(ptr(d), ptr(cy), ptr(cu), ptr(cv))
        assert got == want, d
        seen.add((int(d[BS["frame_type"]]), int(d[BS["mode"]]), int(d[BS["tb_split"]]), int(d[BS["size"]])))
        # the syntax part alone: the same block without levels where that is a legal block (no cbp bit set)
        z = d.copy()
        z[BS["cbp_y"]:BS["cbp_y"] + 3] = 0
        assert orc.orc_block_syntax_bits(ptr(z)) == shim.ref_write_block_bits(ptr(z), None, None, None)
    assert len(seen) > 60


@pytest.mark.parametrize("qp", [15, 17, 18, 20, 32, 51])
def test_deblock_threshold_kat(simd, qp):
    """`d < beta` exactly at the boundary (SURVEY 8c KAT list): edges designed with d = beta-2..beta+2 are filtered by the
    reference iff d < beta, and the oracle agrees pixel for pixel."""
    from thorref import BETA_TABLE, deblock_threshold_case
    w, h = 192, 128
    y, u, v, bi, dv, dh = deblock_threshold_case(w, h, qp)
    a = Picture(w, h)
    a.set_visible(y, u, v)
    b = a.copy()
    fa = a.as_ref_struct()
    ref.deblock_frame_y(C.byref(fa), ptr(bi), w, h, C.c_uint8(qp))
    orc.orc_deblock_frame_y(C.c_void_p(b.origin_ptr(0)), b.stride_y, ptr(bi), w, h, qp)
    out = a.plane(0)
    assert np.array_equal(out, b.plane(0))
    beta = BETA_TABLE[qp]
    # a filtered vertical edge moves q0 of row 0 of the tile to its right (a pixel no horizontal filter... row 0 is
    # touched by the horizontal edge above, so look at row 3, which no horizontal filter writes); likewise column 3
    moved_v = out[3::8, 8::8] != y[3::8, 8::8]
    moved_h = out[8::8, 3::8] != y[8::8, 3::8]
    if qp < 18:                      # beta 0 (qp < 16), or tc 0 (qp < 18: the filter fires and adds a delta clipped to 0)
        assert np.array_equal(out, y)
        return
    assert np.array_equal(moved_v, (dv[:, :-1] < beta)), "vertical edges"
    assert np.array_equal(moved_h, (dh[:-1, :] < beta)), "horizontal edges"
    assert moved_v.any() and not moved_v.all() and moved_h.any() and not moved_h.all()
