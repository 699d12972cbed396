"""GPU parity (bit-exact, through the C ABI) of the temporal interpolation (interpolate_frames,
common/temporal_interp.c:972-1053) against the CPU oracle: the interpolated picture and the full-resolution motion field,
on scenes with global motion, a static band (skip path), an object moving against the background, unrelated pictures
(wandering search, vectors leaving the picture) and motion larger than the pyramid's border (clamped SADs)."""
import ctypes as C

import numpy as np
import pytest

from thorref import Picture, interp_pair, load_oracle, ptr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tb():
    import thor_b200
    thor_b200.init(0)
    yield thor_b200
    thor_b200.shutdown()


@pytest.fixture(scope="module")
def orc():
    return load_oracle()


def oracle_interpolate(orc, p0, p1, ratio, pos):
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
    return out, mv1


def device_pic(tb, p):
    d = tb.Picture(p.width, p.height, p.pad_y, p.pad_c)
    d.upload(p.plane(0), p.plane(1), p.plane(2))
    tb.ops.pad_yuv_frame(d)
    return d


CASES = [(416, 240, "scene", (3, 5)), (192, 136, "scene", (-7, 12)), (128, 128, "noise", None), (320, 192, "scene", (40, -70)),
         (64, 48, "scene", (1, 2)), (200, 72, "noise", None), (416, 240, "scene", (6, 200)), (1920, 1080, "scene", (4, -9)),
         (3840, 2160, "scene", (-6, 11)), (200, 136, "scene", (3, 5)), (200, 136, "scene", (-9, 14)), (200, 136, "noise", None), (136, 200, "scene", (5, -3)),
         (72, 72, "scene", (2, 1)),
         # widths whose coarser levels are not multiples of 4 (200 -> 50, 264 -> 66): the pyramid's right border ends in a partial word
         (264, 200, "noise", None), (200, 136, "scene", (60, -40)), (264, 136, "scene", (-50, 9))]


@pytest.mark.parametrize("case", CASES, ids=[f"{c[0]}x{c[1]}-{c[2]}-{i}" for i, c in enumerate(CASES)])
def test_interpolate_frames(tb, orc, case):
    w, h, kind, motion = case
    p0, p1 = interp_pair(w, h, seed=w + 11, motion=motion or (0, 0), kind=kind)
    d0, d1 = device_pic(tb, p0), device_pic(tb, p1)
    out = tb.Picture(w, h, 96, 48)
    for (ratio, pos) in ([(2, 1), (4, 1), (4, 3), (8, 3), (3, 1)] if w < 1000 else [(2, 1), (8, 5)]):
        want, want_mv = oracle_interpolate(orc, p0, p1, ratio, pos)
        tb.ops.interpolate_frames(out, d0, d1, ratio, pos)
        tb.sync()
        got_mv = tb.ops.interp_field()
        bad = np.argwhere((got_mv != want_mv).any(axis=2))
        assert bad.size == 0, f"ratio {ratio} pos {pos}: {len(bad)} field entries differ, first at (by, bx) = {bad[0]}: " \
                              f"{got_mv[tuple(bad[0])]} vs {want_mv[tuple(bad[0])]}"
        got = out.download()
        for i in range(3):
            assert np.array_equal(got[i], want.plane(i)), f"ratio {ratio} pos {pos} plane {i}"
    for d in (d0, d1, out):
        d.close()


def test_interpolate_rejects(tb):
    a, b, c = tb.Picture(64, 64, 32, 16), tb.Picture(64, 64, 32, 16), tb.Picture(64, 64, 32, 16)
    small = tb.Picture(64, 32, 32, 16)
    nopad = tb.Picture(64, 64)
    for args in [(a, a, b, 2, 1), (a, b, small, 2, 1), (a, b, c, 2, 2), (a, b, c, 1, 1), (a, nopad, nopad, 2, 1)]:
        with pytest.raises(tb.ThorCudaError):
            tb.ops.interpolate_frames(*args)
    for p in (a, b, c, small, nopad):
        p.close()


def test_interpolate_is_repeatable(tb):
    """the wavefront has no run-to-run freedom: ten runs, one result"""
    p0, p1 = interp_pair(416, 240, seed=5, motion=(9, -4))
    d0, d1 = device_pic(tb, p0), device_pic(tb, p1)
    out = tb.Picture(416, 240, 96, 48)
    first = None
    for _ in range(10):
        tb.ops.interpolate_frames(out, d0, d1, 4, 1)
        tb.sync()
        g = [x.copy() for x in out.download()] + [tb.ops.interp_field()]
        if first is None:
            first = g
        else:
            assert all(np.array_equal(a, b) for a, b in zip(first, g))
    for d in (d0, d1, out):
        d.close()
