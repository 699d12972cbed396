"""GPU parity: fused deblock/CLPF/reference kernel vs the CPU oracle (bit-exact), through the C ABI."""
import ctypes as C
import hashlib

import numpy as np
import pytest

from thorref import Picture as HostPic, blocky_frame, load_oracle, ptr, random_blockinfo

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tb():
    import thor_b200
    thor_b200.init(0)
    yield thor_b200
    thor_b200.shutdown()


def oracle_chain(orc, host, org, bi, w, h, qp, stages, bits_in):
    """Run the selected stages on a host picture with the oracle; returns per-SB bits."""
    nsb = (w // 64) * (h // 64)
    bits = np.full(max(nsb, 1), 0xFF, np.uint8)
    if stages & 1:
        orc.orc_deblock_frame_y(C.c_void_p(host.origin_ptr(0)), host.stride_y, ptr(bi), w, h, qp)
    if stages & 2:
        orc.orc_deblock_frame_uv(C.c_void_p(host.origin_ptr(1)), C.c_void_p(host.origin_ptr(2)), host.stride_c, ptr(bi), w, h,
                                 orc.orc_chroma_qp(qp))
    if stages & 4:
        orc.orc_clpf_frame(C.c_void_p(host.origin_ptr(0)), C.c_void_p(host.origin_ptr(1)), C.c_void_p(host.origin_ptr(2)),
                           host.stride_y, host.stride_c, C.c_void_p(org.origin_ptr(0)) if org is not None else None,
                           org.stride_y if org is not None else 0, ptr(bi), w, h, ptr(bits_in) if bits_in is not None else None,
                           ptr(bits))
    return bits[:nsb]


def make_case(w, h, seed):
    rng = np.random.default_rng(seed)
    y, u, v = blocky_frame(w, h, seed)
    org = HostPic(w, h)
    org.set_visible(y, u, v)
    noisy = [np.clip(p.astype(int) + rng.integers(-3, 4, p.shape), 0, 255).astype(np.uint8) for p in (y, u, v)]
    rec = HostPic(w, h)
    rec.set_visible(*noisy)
    bi = random_blockinfo(w, h, rng)
    return rng, org, rec, bi


@pytest.mark.parametrize("wh", [(416, 240), (200, 136), (64, 64), (72, 8), (1920, 1080)])
@pytest.mark.parametrize("qp", [0, 22, 32, 51])
def test_deblock_stages(tb, wh, qp):
    w, h = wh
    if (w, h) == (1920, 1080) and qp not in (32,):
        pytest.skip("full size once")
    orc = load_oracle()
    rng, org, rec, bi = make_case(w, h, seed=qp + w)
    for stages in (1, 2, 3):
        host = rec.copy()
        oracle_chain(orc, host, None, bi, w, h, qp, stages, None)
        dev = tb.Picture(w, h)
        dev.upload(rec.plane(0), rec.plane(1), rec.plane(2))
        tb.ops.blockinfo_upload(dev, bi)
        tb.ops.loopfilter_frame(dev, qp, stages)
        got = dev.download()
        dev.close()
        for i in range(3):
            assert np.array_equal(got[i], host.plane(i)), (stages, "plane", i)
    if qp >= 22:
        assert not np.array_equal(host.plane(0), rec.plane(0))


@pytest.mark.parametrize("wh", [(416, 240), (256, 192), (200, 136), (1920, 1080)])
@pytest.mark.parametrize("enc", [True, False])
def test_clpf(tb, wh, enc):
    w, h = wh
    orc = load_oracle()
    rng, org, rec, bi = make_case(w, h, seed=w + 5 * enc)
    nsb = (w // 64) * (h // 64)
    bits_in = rng.integers(0, 2, max(nsb, 1)).astype(np.uint8)
    host = rec.copy()
    want_bits = oracle_chain(orc, host, org if enc else None, bi, w, h, 0, 4, None if enc else bits_in)
    dev = tb.Picture(w, h)
    dev.upload(rec.plane(0), rec.plane(1), rec.plane(2))
    tb.ops.blockinfo_upload(dev, bi)
    dorg = None
    if enc:
        dorg = tb.Picture(w, h)
        dorg.upload(org.plane(0), org.plane(1), org.plane(2))
    got_bits = tb.ops.clpf_frame(dev, org=dorg, sb_bits_in=None if enc else bits_in)
    got = dev.download()
    assert np.array_equal(got_bits, want_bits)
    for i in range(3):
        assert np.array_equal(got[i], host.plane(i)), ("plane", i)
    if nsb:
        assert (want_bits == 1).any() and not np.array_equal(host.plane(0), rec.plane(0))
    dev.close()
    if dorg:
        dorg.close()


@pytest.mark.parametrize("wh", [(416, 240), (200, 136), (1920, 1080), (3840, 2160)])     # 2160 = 33*64 + 48: the last SB row is partial
def test_fused_frame_tail_and_reference(tb, wh):
    """deblock Y+UV -> CLPF (encoder decision) -> create_reference_frame in one call."""
    w, h = wh
    orc = load_oracle()
    rng, org, rec, bi = make_case(w, h, seed=99 + w)
    qp = 34
    host = rec.copy()
    want_bits = oracle_chain(orc, host, org, bi, w, h, qp, 7, None)
    href = HostPic(w, h, 96, 48, rng=rng)
    for i in range(3):
        href.plane(i)[...] = host.plane(i)
    orc.orc_pad_plane(C.c_void_p(href.origin_ptr(0)), href.stride_y, w, h, 96, 96)
    orc.orc_pad_plane(C.c_void_p(href.origin_ptr(1)), href.stride_c, w // 2, h // 2, 48, 48)
    orc.orc_pad_plane(C.c_void_p(href.origin_ptr(2)), href.stride_c, w // 2, h // 2, 48, 48)

    dev, dorg, dref = tb.Picture(w, h), tb.Picture(w, h), tb.Picture(w, h, 96, 48)
    dev.upload(rec.plane(0), rec.plane(1), rec.plane(2))
    tb.ops.blockinfo_upload(dev, bi)
    dorg.upload(org.plane(0), org.plane(1), org.plane(2))
    got_bits = tb.ops.loopfilter_frame(dev, qp, 15, org=dorg, ref_out=dref)
    assert np.array_equal(got_bits, want_bits)
    got = dev.download()
    for i in range(3):
        assert np.array_equal(got[i], host.plane(i)), ("rec plane", i)
    py, pu, pv = dref.download_padded()
    # compare whole padded planes except the alignment slack right of the padded row
    wy, wc = w + 192, w // 2 + 96
    assert np.array_equal(py[:, :wy], href.buf_y[:, :wy])
    assert np.array_equal(pu[:, :wc], href.buf_u[:, :wc])
    assert np.array_equal(pv[:, :wc], href.buf_v[:, :wc])
    # idempotence of padding (size-independent property)
    tb.ops.pad_yuv_frame(dref)
    py2, _, _ = dref.download_padded()
    assert hashlib.md5(py2.tobytes()).hexdigest() == hashlib.md5(py.tobytes()).hexdigest()
    for p in (dev, dorg, dref):
        p.close()


def test_no_silent_fallback(tb):
    """The operators must refuse to run without uploaded side information (no fallback path)."""
    dev = tb.Picture(128, 64)
    with pytest.raises(tb.ThorCudaError):
        tb.ops.loopfilter_frame(dev, 30, 7)
    dev.close()


def test_batched_frames_one_launch(tb):
    """Four frames of different sizes / qps through ONE thorcuda_loopfilter_frames launch."""
    orc = load_oracle()
    cases = [(416, 240, 30), (200, 136, 40), (128, 64, 22), (416, 240, 51)]
    jobs, want, pics = [], [], []
    for k, (w, h, qp) in enumerate(cases):
        rng, org, rec, bi = make_case(w, h, seed=700 + k)
        host = rec.copy()
        bits = oracle_chain(orc, host, org, bi, w, h, qp, 7, None)
        href = HostPic(w, h, 96, 48)
        for i in range(3):
            href.plane(i)[...] = host.plane(i)
        orc.orc_pad_plane(C.c_void_p(href.origin_ptr(0)), href.stride_y, w, h, 96, 96)
        orc.orc_pad_plane(C.c_void_p(href.origin_ptr(1)), href.stride_c, w // 2, h // 2, 48, 48)
        orc.orc_pad_plane(C.c_void_p(href.origin_ptr(2)), href.stride_c, w // 2, h // 2, 48, 48)
        dev, dorg, dref = tb.Picture(w, h), tb.Picture(w, h), tb.Picture(w, h, 96, 48)
        dev.upload(rec.plane(0), rec.plane(1), rec.plane(2))
        dorg.upload(org.plane(0), org.plane(1), org.plane(2))
        tb.ops.blockinfo_upload(dev, bi)
        jobs.append(dict(rec=dev, qp=qp, stages=15, org=dorg, ref_out=dref))
        want.append((host, bits, href))
        pics += [dev, dorg, dref]
    before = tb.launch_count()
    got_bits = tb.ops.loopfilter_frames(jobs)
    assert tb.launch_count() - before == 2          # fused filter kernel + border kernel, for all four frames
    for j, (host, bits, href), gb in zip(jobs, want, got_bits):
        assert np.array_equal(gb, bits)
        got = j["rec"].download()
        for i in range(3):
            assert np.array_equal(got[i], host.plane(i))
        py, pu, pv = j["ref_out"].download_padded()
        w = host.width
        assert np.array_equal(py[:, :w + 192], href.buf_y[:, :w + 192])
        assert np.array_equal(pu[:, :w // 2 + 96], href.buf_u[:, :w // 2 + 96])
        assert np.array_equal(pv[:, :w // 2 + 96], href.buf_v[:, :w // 2 + 96])
    for p in pics:
        p.close()


def test_fences(tb):
    """thorcuda_fence_record / _wait: a marker behind the work queued so far (library stream + bulk fetches); waiting for
    it makes that work's host outputs valid without a full thorcuda_sync."""
    lib = tb.cdll()
    w, h = 416, 240
    rng = np.random.default_rng(9)
    planes = [rng.integers(0, 256, (h, w), dtype=np.uint8), rng.integers(0, 256, (h // 2, w // 2), dtype=np.uint8),
              rng.integers(0, 256, (h // 2, w // 2), dtype=np.uint8)]
    a, b = tb.Picture(w, h), tb.Picture(w, h)
    a.upload(*planes)
    tb.check(lib.thorcuda_pic_copy(b.handle, a.handle))
    f = lib.thorcuda_fence_record()
    assert 0 <= f < 16
    ids = [lib.thorcuda_fence_record() for _ in range(15)]
    assert all(0 <= i < 16 for i in ids) and len(set(ids + [f])) == 16
    assert lib.thorcuda_fence_record() == -1, "a 17th open fence must be refused"
    tb.check(lib.thorcuda_fence_wait(f))
    for i in ids:
        tb.check(lib.thorcuda_fence_wait(i))
    assert lib.thorcuda_fence_wait(f) != 0, "a fence can be waited for once"
    assert lib.thorcuda_fence_wait(99) != 0
    got = b.download()
    for i in range(3):
        assert np.array_equal(got[i], planes[i])
    a.close(); b.close()


def test_calibrate_absdiff(tb):
    """The register-only VABSDIFF4 loop behind bench.py's roofline.alu: a plausible rate for a B200 (148 SMs x 64 ALU
    lanes x 4 abs-diffs at 1.5-2.1 GHz = 57-80 T/s), and stable between two runs."""
    lib = tb.cdll()
    a = lib.thorcuda_calibrate_absdiff(0)
    b = lib.thorcuda_calibrate_absdiff(8192)
    assert 20e12 < a < 120e12 and 20e12 < b < 120e12
    assert abs(a - b) / max(a, b) < 0.25


@pytest.mark.parametrize("qp", [15, 17, 18, 20, 32, 51])
def test_deblock_threshold_kat(tb, qp):
    """`d < beta` exactly at the boundary: the known-answer picture of thorref.deblock_threshold_case (edges designed with
    d = beta-2..beta+2; pinned against the reference in tests/test_oracle_vs_ref.py), alone and inside the fused tail"""
    from thorref import BETA_TABLE, deblock_threshold_case
    orc = load_oracle()
    w, h = 192, 128
    y, u, v, bi, dv, dh = deblock_threshold_case(w, h, qp)
    rec = HostPic(w, h)
    rec.set_visible(y, u, v)
    for stages in (1, 7):
        host = rec.copy()
        oracle_chain(orc, host, None, bi, w, h, qp, stages, np.ones(6, np.uint8))
        dev = tb.Picture(w, h)
        dev.upload(y, u, v)
        tb.ops.blockinfo_upload(dev, bi)
        if stages == 1:
            tb.ops.loopfilter_frame(dev, qp, 1)
        else:
            tb.ops.loopfilter_frame(dev, qp, 3)
            tb.ops.clpf_frame(dev, org=None, sb_bits_in=np.ones(6, np.uint8))
        got = dev.download()
        dev.close()
        for i in range(3):
            assert np.array_equal(got[i], host.plane(i)), (stages, "plane", i)
        if stages == 1 and qp >= 18:
            moved = got[0][3::8, 8::8] != y[3::8, 8::8]
            assert np.array_equal(moved, dv[:, :-1] < BETA_TABLE[qp])
