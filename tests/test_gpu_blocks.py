"""GPU parity (bit-exact, through the C ABI) of the block-level batched kernels against the CPU oracle:
inter prediction, transform/quantise/reconstruct, SAD tables and the batched motion-search costs."""
import ctypes as C

import numpy as np
import pytest

from thorref import Picture as HostPic, load_oracle, ptr, smooth_frame

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tb():
    import thor_b200
    thor_b200.init(0)
    yield thor_b200
    thor_b200.shutdown()


@pytest.fixture(scope="module")
def orc():
    o = load_oracle()
    for f in ("orc_sad", "orc_widesad", "orc_sad_fasthalf", "orc_sad_fastquarter"):
        getattr(o, f).restype = C.c_uint
    return o


# ----------------------------------------------------------------------------- helpers ----
def _arr(tasks, dt):
    out = np.zeros(len(tasks), dt)
    for i, t in enumerate(tasks):
        out[i] = t[()]
    return out


def partition(width, height, rng, min_size=8, p_split=0.6):
    """Non-overlapping quad-tree partition of the frame into square blocks (x, y, size)."""
    out = []

    def rec(y, x, size):
        if y >= height or x >= width:
            return
        fits = y + size <= height and x + size <= width
        if size > min_size and (not fits or rng.random() < p_split):
            h = size // 2
            for dy, dx in ((0, 0), (0, h), (h, 0), (h, h)):
                rec(y + dy, x + dx, h)
        elif fits:
            out.append((x, y, size))

    for y in range(0, height, 64):
        for x in range(0, width, 64):
            rec(y, x, 64)
    return out


def dev_ref_from_host(tb, hp):
    d = tb.Picture(hp.width, hp.height, hp.pad_y, hp.pad_c)
    d.upload_padded(hp.buf_y, hp.buf_u, hp.buf_v)
    return d


def textured_ref(w, h, seed):
    """Padded reference picture with texture everywhere (borders included) plus extreme samples."""
    rng = np.random.default_rng(seed)
    hp = HostPic(w, h, 96, 48)
    y, u, v = smooth_frame(hp.stride_y, hp.rows_y + 0, seed)
    hp.buf_y[...] = y[:hp.rows_y, :hp.stride_y]
    hp.buf_u[...] = rng.integers(0, 256, hp.buf_u.shape)
    hp.buf_v[...] = rng.integers(0, 256, hp.buf_v.shape)
    hp.buf_y[::7, ::5] = 255
    hp.buf_y[3::11, 1::3] = 0
    return hp


# ------------------------------------------------------------------- inter prediction ----
def oracle_predict(orc, refs, t, bipred):
    """One task through the oracle; returns (luma, u, v) blocks."""
    res = []
    for pl in range(3):
        w, h = (int(t["w"]), int(t["h"])) if pl == 0 else (int(t["w"]) // 2, int(t["h"]) // 2)
        x, y = (int(t["x"]), int(t["y"])) if pl == 0 else (int(t["x"]) // 2, int(t["y"]) // 2)
        blocks = []
        for k in range(2):
            rix = int(t["ref%d" % k])
            if rix < 0:
                break
            r = refs[rix]
            stride = r.stride_y if pl == 0 else r.stride_c
            base = C.c_void_p(r.origin_ptr(pl) + y * stride + x)
            b = np.zeros((h, w), np.uint8)
            mvx, mvy, sg = int(t["mv%dx" % k]), int(t["mv%dy" % k]), int(t["sign%d" % k])
            if pl == 0:
                orc.orc_inter_luma(ptr(b), base, w, h, stride, w, mvx, mvy, sg, bipred)
            else:
                orc.orc_inter_chroma(ptr(b), base, w, h, stride, w, mvx, mvy, sg)
            blocks.append(b)
        if len(blocks) == 2:
            out = np.zeros_like(blocks[0])
            orc.orc_average_trunc(ptr(out), ptr(blocks[0]), ptr(blocks[1]), out.size)
            blocks = [out]
        res.append(blocks[0])
    return res


def make_pred_tasks(tb, w, h, rng, nrefs, dst_index, mv_sigma=24, bi_frac=0.3, rect_frac=0.3, min_size=8):
    tasks = []
    for (x, y, s) in partition(w, h, rng, min_size=min_size):
        parts = [(x, y, s, s)]
        r = rng.random()
        if r < rect_frac / 3:
            parts = [(x, y, s, s // 2), (x, y + s // 2, s, s // 2)]
        elif r < 2 * rect_frac / 3:
            parts = [(x, y, s // 2, s), (x + s // 2, y, s // 2, s)]
        elif r < rect_frac:
            parts = [(x + dx, y + dy, s // 2, s // 2) for dy in (0, s // 2) for dx in (0, s // 2)]
        for (px, py, pw, ph) in parts:
            t = np.zeros((), tb.PRED_TASK)
            t["x"], t["y"], t["w"], t["h"], t["dst"] = px, py, pw, ph, dst_index
            bi = rng.random() < bi_frac and nrefs > 1
            t["ref0"] = rng.integers(0, nrefs)
            t["ref1"] = rng.integers(0, nrefs) if bi else -1
            for k in range(2):
                # keep the block + filter reach inside the 96-pixel border, like clip_mv does (80 px)
                mx = int(np.clip(rng.normal(0, mv_sigma), -4 * (px + 80), 4 * (w + 80 - px - pw)))
                my = int(np.clip(rng.normal(0, mv_sigma), -4 * (py + 80), 4 * (h + 80 - py - ph)))
                sg = int(rng.integers(0, 2))
                t["mv%dx" % k], t["mv%dy" % k], t["sign%d" % k] = (-mx if sg else mx), (-my if sg else my), sg
            tasks.append(t)
    return _arr(tasks, tb.PRED_TASK)


@pytest.mark.parametrize("wh,bipred", [((416, 240), 1), ((416, 240), 0), ((200, 136), 1), ((1920, 1080), 1), ((3840, 2160), 1)])
def test_predict_blocks(tb, orc, wh, bipred):
    w, h = wh
    rng = np.random.default_rng(w + bipred)
    href = [textured_ref(w, h, 31 + k) for k in range(2)]
    dref = [dev_ref_from_host(tb, r) for r in href]
    dst = tb.Picture(w, h)
    big = w >= 1920
    tasks = make_pred_tasks(tb, w, h, rng, 2, 2, min_size=16 if big else 8, rect_frac=0.15 if big else 0.3)
    # force every luma phase (incl. the 2-D centre filter) and every chroma phase to appear
    for i in range(min(64, tasks.size)):
        tasks[i]["mv0x"] = (tasks[i]["mv0x"] & ~7) | (i & 7)
        tasks[i]["mv0y"] = (tasks[i]["mv0y"] & ~7) | (i >> 3)
    tb.ops.predict_blocks(dref + [dst], tasks, bipred)
    got = dst.download()
    step = 7 if big else 1                      # full size: every 7th task through the (slow) ctypes oracle
    for t in tasks[::step]:
        want = oracle_predict(orc, href, t, bipred)
        for pl in range(3):
            x, y = (int(t["x"]), int(t["y"])) if pl == 0 else (int(t["x"]) // 2, int(t["y"]) // 2)
            hh, ww = want[pl].shape
            assert np.array_equal(got[pl][y:y + hh, x:x + ww], want[pl]), (pl, t)
    for p in dref + [dst]:
        p.close()


@pytest.mark.parametrize("wh", [(416, 240), (1920, 1080)])
def test_predict_mv_at_clip(tb, orc, wh):
    """SURVEY 8c KAT: vectors exactly at the +-80 pixel clip of clip_mv (enc/encode_block.c:816-828) -- the block lies
    wholly in the picture's border, 80 pixels outside the frame, at all four corners and the edge midpoints, for every
    block size and every fractional phase that still satisfies the clip (the integer limit itself: phase 0 at the far
    side), uni- and bi-predicted."""
    w, h = wh
    href = [textured_ref(w, h, 77 + k) for k in range(2)]
    dref = [dev_ref_from_host(tb, r) for r in href]
    dst = tb.Picture(w, h)
    tasks = []
    for s in (8, 16, 32, 64):
        xs, ys = (0, (w // 2) & ~63, (w - s) & ~(s - 1)), (0, (h // 2) & ~63, (h - s) & ~(s - 1))
        k = 0
        for py in ys:
            for px in xs:
                lo_x, hi_x = -4 * (px + 80), 4 * (w + 80 - px - s)
                lo_y, hi_y = -4 * (py + 80), 4 * (h + 80 - py - s)
                for (mx, my) in ((lo_x, lo_y), (hi_x, hi_y), (lo_x, hi_y), (hi_x, lo_y), (lo_x + 1 + k % 3, lo_y + 1 + k % 2),
                                 (hi_x - 1 - k % 3, hi_y - 2), (lo_x + 2, hi_y - 3), (hi_x, 0), (0, lo_y)):
                    t = np.zeros((), tb.PRED_TASK)
                    t["x"], t["y"], t["w"], t["h"], t["dst"] = px, py, s, s, 2
                    t["ref0"], t["ref1"] = k % 2, (1 - k % 2) if k % 3 == 0 else -1
                    t["mv0x"], t["mv0y"] = mx, my
                    t["mv1x"], t["mv1y"], t["sign1"] = -hi_x if k % 2 else -lo_x, -lo_y, 1      # negated before use
                    tasks.append(t)
                    k += 1
    tasks = _arr(tasks, tb.PRED_TASK)
    # the destination blocks overlap between tasks (same corner, nine vectors, four sizes): run groups of disjoint blocks
    for i in range(36):
        group = tasks[(tasks["w"] == (8, 16, 32, 64)[i // 9])][i % 9::9]
        tb.ops.predict_blocks(dref + [dst], group, 1)
        got = dst.download()
        for t in group:
            want = oracle_predict(orc, href, t, 1)
            for pl in range(3):
                x, y = (int(t["x"]), int(t["y"])) if pl == 0 else (int(t["x"]) // 2, int(t["y"]) // 2)
                hh, ww = want[pl].shape
                assert np.array_equal(got[pl][y:y + hh, x:x + ww], want[pl]), (pl, t)
    for p in dref + [dst]:
        p.close()


def test_predict_rejects_out_of_range_mv(tb):
    ref = tb.Picture(128, 64, 96, 48)
    dst = tb.Picture(128, 64)
    t = np.zeros(1, tb.PRED_TASK)
    t["w"], t["h"], t["dst"], t["ref0"], t["ref1"], t["mv0x"] = 16, 16, 1, 0, -1, -4 * 120
    with pytest.raises(tb.ThorCudaError):
        tb.ops.predict_blocks([ref, dst], t, 1)
    ref.close()
    dst.close()


# -------------------------------------------------- transform / quantise / reconstruct ----
def oracle_tb(orc, org_pl, pred_pl, t, decode_coeffq=None):
    """encode_and_reconstruct_block_inter for one task; returns (rec block, coeffq corner, cbp, ssd)."""
    n, x, y, qp = int(t["size"]), int(t["x"]), int(t["y"]), int(t["qp"])
    q = min(n, 16)
    pred = np.ascontiguousarray(pred_pl[y:y + n, x:x + n])
    rec = pred.copy()
    if decode_coeffq is None:
        org = np.ascontiguousarray(org_pl[y:y + n, x:x + n])
        res = np.zeros((n, n), np.int16)
        orc.orc_get_residual(ptr(res), ptr(pred), ptr(org), n, n)
        coeff = np.zeros((n, n), np.int16)
        orc.orc_transform(ptr(res), ptr(coeff), n, int(t["fast"]))
        cq = np.zeros((n, n), np.int16)
        cbp = orc.orc_quantize(ptr(coeff), ptr(cq), qp, n, int(t["coeff_type"]))
    else:
        cq = np.zeros((n, n), np.int16)
        cq[:q, :q] = decode_coeffq.reshape(q, q)
        cbp = 1
    if cbp:
        dq = np.zeros((n, n), np.int16)
        orc.orc_dequantize(ptr(cq), ptr(dq), qp, n)
        blk = np.zeros((n, n), np.int16)
        orc.orc_inverse_transform(ptr(dq), ptr(blk), n)
        orc.orc_reconstruct_block(ptr(blk), ptr(pred), ptr(rec), n, n)
    ssd = 0
    if decode_coeffq is None:
        ssd = int(((org.astype(np.int64) - rec.astype(np.int64)) ** 2).sum())
    return rec, cq[:q, :q].copy(), int(cbp), ssd


def make_tb_tasks(tb, w, h, rng, qps, sizes_luma=(8, 64), intra_frame=False):
    tasks = []
    for pl in range(3):
        pw, ph = (w, h) if pl == 0 else (w // 2, h // 2)
        lo = sizes_luma[0] if pl == 0 else max(4, sizes_luma[0] // 2)
        hi = sizes_luma[1] if pl == 0 else sizes_luma[1] // 2
        blocks = []
        for (x, y, s) in partition(pw, ph, rng, min_size=lo):
            if s > hi:      # larger than the biggest transform of this plane: tile it
                blocks += [(x + dx, y + dy, hi) for dy in range(0, s, hi) for dx in range(0, s, hi)]
            else:
                blocks.append((x, y, s))
        for (x, y, s) in blocks:
            t = np.zeros((), tb.TB_TASK)
            t["x"], t["y"], t["size"], t["plane"] = x, y, s, pl
            t["qp"] = rng.choice(qps)
            t["coeff_type"] = (1 if pl else 0) | (2 if intra_frame else 0)
            t["fast"] = int(rng.random() < 0.5)
            t["org"], t["pred"], t["rec"] = 0, 1, 2
            tasks.append(t)
    tasks = _arr(tasks, tb.TB_TASK)
    rng.shuffle(tasks)
    return tasks


def tb_pictures(tb, w, h, seed, noise):
    rng = np.random.default_rng(seed)
    oy, ou, ov = smooth_frame(w, h, seed)
    org = tb.Picture(w, h)
    org.upload(oy, ou, ov)
    ppl = []
    for p in (oy, ou, ov):
        d = rng.normal(0, noise, p.shape) * (rng.random(p.shape) < 0.5)
        # flat areas (zero residual), busy areas, and a few saturated samples
        q = np.clip(p.astype(np.float64) + d, 0, 255).astype(np.uint8)
        q[:p.shape[0] // 5] = p[:p.shape[0] // 5]
        q[-8:, -8:] = 255 - p[-8:, -8:]
        ppl.append(q)
    pred = tb.Picture(w, h)
    pred.upload(*ppl)
    return (oy, ou, ov), ppl, org, pred


@pytest.mark.parametrize("wh,qps,intra", [((256, 192), (22, 32, 37), False), ((256, 192), (0, 13, 51), True),
                                          ((416, 240), (30,), False), ((1920, 1080), (32,), False), ((3840, 2160), (32,), False)])
def test_encode_blocks(tb, orc, wh, qps, intra):
    w, h = wh
    rng = np.random.default_rng(w + qps[0])
    opl, ppl, org, pred = tb_pictures(tb, w, h, 5 + qps[0], noise=14)
    rec = tb.Picture(w, h)
    tb.cdll().thorcuda_pic_copy(rec.handle, pred.handle)     # areas no task covers (frame-edge remainders)
    big = w >= 1920
    tasks = make_tb_tasks(tb, w, h, rng, qps, sizes_luma=(16, 64) if big else (8, 64), intra_frame=intra)
    tb.ops.pool_layout(tasks)
    coeffq, cbp, ssd = tb.ops.encode_blocks([org, pred, rec], tasks)
    got = rec.download()
    n_cbp = 0
    step = 9 if big else 1
    for i in range(0, tasks.size, step):
        t = tasks[i]
        pl, n, x, y = int(t["plane"]), int(t["size"]), int(t["x"]), int(t["y"])
        wrec, wcq, wcbp, wssd = oracle_tb(orc, opl[pl], ppl[pl], t)
        q = min(n, 16)
        off = int(t["coeff_offset"])
        assert cbp[i] == wcbp, (i, t)
        assert np.array_equal(coeffq[off:off + q * q].reshape(q, q), wcq), (i, t)
        assert np.array_equal(got[pl][y:y + n, x:x + n], wrec), (i, t)
        assert ssd[i] == wssd, (i, t)
        n_cbp += wcbp
    assert 0 < n_cbp < len(range(0, tasks.size, step)), "test must exercise both coded and uncoded blocks"
    # decoder leg on the same coefficients reproduces the encoder's reconstruction (enc-rec == dec-out)
    coded = tasks[cbp != 0]
    rec2 = tb.Picture(w, h)
    tb.cdll().thorcuda_pic_copy(rec2.handle, pred.handle)
    dtasks = coded.copy()
    dtasks["rec"] = 2
    tb.ops.reconstruct_blocks([org, pred, rec2], dtasks, coeffq)
    got2 = rec2.download()
    for pl in range(3):
        assert np.array_equal(got2[pl], got[pl]), pl
    for p in (org, pred, rec, rec2):
        p.close()


def test_encode_in_place_and_size4(tb, orc):
    """pred == rec (in place), 4x4 luma blocks (tb_split of 8x8), fast flag irrelevant below 32."""
    w, h = 64, 64
    rng = np.random.default_rng(3)
    opl, ppl, org, pred = tb_pictures(tb, w, h, 77, noise=25)
    tasks = []
    for y in range(0, 64, 4):
        for x in range(0, 64, 4):
            t = np.zeros((), tb.TB_TASK)
            t["x"], t["y"], t["size"], t["plane"], t["qp"], t["fast"] = x, y, 4, 0, 24 + (x // 4) % 3, 1
            t["org"], t["pred"], t["rec"] = 0, 1, 1
            tasks.append(t)
    tasks = _arr(tasks, tb.TB_TASK)
    tb.ops.pool_layout(tasks)
    coeffq, cbp, ssd = tb.ops.encode_blocks([org, pred], tasks)
    got = pred.download()
    for i, t in enumerate(tasks):
        x, y = int(t["x"]), int(t["y"])
        wrec, wcq, wcbp, wssd = oracle_tb(orc, opl[0], ppl[0], t)
        off = int(t["coeff_offset"])
        assert cbp[i] == wcbp and ssd[i] == wssd
        assert np.array_equal(coeffq[off:off + 16].reshape(4, 4), wcq)
        assert np.array_equal(got[0][y:y + 4, x:x + 4], wrec)
    assert cbp.any()
    org.close()
    pred.close()


def test_reconstruct_dense_coefficients(tb, orc):
    """Decoder leg with large / dense coefficients (int16 clips in both inverse passes)."""
    w, h = 128, 128
    rng = np.random.default_rng(8)
    _, ppl, org, pred = tb_pictures(tb, w, h, 12, noise=5)
    rec = tb.Picture(w, h)
    tb.cdll().thorcuda_pic_copy(rec.handle, pred.handle)
    tasks, cqs = [], []
    for (x, y, s) in [(0, 0, 64), (64, 0, 32), (96, 0, 32), (64, 32, 16), (80, 32, 16), (64, 48, 8), (72, 48, 8), (64, 56, 4),
                      (0, 64, 64), (64, 64, 64)]:
        t = np.zeros((), tb.TB_TASK)
        t["x"], t["y"], t["size"], t["plane"], t["qp"] = x, y, s, 0, int(rng.choice([20, 32, 51]))
        t["org"], t["pred"], t["rec"] = 0, 1, 2
        tasks.append(t)
        q = min(s, 16)
        amp = int(rng.choice([3, 300, 6000, 32767]))
        cqs.append(rng.integers(-amp, amp + 1, (q, q)).astype(np.int16))
    tasks = _arr(tasks, tb.TB_TASK)
    total = tb.ops.pool_layout(tasks)
    pool = np.zeros(total, np.int16)
    for t, cq in zip(tasks, cqs):
        pool[int(t["coeff_offset"]):int(t["coeff_offset"]) + cq.size] = cq.reshape(-1)
    tb.ops.reconstruct_blocks([org, pred, rec], tasks, pool)
    got = rec.download()
    for t, cq in zip(tasks, cqs):
        x, y, n = int(t["x"]), int(t["y"]), int(t["size"])
        wrec, _, _, _ = oracle_tb(orc, None, ppl[0], t, decode_coeffq=cq)
        assert np.array_equal(got[0][y:y + n, x:x + n], wrec), t
    for p in (org, pred, rec):
        p.close()


# ------------------------------------------------------------------------ SAD tables ----
def numpy_sad_tables(org_y, ref_pic, R, centres):
    """All four levels by brute force.  ref_pic: host Picture (padded)."""
    h, w = org_y.shape
    n = 2 * R                                     # displacements -R .. R-1
    pad = ref_pic.pad_y
    full = ref_pic.buf_y.astype(np.int16)
    out = {bs: np.zeros((h // bs, w // bs, n, n), np.uint16 if bs <= 16 else np.uint32) for bs in (8, 16, 32, 64)}
    nsbx = (w + 63) // 64
    for y0 in range(0, h, 64):
        for x0 in range(0, w, 64):
            sb = (y0 // 64) * nsbx + x0 // 64
            cx, cy = (0, 0) if centres is None else (int(centres[2 * sb]), int(centres[2 * sb + 1]))
            bh, bw = min(64, h - y0), min(64, w - x0)
            o = org_y[y0:y0 + bh, x0:x0 + bw].astype(np.int16)
            win = full[pad + y0 + cy - R:pad + y0 + cy + R - 1 + bh, pad + x0 + cx - R:pad + x0 + cx + R - 1 + bw]
            sw = np.lib.stride_tricks.sliding_window_view(win, (bh, bw))           # [n, n, bh, bw]
            ad = np.abs(sw - o[None, None]).astype(np.uint32)
            s8 = ad.reshape(n, n, bh // 8, 8, bw // 8, 8).sum(axis=(3, 5))           # [n, n, bh/8, bw/8]
            s8 = np.moveaxis(s8, (2, 3), (0, 1))                                    # [by, bx, n, n]
            out[8][y0 // 8:y0 // 8 + bh // 8, x0 // 8:x0 // 8 + bw // 8] = s8
            for bs, f in ((16, 2), (32, 4), (64, 8)):
                ny, nx = bh // bs, bw // bs
                if ny == 0 or nx == 0:
                    continue
                s = s8[:ny * f, :nx * f].reshape(ny, f, nx, f, n, n).sum(axis=(1, 3))
                out[bs][y0 // bs:y0 // bs + ny, x0 // bs:x0 // bs + nx] = s
    return out


@pytest.mark.parametrize("wh,R,with_centres", [((416, 240), 8, False), ((200, 136), 16, True), ((128, 128), 32, True),
                                               ((1920, 1080), 16, False), ((3840, 2160), 16, False)])
def test_sad_table(tb, orc, wh, R, with_centres):
    w, h = wh
    rng = np.random.default_rng(w + R)
    href = textured_ref(w, h, 3)
    oy, ou, ov = smooth_frame(w, h, 3, shift=(2, -3))
    dorg = tb.Picture(w, h)
    dorg.upload(oy, ou, ov)
    dref = dev_ref_from_host(tb, href)
    nsb = ((w + 63) // 64) * ((h + 63) // 64)
    centres = None
    if with_centres:
        lim = 96 - R                                # window must stay inside the 96-pixel border
        centres = rng.integers(-lim, lim + 1, 2 * nsb).astype(np.int16)
    got = tb.ops.sad_table(dorg, dref, R, centres)
    if w >= 1920:
        # full size: spot-check against sad_calc of the oracle + the additivity property between levels
        n = 2 * R
        for _ in range(300):
            bs = int(rng.choice([8, 16, 32, 64]))
            by, bx = int(rng.integers(0, h // bs)), int(rng.integers(0, w // bs))
            dy, dx = int(rng.integers(-R, R)), int(rng.integers(-R, R))
            a = np.ascontiguousarray(oy[by * bs:(by + 1) * bs, bx * bs:(bx + 1) * bs])
            b = C.c_void_p(href.origin_ptr(0) + (by * bs + dy) * href.stride_y + bx * bs + dx)
            assert got[bs][by, bx, dy + R, dx + R] == orc.orc_sad(ptr(a), b, bs, href.stride_y, bs, bs)
        s8 = got[8].astype(np.uint32)
        hy, hx = (h // 16) * 2, (w // 16) * 2
        s16 = s8[:hy, :hx].reshape(hy // 2, 2, hx // 2, 2, n, n).sum(axis=(1, 3))
        assert np.array_equal(s16, got[16].astype(np.uint32))
    else:
        want = numpy_sad_tables(oy, href, R, centres)
        for bs in (8, 16, 32, 64):
            assert np.array_equal(got[bs], want[bs]), bs
        # pin the numpy brute force itself to the oracle's sad_calc on a few entries
        for _ in range(20):
            bs = int(rng.choice([8, 16]))
            by, bx = int(rng.integers(0, h // bs)), int(rng.integers(0, w // bs))
            dy, dx = int(rng.integers(-R, R)), int(rng.integers(-R, R))
            sb = (by * bs // 64) * ((w + 63) // 64) + bx * bs // 64
            cx, cy = (0, 0) if centres is None else (int(centres[2 * sb]), int(centres[2 * sb + 1]))
            a = np.ascontiguousarray(oy[by * bs:(by + 1) * bs, bx * bs:(bx + 1) * bs])
            b = C.c_void_p(href.origin_ptr(0) + (by * bs + cy + dy) * href.stride_y + bx * bs + cx + dx)
            assert want[bs][by, bx, dy + R, dx + R] == orc.orc_sad(ptr(a), b, bs, href.stride_y, bs, bs)
    dorg.close()
    dref.close()


def test_sad_table_batch_two_pairs(tb):
    """Two (org, ref) pairs of different size in one launch land in their own slots."""
    res = []
    pics = []
    for k, (w, h) in enumerate(((192, 128), (128, 64))):
        href = textured_ref(w, h, 40 + k)
        oy, ou, ov = smooth_frame(w, h, 40 + k, shift=(1, 1))
        dorg = tb.Picture(w, h)
        dorg.upload(oy, ou, ov)
        dref = dev_ref_from_host(tb, href)
        pics += [dorg, dref]
        res.append((oy, href))
    before = tb.launch_count()
    tb.ops.sad_table_build([5, 9], [pics[0], pics[2]], [pics[1], pics[3]], 8)
    assert tb.launch_count() - before == 1
    for slot, (oy, href) in zip((5, 9), res):
        h, w = oy.shape
        got = tb.ops.sad_table_fetch(slot, w, h, 8)
        want = numpy_sad_tables(oy, href, 8, None)
        for bs in (8, 16, 32, 64):
            assert np.array_equal(got[bs], want[bs])
    for p in pics:
        p.close()


# ------------------------------------------------------------------------ cost batch ----
def test_cost_batch_all_kinds(tb, orc):
    w, h = 416, 240
    rng = np.random.default_rng(17)
    href = textured_ref(w, h, 9)
    oy, ou, ov = smooth_frame(w, h, 9, shift=(1, 2))
    dorg = tb.Picture(w, h)
    dorg.upload(oy, ou, ov)
    dref = dev_ref_from_host(tb, href)
    shapes = [(8, 8), (16, 16), (32, 32), (64, 64), (16, 8), (8, 16), (4, 8), (8, 4), (4, 4), (32, 16), (64, 32)]
    tasks = []
    for i in range(600):
        bw, bh = shapes[i % len(shapes)]
        t = np.zeros((), tb.COST_TASK)
        t["w"], t["h"] = bw, bh
        t["x"] = int(rng.integers(0, (w - bw) // 4 + 1)) * 4
        t["y"] = int(rng.integers(0, (h - bh) // 4 + 1)) * 4
        t["org"], t["ref"] = 0, 1
        t["ix"] = int(np.clip(rng.normal(0, 12), -(int(t["x"]) + 80), w + 80 - int(t["x"]) - bw))
        t["iy"] = int(np.clip(rng.normal(0, 12), -(int(t["y"]) + 80), h + 80 - int(t["y"]) - bh))
        t["hx"], t["hy"] = int(rng.choice([-2, 0, 2])), int(rng.choice([-2, 0, 2]))
        tasks.append(t)
    tasks = _arr(tasks, tb.COST_TASK)
    oyc = np.ascontiguousarray(oy)
    for kind in (tb.COST_SAD, tb.COST_WIDESAD, tb.COST_FASTHALF, tb.COST_FASTQUARTER, tb.COST_SSD):
        got = tb.ops.cost_batch(kind, [dorg, dref], tasks)
        for t, g in zip(tasks, got):
            x, y, bw, bh = int(t["x"]), int(t["y"]), int(t["w"]), int(t["h"])
            a = C.c_void_p(oyc.ctypes.data + y * w + x)
            b = C.c_void_p(href.origin_ptr(0) + (y + int(t["iy"])) * href.stride_y + x + int(t["ix"]))
            rx, ry = C.c_int(int(t["hx"])), C.c_int(int(t["hy"]))
            if kind == tb.COST_SAD:
                want = (orc.orc_sad(a, b, w, href.stride_y, bw, bh), 0, 0)
            elif kind == tb.COST_SSD:
                want = (orc.orc_ssd(a, b, w, href.stride_y, bw, bh), 0, 0)
            elif kind == tb.COST_WIDESAD:
                want = (orc.orc_widesad(a, b, w, href.stride_y, bw, bh, C.byref(rx)), rx.value, 0)
            elif kind == tb.COST_FASTHALF:
                want = (orc.orc_sad_fasthalf(a, b, w, href.stride_y, bw, bh, C.byref(rx), C.byref(ry)), rx.value, ry.value)
            else:
                want = (orc.orc_sad_fastquarter(a, b, w, href.stride_y, bw, bh, C.byref(rx), C.byref(ry)), rx.value, ry.value)
            assert (int(g["cost"]), int(g["x"]), int(g["y"])) == want, (kind, t)
    dorg.close()
    dref.close()


def test_resident_list_reuse(tb, orc):
    """A list made resident once gives identical results on every run (the device-resident bench path)."""
    w, h = 128, 64
    rng = np.random.default_rng(4)
    href = textured_ref(w, h, 2)
    oy, ou, ov = smooth_frame(w, h, 2)
    dorg = tb.Picture(w, h)
    dorg.upload(oy, ou, ov)
    dref = dev_ref_from_host(tb, href)
    tasks = np.zeros(32, tb.COST_TASK)
    tasks["w"], tasks["h"], tasks["org"], tasks["ref"] = 16, 16, 0, 1
    tasks["x"] = (np.arange(32) % 8) * 16
    tasks["y"] = (np.arange(32) // 8) * 16
    tasks["ix"] = rng.integers(-8, 9, 32)
    tasks["iy"] = rng.integers(-8, 9, 32)
    lst = tb.ops.TaskList(tb.LIST_COST, [dorg, dref], tasks)
    a = np.zeros(32, tb.COST_RESULT)
    b = np.zeros(32, tb.COST_RESULT)
    lst.cost(tb.COST_SAD, a)
    lst.cost(tb.COST_SAD, b)
    tb.sync()
    assert np.array_equal(a, b) and a["cost"].any()
    with pytest.raises(tb.ThorCudaError):
        lst.predict(1)            # wrong kind of list: refused, not silently ignored
    lst.close()
    dorg.close()
    dref.close()


# ------------------------------------------------------------------- early-skip test (a19) ----
@pytest.mark.parametrize("wh,qp,thr,noise", [((256, 192), 32, 1.0, 3), ((416, 240), 24, 0.8, 2), ((1920, 1080), 37, 1.3, 6)])
def test_early_skip_batch(tb, orc, wh, qp, thr, noise):
    """check_early_skip_sub_block / _sub_blockC (enc/encode_block.c:2507-2610) for every 16x16 and 8x8 sub-block
    of a frame; thresholds from the library's host helper against the oracle's (the reference's double arithmetic)."""
    from thor_b200.lib import SKIP_TASK
    from thor_b200.ops import early_skip_batch, early_skip_thresholds
    w, h = wh
    opl, ppl, org, pred = tb_pictures(tb, w, h, 90 + qp, noise=noise)
    rng = np.random.default_rng(qp)
    for i, p in enumerate(ppl):            # region-wise DC errors so that the (one-sided) chroma test fires in places
        gy, gx = (p.shape[0] + 15) // 16, (p.shape[1] + 15) // 16
        dc = np.kron(rng.integers(-3 * noise, 3 * noise + 1, (gy, gx)) * (rng.random((gy, gx)) < 0.3), np.ones((16, 16), int))
        ppl[i] = np.clip(p.astype(int) + dc[:p.shape[0], :p.shape[1]], 0, 255).astype(np.uint8)
    pred.upload(*ppl)
    orc.orc_early_skip_threshold.argtypes = [C.c_int, C.c_int, C.c_double]
    for f in (orc.orc_early_skip_luma, orc.orc_early_skip_chroma):
        f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_float]
    thr32 = float(np.float32(thr))
    tasks = []
    for size in (16, 8):
        ty, tc = early_skip_thresholds(qp, size, thr32)
        assert ty == orc.orc_early_skip_threshold(qp, size // 2, 0.5 * thr32)
        for y in range(0, h - size + 1, size):
            for x in range(0, w - size + 1, size):
                tasks.append((x, y, size, 0, 1, 0, ty, tc))
    if w * h > 500000:                                     # oracle loop in Python: bound the big frame
        sel = np.random.default_rng(1).choice(len(tasks), 6000, replace=False)
        tasks = [tasks[i] for i in sorted(sel)]
    t = np.array(tasks, SKIP_TASK)
    got = early_skip_batch([org, pred], t)
    qpc = orc.orc_chroma_qp(qp)
    want = np.zeros(len(tasks), np.uint8)
    for i, (x, y, size, *_r) in enumerate(tasks):
        pb = np.ascontiguousarray(ppl[0][y:y + size, x:x + size])
        f = orc.orc_early_skip_luma(ptr(opl[0][y:, x:]), opl[0].strides[0], size, qp, ptr(pb), thr32)
        cs, cx, cy = size // 2, x // 2, y // 2
        for pl in (1, 2):
            pc = np.ascontiguousarray(ppl[pl][cy:cy + cs, cx:cx + cs])
            f |= orc.orc_early_skip_chroma(ptr(opl[pl][cy:, cx:]), opl[pl].strides[0], cs, qpc, ptr(pc), thr32) << pl
        want[i] = f
    assert np.array_equal(got, want)
    for bit in (1, 2, 4):                                    # both outcomes occur for every plane
        assert 0 < np.count_nonzero(want & bit) < len(want)
    org.close(); pred.close()


@pytest.mark.parametrize("qp", [0, 20, 32, 51])
def test_encode_blocks_extreme_residuals(tb, orc, qp):
    """SURVEY 8c's adversarial inputs for the transform chain: residuals of +255 / -255 everywhere, a +-255 checkerboard,
    +-255 columns, one impulse and zero, for every transform size and both quantiser types -- the int16 intermediate
    clips, the dead zone at qp 51 and the full coefficient range at qp 0."""
    w, h = 512, 128
    sizes = [(0, 0, 64), (64, 0, 32), (96, 0, 16), (112, 0, 8), (120, 0, 4), (64, 32, 32), (96, 16, 16), (112, 8, 8), (120, 4, 4)]
    yy, xx = np.mgrid[0:h, 0:w]
    patterns = []                                    # (org, pred) luma planes; residual = org - pred
    patterns.append((np.full((h, w), 255), np.zeros((h, w))))
    patterns.append((np.zeros((h, w)), np.full((h, w), 255)))
    cb = ((xx + yy) & 1) * 255
    patterns.append((cb, 255 - cb))
    col = (xx & 1) * 255
    patterns.append((col, 255 - col))
    imp = np.zeros((h, w)); imp[::4, ::4] = 255
    patterns.append((imp, np.zeros((h, w))))
    patterns.append((np.full((h, w), 77), np.full((h, w), 77)))
    total = 0
    for k, (o, p) in enumerate(patterns):
        oy, py = o.astype(np.uint8), p.astype(np.uint8)
        oc, pc = np.ascontiguousarray(oy[::2, ::2]), np.ascontiguousarray(py[::2, ::2])
        org, pred, rec = tb.Picture(w, h), tb.Picture(w, h), tb.Picture(w, h)
        org.upload(oy, oc, oc)
        pred.upload(py, pc, pc)
        tasks = []
        for rep, (x0, fast, ctype) in enumerate([(0, 0, 0), (128, 1, 2)]):       # inter / intra quantiser, fast flag for 32 and 64
            for (x, y, s) in sizes:
                t = np.zeros((), tb.TB_TASK)
                t["x"], t["y"], t["size"], t["plane"], t["qp"], t["coeff_type"], t["fast"] = x0 + x, y, s, 0, qp, ctype, fast
                t["org"], t["pred"], t["rec"] = 0, 1, 2
                tasks.append(t)
        for (x, y, s) in [(0, 0, 32), (32, 0, 16), (48, 0, 8), (56, 0, 4)]:      # chroma
            t = np.zeros((), tb.TB_TASK)
            t["x"], t["y"], t["size"], t["plane"], t["qp"], t["coeff_type"] = x, y, s, 1, qp, 1
            t["org"], t["pred"], t["rec"] = 0, 1, 2
            tasks.append(t)
        tasks = _arr(tasks, tb.TB_TASK)
        tb.ops.pool_layout(tasks)
        coeffq, cbp, ssd = tb.ops.encode_blocks([org, pred, rec], tasks)
        got = rec.download()
        for i, t in enumerate(tasks):
            pl, n, x, y = int(t["plane"]), int(t["size"]), int(t["x"]), int(t["y"])
            wrec, wcq, wcbp, wssd = oracle_tb(orc, (oy, oc, oc)[pl], (py, pc, pc)[pl], t)
            q, off = min(n, 16), int(t["coeff_offset"])
            assert cbp[i] == wcbp and ssd[i] == wssd, (k, i, t)
            assert np.array_equal(coeffq[off:off + q * q].reshape(q, q), wcq), (k, i, t)
            assert np.array_equal(got[pl][y:y + n, x:x + n], wrec), (k, i, t)
            total += 1
        for pic in (org, pred, rec):
            pic.close()
    assert total == len(patterns) * 22


def test_list_outliving_its_pictures_is_refused(tb):
    """a resident list names pictures by handle and the kernels never clamp: once a picture of the list is destroyed (or its
    handle recycled by a smaller one) every launch of the list must fail instead of touching freed memory"""
    ref, dst = tb.Picture(128, 64, 96, 48), tb.Picture(128, 64)
    tasks = np.zeros(2, tb.PRED_TASK)
    tasks[0] = (0, 0, 16, 16, 0, -1, 40, -20, 0, 0, 0, 0, 1, 0)
    tasks[1] = (64, 32, 16, 16, 0, -1, 300, 100, 0, 0, 0, 0, 1, 0)       # legal only with the 96-sample border
    lst = tb.ops.TaskList(tb.LIST_PRED, [ref, dst], tasks)
    lst.predict(1)
    tb.sync()
    ref.close()
    with pytest.raises(tb.ThorCudaError, match="destroyed or recreated"):
        lst.predict(1)
    small = tb.Picture(64, 64, 16, 8)                                    # takes the freed handle, smaller geometry
    with pytest.raises(tb.ThorCudaError, match="destroyed or recreated"):
        lst.predict(1)
    lst.close(); small.close(); dst.close()
