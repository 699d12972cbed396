"""GPU parity (bit-exact, through the C ABI) of the intra prediction kernel against the CPU oracle, and a whole
intra frame reconstructed on the device in wavefront order against the oracle's serial decode order
(dec/decode_block.c:48-88: neighbours -> prediction -> dequantise -> inverse transform -> reconstruct, block by block)."""
import ctypes as C

import numpy as np
import pytest

from thorref import load_oracle, ptr

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


def oracle_intra(orc, plane, x, y, cb_x, size, mode, flags):
    left, top = np.zeros(2 * size + 2, np.uint8), np.zeros(2 * size + 2, np.uint8)
    tl = C.c_uint8(0)
    orc.orc_intra_neighbours(ptr(plane[y:, x:]), plane.strides[0], x, y, cb_x, size, flags & 1, (flags >> 1) & 1, ptr(left), ptr(top), C.byref(tl))
    out = np.zeros((size, size), np.uint8)
    orc.orc_intra_predict(ptr(left), ptr(top), tl.value, y, x, size, ptr(out), mode)
    return out


@pytest.mark.parametrize("wh", [(192, 128), (416, 240)])
def test_intra_predict_blocks(tb, orc, wh):
    """every size x mode (incl. an unknown one) x plane, blocks on all four picture edges, every legal flag combination"""
    from thor_b200.lib import INTRA_TASK
    w, h = wh
    rng = np.random.default_rng(w)
    planes = [rng.integers(0, 256, (h, w), dtype=np.uint8), rng.integers(0, 256, (h // 2, w // 2), dtype=np.uint8),
              rng.integers(0, 256, (h // 2, w // 2), dtype=np.uint8)]
    src, dst = tb.Picture(w, h), tb.Picture(w, h)
    src.upload(*planes)
    n = 0
    for size in (4, 8, 16, 32, 64):
        tasks = []
        for pl in range(3):
            pw, ph = (w, h) if pl == 0 else (w // 2, h // 2)
            k = 0
            for y in range(0, ph - size + 1, size):
                for x in range(0, pw - size + 1, size):
                    up_ok, dl_ok = y > 0 and x + size < pw, x > 0 and y + size < ph
                    flags = (int(rng.integers(0, 2)) if up_ok else 0) | ((int(rng.integers(0, 2)) << 1) if dl_ok else 0)
                    cb_x = x if rng.random() < 0.7 else int(rng.integers(0, x // 4 + 1)) * 4      # incl. the "CB at the left edge" corner rule
                    tasks.append((x, y, cb_x, size, pl, (k + pl) % 11, flags, 0, 1))
                    k += 1
        t = np.array(tasks, INTRA_TASK)
        tb.ops.intra_predict_blocks([src, dst], t)
        tb.sync()
        got = dst.download()
        for (x, y, cb_x, s, pl, mode, flags, _a, _b) in tasks:
            want = oracle_intra(orc, planes[pl], x, y, cb_x, s, mode, flags)
            assert np.array_equal(got[pl][y:y + s, x:x + s], want), (x, y, cb_x, s, pl, mode, flags)
            n += 1
    assert n > 500
    src.close(); dst.close()


def test_intra_rejects_bad_flags(tb):
    from thor_b200.lib import INTRA_TASK
    p = tb.Picture(64, 64)
    for bad in [(56, 8, 56, 8, 0, 0, 1, 0, 0),        # up-right beyond the right edge
                (8, 56, 8, 8, 0, 0, 2, 0, 0),         # down-left beyond the bottom edge
                (8, 8, 12, 8, 0, 0, 0, 0, 0),         # coding block right of its own sub-block
                (8, 8, 8, 12, 0, 0, 0, 0, 0)]:        # size
        with pytest.raises(tb.ThorCudaError):
            tb.ops.intra_predict_blocks([p], np.array([bad], INTRA_TASK))
    p.close()


# --------------------------------------------------------------------- whole intra frame ----
def zorder_cbs(x, y, size, w, h, rng, out):
    """coding blocks of one SB in decode order (quad-tree, z-scan), random split decisions"""
    if x >= w or y >= h:
        return
    if size > 8 and (x + size > w or y + size > h or rng.random() < (0.75 if size == 64 else 0.45)):
        half = size // 2
        for dy in (0, half):
            for dx in (0, half):
                zorder_cbs(x + dx, y + dy, half, w, h, rng, out)
    else:
        out.append((x, y, size))


def intra_frame_tbs(orc, w, h, rng):
    """transform blocks of a synthetic intra frame in decode order: (plane, x, y, cb_x, size, mode, flags, qp)"""
    tbs = []
    for sy in range(0, h, 64):
        for sx in range(0, w, 64):
            cbs = []
            zorder_cbs(sx, sy, 64, w, h, rng, cbs)
            for (x, y, s) in cbs:
                mode = int(rng.integers(0, 10))
                up, dl = orc.orc_upright_available(y, x, s, w), orc.orc_downleft_available(y, x, s, h)
                split = s <= 32 and rng.random() < 0.35          # tb_split: four sub-blocks in raster order
                qp = int(rng.choice([22, 32, 40]))
                for pl in range(3):
                    px, py, ps = (x, y, s) if pl == 0 else (x // 2, y // 2, s // 2)
                    psplit = split if pl == 0 else (split and s > 8)         # dec/decode_block.c:238
                    pqp = qp if pl == 0 else orc.orc_chroma_qp(qp)
                    if psplit:
                        s2 = ps // 2
                        for i in (0, s2):
                            for j in (0, s2):
                                f = (1 if (j == 0 or (i == 0 and up)) else 0) | ((1 if (j == 0 and (i == 0 or dl)) else 0) << 1)
                                tbs.append((pl, px + j, py + i, px, s2, mode, f, pqp))
                    else:
                        tbs.append((pl, px, py, px, ps, mode, up | (dl << 1), pqp))
    return tbs


def test_intra_frame_wavefront(tb, orc):
    from thor_b200.lib import INTRA_TASK
    w, h = 192, 128
    rng = np.random.default_rng(2024)
    tbs = intra_frame_tbs(orc, w, h, rng)
    assert {t[4] for t in tbs} >= {4, 8, 16, 32}
    # coefficients: sparse low-frequency levels (some blocks empty, some dense)
    coefs = []
    for (pl, x, y, cbx, s, mode, f, qp) in tbs:
        q = min(s, 16)
        c = np.zeros((q, q), np.int16)
        r = rng.random()
        if r > 0.25:
            k = q if r > 0.9 else max(1, q // 4)
            c[:k, :k] = rng.integers(-6, 7, (k, k)) * (rng.random((k, k)) < 0.5)
            c[0, 0] = rng.integers(-40, 41)
        coefs.append(c)

    # ---- oracle: serial decode order ----
    planes = [np.zeros((h, w), np.uint8), np.zeros((h // 2, w // 2), np.uint8), np.zeros((h // 2, w // 2), np.uint8)]
    for (pl, x, y, cbx, s, mode, f, qp), cq in zip(tbs, coefs):
        P = planes[pl]
        pred = oracle_intra(orc, P, x, y, cbx, s, mode, f)
        full = np.zeros((s, s), np.int16)
        full[:cq.shape[0], :cq.shape[1]] = cq
        dq, blk = np.zeros((s, s), np.int16), np.zeros((s, s), np.int16)
        orc.orc_dequantize(ptr(full), ptr(dq), qp, s)
        orc.orc_inverse_transform(ptr(dq), ptr(blk), s)
        rec = np.zeros((s, s), np.uint8)
        orc.orc_reconstruct_block(ptr(blk), ptr(pred), ptr(rec), s, s)
        P[y:y + s, x:x + s] = rec

    # ---- device: wavefront levels.  level(block) = 1 + max level of the blocks that own the samples it reads ----
    lvl_map = [np.zeros((h // 4, w // 4), np.int32), np.zeros((h // 8, w // 8), np.int32), np.zeros((h // 8, w // 8), np.int32)]
    levels = {}
    for idx, (pl, x, y, cbx, s, mode, f, qp) in enumerate(tbs):
        M = lvl_map[pl]
        u = s // 4                                   # block size in 4-sample units
        ux, uy = x // 4, y // 4
        dep = 0
        if y > 0:
            x1 = ux + u + (1 if f & 1 else 0)
            dep = max(dep, int(M[uy - 1, max(ux - 1, 0):x1].max()))
        if x > 0:
            y1 = uy + u + (1 if f & 2 else 0)
            dep = max(dep, int(M[max(uy - 1, 0):y1, ux - 1].max()))
        lv = dep + 1
        M[uy:uy + u, ux:ux + u] = lv
        levels.setdefault(lv, []).append(idx)
    assert len(levels) < len(tbs) / 3, "the wavefront must expose parallelism"

    rec = tb.Picture(w, h)
    rec.upload(*[np.zeros_like(p) for p in planes])
    for lv in sorted(levels):
        idxs = levels[lv]
        it = np.array([(tbs[i][1], tbs[i][2], tbs[i][3], tbs[i][4], tbs[i][0], tbs[i][5], tbs[i][6], 0, 0) for i in idxs], INTRA_TASK)
        tb.ops.intra_predict_blocks([rec], it)
        tt = np.zeros(len(idxs), tb.TB_TASK)
        for k, i in enumerate(idxs):
            pl, x, y, cbx, s, mode, f, qp = tbs[i]
            tt[k]["x"], tt[k]["y"], tt[k]["size"], tt[k]["plane"], tt[k]["qp"] = x, y, s, pl, qp
        total = tb.ops.pool_layout(tt)
        pool = np.zeros(total, np.int16)
        for k, i in enumerate(idxs):
            o = int(tt[k]["coeff_offset"])
            pool[o:o + coefs[i].size] = coefs[i].reshape(-1)
        tb.ops.reconstruct_blocks([rec], tt, pool)           # in place: pred = rec = the frame being decoded
    got = rec.download()
    for pl in range(3):
        assert np.array_equal(got[pl], planes[pl]), f"plane {pl}"

    # ---- the same frame in ONE call: tasks and coefficients uploaded once, two launches per level ----
    rec.upload(*[np.zeros_like(p) for p in planes])
    order = [i for lv in sorted(levels) for i in levels[lv]]
    it = np.array([(tbs[i][1], tbs[i][2], tbs[i][3], tbs[i][4], tbs[i][0], tbs[i][5], tbs[i][6], 0, 0) for i in order], INTRA_TASK)
    coded = [i for i in order if coefs[i].any()]                   # blocks without levels: the prediction is the reconstruction
    tt = np.zeros(len(coded), tb.TB_TASK)
    for k, i in enumerate(coded):
        pl, x, y, cbx, s, mode, f, qp = tbs[i]
        tt[k]["x"], tt[k]["y"], tt[k]["size"], tt[k]["plane"], tt[k]["qp"] = x, y, s, pl, qp
    total = tb.ops.pool_layout(tt)
    pool = np.zeros(total, np.int16)
    for k, i in enumerate(coded):
        o = int(tt[k]["coeff_offset"])
        pool[o:o + coefs[i].size] = coefs[i].reshape(-1)
    ifirst, tfirst = [0], [0]
    for lv in sorted(levels):
        ifirst.append(ifirst[-1] + len(levels[lv]))
        tfirst.append(tfirst[-1] + sum(1 for i in levels[lv] if coefs[i].any()))
    assert tfirst[-1] == len(coded) and 0 < len(coded) < len(order)
    n0 = tb.launch_count()
    tb.ops.intra_reconstruct_levels([rec], it, ifirst, tt, tfirst, pool)
    assert tb.launch_count() - n0 <= 2 * len(levels)
    got = rec.download()
    for pl in range(3):
        assert np.array_equal(got[pl], planes[pl]), f"one call, plane {pl}"
    rec.close()


def test_intra_trial_costs(tb, orc):
    """The encoder's intra trial loop (enc/encode_block.c:1230-1329, 2224-2310: every mode of a block is predicted,
    trial-encoded and priced) as three batched calls on resident lists: intra_predict_blocks -> encode_list (ssd) ->
    coeff_bits_list (write_coeff's bits).  Ten modes of every block at once, mode m of all blocks in scratch picture m.
    SSD, bits and therefore cost_calc's ssd + (int)(lambda * nbits + 0.5) ranking agree with the oracle for every trial."""
    from test_gpu_blocks import oracle_tb
    from thor_b200.lib import INTRA_TASK
    w, h, qp, lam = 256, 192, 30, 38.5
    rng = np.random.default_rng(11)
    yy, xx = np.mgrid[0:h, 0:w]
    # horizontal stripes under the 8x8 blocks, vertical stripes under the 16x16 ones, a smooth ramp under the 32x32 ones
    base = np.where(xx < 56, 128 + 60 * np.sin(yy / 2.5), np.where(xx < 136, 128 + 60 * np.sin(xx / 2.5), 60 + 0.5 * xx + 0.3 * yy)).astype(np.float64)
    org = [np.clip(base + rng.normal(0, 6, (h, w)), 0, 255).astype(np.uint8),
           rng.integers(90, 170, (h // 2, w // 2)).astype(np.uint8), rng.integers(90, 170, (h // 2, w // 2)).astype(np.uint8)]
    rec = [np.clip(o.astype(np.int32) + rng.integers(-3, 4, o.shape), 0, 255).astype(np.uint8) for o in org]   # "already coded" surroundings
    d_org, d_rec = tb.Picture(w, h), tb.Picture(w, h)
    d_org.upload(*org)
    d_rec.upload(*rec)
    scratch = [tb.Picture(w, h) for _ in range(10)]
    # disjoint blocks (their trials share the scratch pictures), none on a picture edge
    blocks = [(8, 8, 8), (24, 8, 8), (8, 24, 8), (24, 24, 8), (64, 16, 16), (96, 16, 16), (64, 48, 16), (96, 48, 16),
              (144, 32, 32), (208, 32, 32), (144, 96, 32), (208, 96, 32)]
    # one intra task + one transform task per (block, mode)
    it, tt = [], []
    for (x, y, s) in blocks:
        for m in range(10):
            it.append((x, y, x, s, 0, m, 1 | 2, 0, 1 + m % 5))        # src = pics[0] (rec), dst = the mode's scratch picture
            t = np.zeros((), tb.TB_TASK)
            t["x"], t["y"], t["size"], t["plane"], t["qp"], t["coeff_type"], t["fast"] = x, y, s, 0, qp, 2, 0
            t["org"], t["pred"], t["rec"] = 0, 1 + m, 1 + m              # pics = [org, scratch 0..9]; reconstruct in place
            tt.append(t)
    it = np.array(it, INTRA_TASK)
    for half in (0, 1):                                                 # at most 8 pictures per intra call: five modes each
        sel = np.array([k for k in range(len(it)) if (k % 10) // 5 == half])
        tb.ops.intra_predict_blocks([d_rec] + scratch[5 * half:5 * half + 5], it[sel])
    tasks = np.zeros(len(tt), tb.TB_TASK)
    for k, t in enumerate(tt):
        tasks[k] = t[()]
    tb.ops.pool_layout(tasks)
    lst = tb.ops.TaskList(tb.LIST_TB, [d_org] + scratch, tasks)
    ssd, cbp, bits = np.zeros(lst.ntasks, np.uint32), np.zeros(lst.ntasks, np.uint8), np.zeros(lst.ntasks, np.uint32)
    lst.encode(None, cbp, ssd)
    lst.coeff_bits(bits)
    tb.sync()
    lst.close()
    k = 0
    winners = set()
    for (x, y, s) in blocks:
        cost_gpu, cost_cpu = [], []
        for m in range(10):
            pred = oracle_intra(orc, rec[0], x, y, x, s, m, 3)
            pred_plane = np.zeros((h, w), np.uint8)
            pred_plane[y:y + s, x:x + s] = pred
            _, wcq, wcbp, wssd = oracle_tb(orc, org[0], pred_plane, tasks[k])
            q = min(s, 16)
            wbits = orc.orc_coeff_bits(ptr(np.ascontiguousarray(wcq)), q, s, 2) if wcbp else 0
            assert (int(ssd[k]), int(cbp[k]), int(bits[k])) == (wssd, wcbp, wbits), (x, y, s, m)
            cost_gpu.append(int(ssd[k]) + int(lam * int(bits[k]) + 0.5))
            cost_cpu.append(wssd + int(lam * wbits + 0.5))
            k += 1
        assert cost_gpu == cost_cpu
        winners.add(int(np.argmin(cost_gpu)))
    assert len(winners) >= 3, "the trial costs must discriminate between modes"
    for p in [d_org, d_rec] + scratch:
        p.close()
