#!/usr/bin/env python
"""Generate tests/golden/*.  Run in the BUILD container, where /root/reference is mounted and
`make -C oracle ref` has compiled the unmodified reference into oracle/_ref/ :

    python tests/golden/make_golden.py

The reference (awakecoding/thor) ships no tests, known-answer vectors or fixtures of its own, so the
golden data are outputs of the reference itself on seeded inputs:

  kernels.npz   inputs AND reference outputs of every hot-path function, produced by
                oracle/_ref/libthorref.so with its SIMD dispatch on (the plain-C path is checked equal here);
  extras.npz    the same for row a19 (early-skip test) and SURVEY 8f #1 (intra prediction); `--extras-only`
                regenerates just this file;
  bits.npz      SURVEY 8f #3 (bit-cost model): bits the reference's write_coeff / write_mv put into a stream and
                quote_vlc's lengths, on seeded inputs; `--bits-only` regenerates just this file (`--streams-only`: just streams.json);
  streams.json  md5 of the bitstream and of the reconstruction that oracle/_ref/Thorenc writes for three shipped
                operating points on seeded synthetic clips (+ that Thordec reproduces the reconstruction).

tests/test_golden.py checks the oracle port (CPU) and libthorcuda (GPU) against these files; nothing there needs
/root/reference.
"""
import ctypes as C
import hashlib
import json
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from thorref import (BLOCKINFO, COEFF_KINDS, DECODER_STREAMS, THOR_CONFIGS, Picture, blocky_frame, coeff_block, load_ref, ptr, random_blockinfo,  # noqa: E402
                     ref_bitpos, ref_stream, run_thordec, run_thorenc, set_ref_simd, smooth_frame, write_clip)


class Mv(C.Structure):
    _fields_ = [("x", C.c_int16), ("y", C.c_int16)]


def aligned(a):
    raw = np.zeros(a.nbytes + 64, np.uint8)
    off = (-raw.ctypes.data) % 64
    out = raw[off:off + a.nbytes].view(a.dtype).reshape(a.shape)
    out[...] = a
    return out


def residuals(rng, n):
    cb = ((np.add.outer(np.arange(n), np.arange(n)) & 1) * 510 - 255).astype(np.int16)
    return [np.full((n, n), 255, np.int16), np.full((n, n), -255, np.int16), cb,
            np.clip(rng.normal(0, 20, (n, n)).round(), -255, 255).astype(np.int16),
            rng.integers(-255, 256, (n, n)).astype(np.int16)]


def kernels(ref):
    out = {}
    rng = np.random.default_rng(20261104)
    # ---- transforms / quantisation (a10-a13) ----
    for n in (4, 8, 16, 32, 64):
        q = min(n, 16)
        blocks = np.stack(residuals(rng, n))
        out[f"tr{n}_in"] = blocks
        fw, qq, cbp, rec = [], [], [], []
        for blk in blocks:
            blk = aligned(blk)
            for fast in ((0, 1) if n > 16 else (0,)):
                co = aligned(np.zeros((n, n), np.int16))
                ref.transform(ptr(blk), ptr(co), n, fast)
                fw.append(co[:q, :q].copy())
                for qp in (0, 20, 32, 51):
                    for cbt in (0, 1, 2, 3):
                        cq = aligned(np.zeros((n, n), np.int16))
                        cbp.append(ref.quantize(ptr(co), ptr(cq), qp, n, cbt, 0))
                        qq.append(cq[:q, :q].copy())
                    dq = aligned(np.zeros((n, n), np.int16))
                    ref.dequantize(ptr(cq), ptr(dq), qp, n)
                    ib = aligned(np.zeros((n, n), np.int16))
                    ref.inverse_transform(ptr(dq), ptr(ib), n)
                    rec.append(ib.copy())
        out[f"tr{n}_fwd"], out[f"tr{n}_q"], out[f"tr{n}_cbp"], out[f"tr{n}_rec"] = np.stack(fw), np.stack(qq), np.array(cbp, np.int32), np.stack(rec)
    # ---- inter prediction (a6-a8') ----
    src = rng.integers(0, 256, (48, 48), dtype=np.uint8)
    src[::7, ::5] = 255
    src[3::11, 1::3] = 0
    out["pred_src"] = src
    src = aligned(src)
    base = src.ctypes.data + 16 * 48 + 16
    luma = np.zeros((2, 4, 4, 16, 16), np.uint8)
    for bip in (0, 1):
        for fy in range(4):
            for fx in range(4):
                mv = Mv(fx - 8, fy + 4)          # integer part (-2, +1)
                o = aligned(np.zeros((16, 16), np.uint8))
                ref.get_inter_prediction_luma(ptr(o), C.c_void_p(base), 16, 16, 48, 16, C.byref(mv), 0, bip)
                luma[bip, fy, fx] = o
    out["pred_luma"] = luma
    chroma = np.zeros((8, 8, 8, 8), np.uint8)
    for fy in range(8):
        for fx in range(8):
            mv = Mv(fx + 8, fy - 16)             # integer part (+1, -2)
            o = aligned(np.zeros((8, 8), np.uint8))
            ref.get_inter_prediction_chroma(ptr(o), C.c_void_p(base), 8, 8, 48, 8, C.byref(mv), 0)
            chroma[fy, fx] = o
    out["pred_chroma"] = chroma
    # ---- SAD family (a1-a5) ----
    win = smooth_frame(64, 64, seed=4)[0]
    out["sad_ref"] = win
    win = aligned(win)
    res = []
    orgs = []
    for k, (w, h) in enumerate(((8, 8), (16, 16), (32, 32), (16, 8), (8, 16))):
        oy, ox = 12 + k, 14 + 2 * k
        org = aligned(np.clip(win[oy:oy + h, ox:ox + w].astype(int) + rng.integers(-9, 10, (h, w)), 0, 255).astype(np.uint8))
        padded = np.zeros((32, 32), np.uint8)
        padded[:h, :w] = org
        orgs.append(padded)
        b = C.c_void_p(win.ctypes.data + (oy + 1) * 64 + ox - 1)
        for f in ("sad_calc", "widesad_calc", "sad_calc_fasthalf", "sad_calc_fastquarter"):
            getattr(ref, f).restype = C.c_uint
        row = [w, h, oy + 1, ox - 1, ref.sad_calc(ptr(org), b, w, 64, w, h)]
        x = C.c_int(0)
        row += [ref.widesad_calc(ptr(org), b, w, 64, w, h, C.byref(x)), x.value]
        x, y = C.c_int(0), C.c_int(0)
        row += [ref.sad_calc_fasthalf(ptr(org), b, w, 64, w, h, C.byref(x), C.byref(y)), x.value, y.value]
        for (px, py) in ((0, 0), (2, 0), (0, 2), (2, 2)):
            x, y = C.c_int(px), C.c_int(py)
            row += [ref.sad_calc_fastquarter(ptr(org), b, w, 64, w, h, C.byref(x), C.byref(y)), x.value, y.value]
        rb = aligned(win[oy + 1:oy + 1 + h, ox - 1:ox - 1 + w].copy())
        row.append(ref.ssd_calc(ptr(org), ptr(rb), w, w, w, h))
        res.append(row)
    out["sad_org"] = np.stack(orgs)
    out["sad_res"] = np.array(res, np.int64)
    # ---- clip_mv (a20) ----
    cm = []
    for _ in range(64):
        fw_, fh_ = (416, 240) if rng.random() < 0.5 else (1920, 1080)
        size = int(rng.choice([8, 16, 32, 64]))
        ypos, xpos = int(rng.integers(0, fh_ // 8)) * 8, int(rng.integers(0, fw_ // 8)) * 8
        sign = int(rng.integers(0, 2))
        mx, my = int(rng.integers(-2000, 2000)), int(rng.integers(-2000, 2000))
        m = Mv(mx, my)
        ref.clip_mv(C.byref(m), ypos, xpos, fw_, fh_, size, sign)
        cm.append([mx, my, ypos, xpos, fw_, fh_, size, sign, m.x, m.y])
    out["clip_mv"] = np.array(cm, np.int32)
    # ---- frame tail: deblock Y/UV, reference creation (a14, a15, a17) on a 192x136 frame (partial SB row/col) ----
    w, h, qp = 192, 136, 32
    y, u, v = blocky_frame(w, h, seed=33)
    bi = random_blockinfo(w, h, np.random.default_rng(7))
    pic = Picture(w, h)
    pic.set_visible(y, u, v)
    f = pic.as_ref_struct()
    chroma_qp = (C.c_int * 52).in_dll(ref, "chroma_qp")
    ref.deblock_frame_y(C.byref(f), ptr(bi), w, h, C.c_uint8(qp))
    ref.deblock_frame_uv(C.byref(f), ptr(bi), w, h, C.c_uint8(chroma_qp[qp]))
    out["lf_in_y"], out["lf_in_u"], out["lf_in_v"] = y, u, v
    out["lf_blockinfo"] = bi.view(np.uint8).reshape(-1, BLOCKINFO.itemsize)
    out["lf_qp"] = np.array([qp, chroma_qp[qp]], np.int32)
    out["lf_out_y"], out["lf_out_u"], out["lf_out_v"] = pic.plane(0).copy(), pic.plane(1).copy(), pic.plane(2).copy()
    refpic = Picture(w, h, 96, 48)
    fr = refpic.as_ref_struct()
    ref.create_reference_frame(C.byref(fr), C.byref(f))
    out["ref_md5"] = np.frombuffer(hashlib.md5(refpic.buf_y[:, :w + 192].tobytes() + refpic.buf_u[:, :w // 2 + 96].tobytes() +
                                               refpic.buf_v[:, :w // 2 + 96].tobytes()).digest(), np.uint8)
    return out


def streams():
    res = {}
    with tempfile.TemporaryDirectory() as d:
        for cfg, w, h, n, qp, seed in (("LDB_low_complexity", 416, 240, 10, 32, 1234), ("LDB_medium_complexity", 192, 136, 6, 30, 7),
                                       ("RA_high_efficiency", 128, 64, 9, 34, 7), ("LDB_low_complexity", 64, 64, 3, 32, 1234),
                                       ("LDB_medium_complexity", 72, 64, 3, 28, 1234), ("RA_high_efficiency", 64, 64, 3, 34, 1234)) + \
                tuple(c + (77,) for c in DECODER_STREAMS):
            yuv = os.path.join(d, "clip.yuv")
            write_clip(yuv, w, h, n, seed)
            bit, rec, _ = run_thorenc("Thorenc", cfg, yuv, w, h, n, qp, os.path.join(d, "o"))
            dec = run_thordec("Thordec", os.path.join(d, "o.bit"), os.path.join(d, "dec.yuv"))
            assert dec == rec, "reference decoder does not reproduce the encoder's reconstruction"
            res[f"{cfg}_{w}x{h}x{n}_qp{qp}_seed{seed}"] = {
                "config": cfg, "width": w, "height": h, "frames": n, "qp": qp, "seed": seed,
                "clip_md5": hashlib.md5(open(yuv, "rb").read()).hexdigest(),
                "bit_bytes": len(bit), "bit_md5": hashlib.md5(bit).hexdigest(), "rec_md5": hashlib.md5(rec).hexdigest()}
    return res


def coding_orders():
    """(frame_num, type) per coded frame as printed by the reference's Thorenc, for thor_b200/segments.py"""
    import subprocess
    from thorref import REF_DIR
    res = {}
    with tempfile.TemporaryDirectory() as d:
        yuv = os.path.join(d, "clip.yuv")
        write_clip(yuv, 64, 64, 70, 3)
        for name, cfg, n, extra in (("RA_n20", "RA_high_efficiency", 20, []), ("RA_p16_n53", "RA_high_efficiency", 53, ["-intra_period", "16"]),
                                    ("RA_n70", "RA_high_efficiency", 70, []), ("LDB_n6", "LDB_low_complexity", 6, []),
                                    ("RA_p16_n48", "RA_high_efficiency", 48, ["-intra_period", "16"])):
            flags = THOR_CONFIGS[cfg]
            cmd = [os.path.join(REF_DIR, "Thorenc"), "-if", yuv, "-of", os.path.join(d, "o.bit"), "-stat", os.path.join(d, "s.txt"),
                   "-width", "64", "-height", "64", "-n", str(n), "-f", "30", "-qp", "40"] + flags + extra
            out = subprocess.run(cmd, capture_output=True, text=True, check=True).stdout
            frames = []
            for ln in out.splitlines():
                f = ln.split()
                if len(f) > 3 and f[0].isdigit() and f[1] in ("I", "P", "B"):
                    frames.append([int(f[0]), f[1]])
            reorder = int(flags[flags.index("-num_reorder_pics") + 1]) if "-num_reorder_pics" in flags else 0
            ip = int((flags + extra)[len(flags + extra) - 1 - (flags + extra)[::-1].index("-intra_period") + 1]) if "-intra_period" in flags + extra else 0
            res[name] = {"num_frames": n, "num_reorder_pics": reorder, "intra_period": ip, "frames": frames}
    return res


def extras(ref):
    """rows a19 (early-skip test) and 8f #1 (intra prediction): inputs and reference outputs"""
    out = {}
    rng = np.random.default_rng(20261105)
    # ---- intra prediction: make_top_and_left + get_intra_prediction on one 128x128 plane ----
    W, H = 128, 128
    plane = rng.integers(0, 256, (H, W), dtype=np.uint8)
    plane[40:, 70:] = (plane[40:, 70:] // 32) * 32 + 100 // 4          # a flatter region
    cases, preds = [], []
    for size in (4, 8, 16, 32, 64):
        pos = [(0, 0), (size, 0), (0, size), (size, size)]
        if size <= 32:
            pos += [(W - size, size), (size, H - size)]
        for (x, y) in pos:
            for mode in range(11):
                up = int(y > 0 and x + size < W and (mode + x // size) % 2 == 0)
                dl = int(x > 0 and y + size < H and (mode + y // size) % 3 != 0)
                left, top = aligned(np.zeros(2 * size + 2, np.uint8)), aligned(np.zeros(2 * size + 2, np.uint8))
                tl = C.c_uint8(0)
                rec = plane[y:, x:]
                ref.make_top_and_left(ptr(left), ptr(top), C.byref(tl), ptr(rec), W, None, 0, 0, 0, y, x, size, up, dl, 0)
                pb = aligned(np.zeros((size, size), np.uint8))
                ref.get_intra_prediction(ptr(left), ptr(top), tl.value, y, x, size, ptr(pb), mode)
                cases.append((x, y, x, size, mode, up | (dl << 1)))
                preds.append(pb.reshape(-1).copy())
        if size <= 32:
            # the four sub-blocks of a tb_split coding block at (64, 32) and at the left picture edge (0, 32)
            for cbx in (64 if size < 32 else 0, 0):
                cby = 32 if 2 * size <= 64 else 0
                for i in (0, size):
                    for j in (0, size):
                        for mode in (1, 4, 7, 8, 0):
                            left, top = aligned(np.zeros(2 * size + 2, np.uint8)), aligned(np.zeros(2 * size + 2, np.uint8))
                            tl = C.c_uint8(0)
                            ref.make_top_and_left(ptr(left), ptr(top), C.byref(tl), ptr(plane[cby:, cbx:]), W, ptr(plane[cby + i:, cbx + j:]), W,
                                                  i, j, cby, cbx, size, 1, 0, 1)
                            pb = aligned(np.zeros((size, size), np.uint8))
                            ref.get_intra_prediction(ptr(left), ptr(top), tl.value, cby + i, cbx + j, size, ptr(pb), mode)
                            up = 1 if (j == 0 or i == 0) else 0          # cb_upright_available = 1, cb_downleft_available = 0
                            dl = 1 if (j == 0 and i == 0) else 0
                            cases.append((cbx + j, cby + i, cbx, size, mode, up | (dl << 1)))
                            preds.append(pb.reshape(-1).copy())
    out["intra_plane"] = plane
    out["intra_cases"] = np.array(cases, np.int32)          # x, y, cb_x, size, mode, flags(bit0 upright, bit1 downleft)
    out["intra_pred"] = np.concatenate(preds)

    # ---- early-skip test: one luma sub-block (16 or 8) + its U and V sub-blocks per case ----
    for f in (ref.check_early_skip_sub_block, ref.check_early_skip_sub_blockC):
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_float]
    chroma_qp = np.ctypeslib.as_array((C.c_int * 52).in_dll(ref, "chroma_qp"))
    rows, org, prd = [], [], []
    for k in range(160):
        size = 16 if k & 1 else 8
        qp = int(rng.integers(16, 48))
        thr = float(np.float32(rng.choice([0.5, 0.8, 1.0, 1.04])))
        amp = int(rng.choice([1, 2, 3, 5, 8, 14]))
        cell_o = rng.integers(0, 256, (3, 16, 16), dtype=np.uint8)
        cell_p = np.clip(cell_o.astype(int) + rng.integers(-amp, amp + 1, (3, 16, 16)) + rng.integers(-amp, amp + 1, (3, 1, 1)), 0, 255).astype(np.uint8)
        flags = 0
        for pl in range(3):
            n = size if pl == 0 else size // 2
            o = aligned(np.ascontiguousarray(cell_o[pl, :n, :n]))
            p = aligned(np.ascontiguousarray(cell_p[pl, :n, :n]))
            if pl == 0:
                r = ref.check_early_skip_sub_block(None, ptr(o), n, n, qp, ptr(p), thr)
            else:
                r = ref.check_early_skip_sub_blockC(None, ptr(o), n, n, int(chroma_qp[qp]), ptr(p), thr)
            flags |= (1 if r else 0) << pl
        rows.append((size, qp, flags))
        org.append(cell_o); prd.append(cell_p)
        rows[-1] = rows[-1] + (int(np.float32(thr).view(np.uint32)),)
    out["skip_cases"] = np.array(rows, np.int64)             # size, qp_y, flags (bit0 Y, bit1 U, bit2 V), float32 bits of the threshold
    out["skip_org"], out["skip_pred"] = np.stack(org), np.stack(prd)
    assert {int(r[2]) & 1 for r in rows} == {0, 1} and any(r[2] & 6 for r in rows) and any(not (r[2] & 6) for r in rows)
    return out


def bits(ref):
    """SURVEY 8f #3, first slice: bit counts of the reference's entropy coder.  write_coeff into a real stream for seeded
    level blocks of every size / type (levels stored in libthorcuda's pool layout: the low q x q corner, q = min(n, 16)),
    quote_vlc for every table, write_mv for seeded vector pairs."""
    rng = np.random.default_rng(20261106)
    sizes, types, want, levels = [], [], [], []
    for size in (4, 8, 16, 32, 64):
        q = min(size, 16)
        for kind in COEFF_KINDS:
            for rep in range(6 if size <= 16 else 3):
                c = coeff_block(rng, size, kind)
                for typ in range(4):
                    s = ref_stream()
                    ref.write_coeff(C.byref(s), ptr(c), size, typ)
                    sizes.append(size); types.append(typ); want.append(ref_bitpos(s))
                    levels.append(c[:q, :q].reshape(-1).copy())
    ref.quote_vlc.restype = C.c_int
    vlc = [(n, cn, ref.quote_vlc(n, cn)) for n in range(14) for cn in (list(range(3)) if n == 8 else list(range(0, 200)) + [255, 256, 1023, 1024, 40000])]
    mvs = []
    for _ in range(200):
        mv, mvp = rng.integers(-600, 601, 2), rng.integers(-600, 601, 2)
        s = ref_stream()
        a, b = Mv(int(mv[0]), int(mv[1])), Mv(int(mvp[0]), int(mvp[1]))
        ref.write_mv(C.byref(s), C.byref(a), C.byref(b))
        mvs.append((int(mv[0]), int(mv[1]), int(mvp[0]), int(mvp[1]), ref_bitpos(s)))
    # write_block through oracle/bits_shim.c: 1500 seeded blocks; the levels are kept in pool layout (pack_blocks)
    from thorref import BS_FIELDS, load_ref_bits, pack_blocks, random_block_syntax
    shim = load_ref_bits()
    blocks = [random_block_syntax(rng) for _ in range(1500)]
    blk_bits = [shim.ref_write_block_bits(ptr(d), ptr(cy), ptr(cu), ptr(cv)) for d, cy, cu, cv in blocks]
    tb_dtype = np.dtype([("size", "u1"), ("coeff_type", "u1"), ("coeff_offset", "<u4")])
    blk_dtype = np.dtype([("tb_first", "<u4")])
    recs, tbs, pool = pack_blocks(blocks, blk_dtype, tb_dtype)
    blk = {"blk_records": np.stack([b[0] for b in blocks]).astype(np.int32), "blk_bits": np.array(blk_bits, np.uint32),
           "blk_tb_first": recs["tb_first"].copy(), "blk_tb_size": tbs["size"].copy(), "blk_tb_type": tbs["coeff_type"].copy(),
           "blk_tb_offset": tbs["coeff_offset"].copy(), "blk_pool": pool}
    assert blk["blk_records"].shape[1] == BS_FIELDS
    return {**blk, "coeff_size": np.array(sizes, np.uint8), "coeff_type": np.array(types, np.uint8), "coeff_bits": np.array(want, np.uint32),
            "coeff_levels": np.concatenate(levels).astype(np.int16), "vlc": np.array(vlc, np.int32), "mv": np.array(mvs, np.int32)}


def main():
    ref = load_ref()
    if ref is None:
        raise SystemExit("oracle/_ref/libthorref.so missing: run `make -C oracle ref` where /root/reference is mounted")
    if "--streams-only" in sys.argv:
        json.dump(streams(), open(os.path.join(HERE, "streams.json"), "w"), indent=1, sort_keys=True)
        print("wrote", os.path.join(HERE, "streams.json"))
        return
    np.savez_compressed(os.path.join(HERE, "bits.npz"), **bits(ref))
    if "--bits-only" in sys.argv:
        print("wrote", os.path.join(HERE, "bits.npz"), os.path.getsize(os.path.join(HERE, "bits.npz")), "bytes")
        return
    set_ref_simd(ref, 1)
    e1 = extras(ref)
    set_ref_simd(ref, 0)
    e0 = extras(ref)
    for k in e1:
        assert np.array_equal(e1[k], e0[k]), f"reference SIMD and plain-C paths disagree on {k}"
    np.savez_compressed(os.path.join(HERE, "extras.npz"), **e1)
    if "--extras-only" in sys.argv:
        print("wrote", os.path.join(HERE, "extras.npz"), os.path.getsize(os.path.join(HERE, "extras.npz")), "bytes")
        return
    set_ref_simd(ref, 1)
    a = kernels(ref)
    set_ref_simd(ref, 0)
    b = kernels(ref)
    for k in a:
        assert np.array_equal(a[k], b[k]), f"reference SIMD and plain-C paths disagree on {k}"
    set_ref_simd(ref, 1)
    np.savez_compressed(os.path.join(HERE, "kernels.npz"), **a)
    json.dump(streams(), open(os.path.join(HERE, "streams.json"), "w"), indent=1, sort_keys=True)
    json.dump(coding_orders(), open(os.path.join(HERE, "coding_order.json"), "w"), sort_keys=True)
    print("wrote", os.path.join(HERE, "kernels.npz"), os.path.getsize(os.path.join(HERE, "kernels.npz")), "bytes;",
          os.path.join(HERE, "streams.json"))


if __name__ == "__main__":
    main()
