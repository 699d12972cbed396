"""Host-side mirror of the reference's operators for the hot path, same names and argument meaning,
executing on the B200 through libthorcuda (no CPU fallback).

  deblock_frame_y / deblock_frame_uv   reference common/common_frame.c:46-321
  clpf_frame                           reference common/common_frame.c:485-557
  create_reference_frame, pad_yuv_frame reference common/common_frame.c:405-483
  loopfilter_frame(s)                  the fused frame tail of enc/encode_frame.c:150-177 /
                                       dec/decode_frame.c:124-147
  predict_blocks                       get_inter_prediction_luma/chroma + bi-pred average
                                       (common/inter_prediction.c:72-180, enc/encode_block.c:1534-1588)
  encode_blocks / reconstruct_blocks   encode_and_reconstruct_block_inter (enc/encode_block.c:1469-1532),
                                       decode_and_reconstruct_block_inter (dec/decode_block.c:90-120)
  intra_predict_blocks                 make_top_and_left + get_intra_prediction (common/intra_prediction.c:57-389)
  interpolate_frames                   interpolate_frames (common/temporal_interp.c:972-1053)
  coeff_bits / vlc_len / mv_bits       write_coeff, quote_vlc, write_mv as bit counts (enc/write_bits.c, enc/putvlc.c)
  early_skip_thresholds / early_skip_batch  check_early_skip_sub_block(C) (enc/encode_block.c:2507-2610)
  sad_table / cost_batch               sad_calc, widesad_calc, sad_calc_fasthalf/fastquarter, ssd_calc
                                       (enc/encode_block.c:497-797)
"""
import ctypes as C

import numpy as np

from .lib import (LF_CLPF, LF_DEBLOCK_UV, LF_DEBLOCK_Y, LF_MAKE_REF, LIST_COST, LIST_PRED, LIST_TB, COST_RESULT, COST_TASK,
                  BLOCK_SYNTAX, INTRA_TASK, PRED_TASK, SKIP_TASK, TB_TASK, ThorCudaError, _p, cdll, check, sync)


class LfJob(C.Structure):  # thorcuda_lf_job
    _fields_ = [("rec", C.c_int), ("org", C.c_int), ("ref_out", C.c_int), ("qp", C.c_int), ("stages", C.c_int),
                ("sb_bits_in", C.c_void_p), ("sb_bits_out", C.c_void_p)]


def blockinfo_upload(rec, deblock_data):
    """deblock_data: numpy array of the reference's deblock_data_t (44-byte records), raster order;
    it describes picture `rec`."""
    a = np.ascontiguousarray(deblock_data)
    assert a.dtype.itemsize == 44 and a.size == (rec.width // 4) * (rec.height // 4)
    check(cdll().thorcuda_blockinfo_upload(rec.handle, _p(a)))


def loopfilter_frames(jobs, wait=True):
    """jobs: list of dicts(rec, qp, stages, org=None, ref_out=None, sb_bits_in=None).  One launch for all
    frames.  Returns the per-SB CLPF bits of every frame (0xFF: not signalled)."""
    arr = (LfJob * len(jobs))()
    outs, keep = [], []
    for i, j in enumerate(jobs):
        rec = j["rec"]
        nsb = (rec.width // 64) * (rec.height // 64)
        out = np.full(max(nsb, 1), 0xFF, np.uint8)
        bin_ = None if j.get("sb_bits_in") is None else np.ascontiguousarray(j["sb_bits_in"], np.uint8)
        keep.append(bin_)
        outs.append((out, nsb))
        org, ref = j.get("org"), j.get("ref_out")
        arr[i] = LfJob(rec.handle, -1 if org is None else org.handle, -1 if ref is None else ref.handle, int(j["qp"]),
                       int(j["stages"]), None if bin_ is None else bin_.ctypes.data, out.ctypes.data)
    check(cdll().thorcuda_loopfilter_frames(len(jobs), C.byref(arr)))
    if wait:
        sync()
    return [o[:n] for o, n in outs]


def loopfilter_frame(rec, qp, stages, org=None, ref_out=None, sb_bits_in=None):
    """Fused deblock -> CLPF -> reference creation.  Returns the per-SB CLPF bits (0xFF: not signalled)."""
    return loopfilter_frames([dict(rec=rec, qp=qp, stages=stages, org=org, ref_out=ref_out, sb_bits_in=sb_bits_in)])[0]


def deblock_frame_y(rec, qp):
    loopfilter_frame(rec, qp, LF_DEBLOCK_Y)


def deblock_frame_uv(rec, qp):
    """qp is the LUMA qp; the chroma mapping chroma_qp[qp] (enc/encode_frame.c:152) is applied inside."""
    loopfilter_frame(rec, qp, LF_DEBLOCK_UV)


def clpf_frame(rec, org=None, sb_bits_in=None):
    return loopfilter_frame(rec, 0, LF_CLPF, org=org, sb_bits_in=sb_bits_in)


def create_reference_frame(ref, rec):
    loopfilter_frame(rec, 0, LF_MAKE_REF, ref_out=ref)


def pad_yuv_frame(pic):
    check(cdll().thorcuda_pad_frame(pic.handle))
    sync()


def _handles(pics):
    return (C.c_int * len(pics))(*[p.handle for p in pics])


class TaskList:
    """A task list resident in HBM (thorcuda_list_create)."""
    _dtypes = {LIST_PRED: PRED_TASK, LIST_TB: TB_TASK, LIST_COST: COST_TASK}

    def __init__(self, kind, pics, tasks):
        t = np.ascontiguousarray(tasks, self._dtypes[kind])
        self.kind, self.ntasks = kind, t.size
        self.handle = cdll().thorcuda_list_create(kind, _handles(pics), len(pics), _p(t), t.size)
        if self.handle < 0:
            raise ThorCudaError(cdll().thorcuda_last_error().decode())
        self.pool_entries = cdll().thorcuda_list_pool_entries(self.handle) if kind == LIST_TB else 0

    def close(self):
        if self.handle >= 0:
            cdll().thorcuda_list_destroy(self.handle)
            self.handle = -1

    # -- execution -----------------------------------------------------------------------------
    def predict(self, bipred_filter, planes=3):
        check(cdll().thorcuda_predict_list(self.handle, int(bipred_filter), int(planes)))

    def encode(self, coeffq=None, cbp=None, ssd=None):
        check(cdll().thorcuda_encode_list(self.handle, _p(coeffq), _p(cbp), _p(ssd)))

    def reconstruct(self, coeffq=None):
        check(cdll().thorcuda_reconstruct_list(self.handle, _p(coeffq)))

    def coeff_bits(self, bits=None, coeffq=None):
        check(cdll().thorcuda_coeff_bits_list(self.handle, _p(coeffq), _p(bits)))

    def cost(self, kind, results=None):
        check(cdll().thorcuda_cost_list(self.handle, int(kind), _p(results)))


def predict_blocks(pics, tasks, bipred_filter, planes=3):
    """tasks: numpy array of PRED_TASK whose dst/ref0/ref1 index into pics."""
    t = np.ascontiguousarray(tasks, PRED_TASK)
    check(cdll().thorcuda_predict_blocks(_handles(pics), len(pics), _p(t), t.size, int(bipred_filter), int(planes)))
    sync()


def pool_layout(tasks):
    """Assign coeff_offset to every TB task (q*q entries each, 8-entry aligned).  Returns total entries."""
    off = 0
    for t in tasks:
        q = min(int(t["size"]), 16)
        t["coeff_offset"] = off
        off += (q * q + 7) & ~7
    return off


def encode_blocks(pics, tasks, want_ssd=True):
    """Batched encode_and_reconstruct_block_inter.  Returns (coeffq pool int16, cbp uint8, ssd uint32)."""
    lst = TaskList(LIST_TB, pics, tasks)
    coeffq = np.zeros(lst.pool_entries, np.int16)
    cbp = np.zeros(lst.ntasks, np.uint8)
    ssd = np.zeros(lst.ntasks, np.uint32) if want_ssd else None
    lst.encode(coeffq, cbp, ssd)
    sync()
    lst.close()
    return coeffq, cbp, ssd


def reconstruct_blocks(pics, tasks, coeffq):
    """Batched decode_and_reconstruct_block_inter."""
    lst = TaskList(LIST_TB, pics, tasks)
    cq = np.zeros(lst.pool_entries, np.int16)
    cq[:min(cq.size, coeffq.size)] = coeffq[:cq.size]
    lst.reconstruct(cq)
    sync()
    lst.close()


def coeff_bits(pics, tasks, coeffq):
    """write_coeff (enc/write_bits.c:110-253) bit counts of the blocks of a TB task list whose levels are `coeffq`
    (pool layout: q*q int16 per task at coeff_offset, q = min(size, 16)).  Returns uint32 per task."""
    lst = TaskList(LIST_TB, pics, tasks)
    cq = np.zeros(lst.pool_entries, np.int16)
    cq[:min(cq.size, coeffq.size)] = coeffq[:cq.size]
    bits = np.zeros(lst.ntasks, np.uint32)
    lst.coeff_bits(bits, cq)
    sync()
    lst.close()
    return bits


def block_bits(blocks, tb_list=None):
    """write_block (enc/write_bits.c:364-650) as a length for an array of BLOCK_SYNTAX records; tb_list: a TaskList
    (LIST_TB) on which coeff_bits() has run, whose tasks tb_first indexes, or None for the syntax bits alone."""
    b = np.ascontiguousarray(blocks, BLOCK_SYNTAX)
    bits = np.zeros(b.size, np.uint32)
    check(cdll().thorcuda_block_bits(tb_list.handle if tb_list is not None else -1, _p(b), b.size, _p(bits)))
    sync()
    return bits


def block_bits_static(blocks):
    """thorcuda_block_bits_static: the syntax bits of BLOCK_SYNTAX records through the entry point that allocates nothing
    (scratch in a static device array, at most lib BLOCK_BITS_STATIC_MAX records per call); synchronous"""
    b = np.ascontiguousarray(blocks, BLOCK_SYNTAX)
    bits = np.zeros(b.size, np.uint32)
    check(cdll().thorcuda_block_bits_static(_p(b), b.size, _p(bits)))
    return bits


def vlc_len(n, cn):
    """quote_vlc (enc/putvlc.c:133-229)"""
    return int(cdll().thorcuda_vlc_len(int(n), int(cn)))


def mv_bits(mv, mvp):
    """write_mv / quote_mv_bits: bits of the vector difference, mv and mvp as (x, y)"""
    return int(cdll().thorcuda_mv_bits(int(mv[0]), int(mv[1]), int(mvp[0]), int(mvp[1])))


def cost_batch(kind, pics, tasks):
    """Batched sad_calc / widesad_calc / sad_calc_fasthalf / sad_calc_fastquarter / ssd_calc.
    Returns a COST_RESULT array."""
    t = np.ascontiguousarray(tasks, COST_TASK)
    out = np.zeros(t.size, COST_RESULT)
    check(cdll().thorcuda_cost_batch(int(kind), _handles(pics), len(pics), _p(t), t.size, _p(out)))
    sync()
    return out


def intra_predict_blocks(pics, tasks):
    """Intra prediction of independent blocks (one wavefront level): neighbours from pics[src], prediction into pics[dst]."""
    t = np.ascontiguousarray(tasks, INTRA_TASK)
    check(cdll().thorcuda_intra_predict_blocks(_handles(pics), len(pics), _p(t), t.size))


def intra_reconstruct_levels(pics, intra_tasks, intra_first, tb_tasks, tb_first, coeffq):
    """A frame's intra blocks in one call: per wavefront level the predictions, then the coded transform blocks in place
    (thorcuda_intra_reconstruct_levels).  intra_first / tb_first: nlevels + 1 offsets into the level-ordered task arrays."""
    it = np.ascontiguousarray(intra_tasks, INTRA_TASK)
    tt = np.ascontiguousarray(tb_tasks, TB_TASK)
    fi, ft = np.ascontiguousarray(intra_first, np.int32), np.ascontiguousarray(tb_first, np.int32)
    cq = np.ascontiguousarray(coeffq, np.int16)
    check(cdll().thorcuda_intra_reconstruct_levels(_handles(pics), len(pics), _p(it), _p(fi), _p(tt), _p(ft), fi.size - 1, _p(cq)))
    sync()


def interpolate_frames(new_frame, ref0, ref1, ratio, pos):
    """interpolate_frames (common/temporal_interp.c:972): the picture at `pos` of `ratio` between two padded references
    into new_frame's visible area (pad it afterwards like decode_frame.c:107).  Queued; no sync."""
    check(cdll().thorcuda_interpolate_frames(new_frame.handle, ref0.handle, ref1.handle, int(ratio), int(pos)))


def interp_field():
    """the level-0 motion field of the last interpolate_frames call: int16 [bh, bw, 2] (x, y; 1/8 pel, into the farther picture)"""
    bw, bh = C.c_int(), C.c_int()
    check(cdll().thorcuda_interp_field(None, C.byref(bw), C.byref(bh)))
    mv = np.zeros((bh.value, bw.value, 2), np.int16)
    check(cdll().thorcuda_interp_field(_p(mv), None, None))
    return mv


def clip_mv(mv, ypos, xpos, fwidth, fheight, size, sign):
    """clip_mv (enc/encode_block.c:816-828) on (x, y) quarter-pel; returns the clipped (x, y).  Host only."""
    m = (C.c_int16 * 2)(int(mv[0]), int(mv[1]))
    cdll().thorcuda_clip_mv(m, ypos, xpos, fwidth, fheight, size, int(sign))
    return m[0], m[1]


def sad_table_index(fwidth, bs, xpos, ypos, mv, sign=0, centre=(0, 0), search_range=16):
    """Index of motion_estimate's integer-pel probe at mv inside one table level, or -1 (outside the window)."""
    m = (C.c_int16 * 2)(int(mv[0]), int(mv[1]))
    return int(cdll().thorcuda_sad_table_index(fwidth, bs, xpos, ypos, m, int(sign), int(centre[0]), int(centre[1]), search_range))


def early_skip_thresholds(qp, size, early_skip_thr):
    """(thr_y, thr_c) of check_early_skip_sub_block / _sub_blockC for a size x size luma sub-block at frame qp."""
    ty, tc = C.c_int(), C.c_int()
    cdll().thorcuda_early_skip_thresholds(int(qp), int(size), float(early_skip_thr), C.byref(ty), C.byref(tc))
    return ty.value, tc.value


def early_skip_batch(pics, tasks):
    """Significance flags (bit0 Y, bit1 U, bit2 V) of each sub-block; 0 = the skip candidate stands."""
    t = np.ascontiguousarray(tasks, SKIP_TASK)
    out = np.zeros(t.size, np.uint8)
    check(cdll().thorcuda_early_skip_batch(_handles(pics), len(pics), _p(t), t.size, _p(out)))
    sync()
    return out


def sad_table_build(slots, orgs, refs, search_range, centres=None):
    """Queue the table build for n (org, ref) pairs (one launch)."""
    n = len(slots)
    sl = (C.c_int * n)(*slots)
    cptr, keep = None, []
    if centres is not None:
        arr = (C.c_void_p * n)()
        for i, c in enumerate(centres):
            if c is None:
                arr[i] = None
            else:
                a = np.ascontiguousarray(c, np.int16)
                keep.append(a)
                arr[i] = a.ctypes.data
        cptr = C.cast(arr, C.c_void_p)
    check(cdll().thorcuda_sad_table_build(n, sl, _handles(orgs), _handles(refs), int(search_range), cptr))


def sad_table_fetch(slot, width, height, search_range, sizes=(8, 16, 32, 64)):
    """dict size -> ndarray [blocks_y, blocks_x, 2R, 2R] (displacements -R..R-1; uint16 for 8/16, uint32 for 32/64)."""
    out = {}
    n = 2 * search_range
    for bs in sizes:
        nbytes = cdll().thorcuda_sad_table_bytes(slot, bs)
        dt = np.uint16 if bs <= 16 else np.uint32
        a = np.empty(nbytes // np.dtype(dt).itemsize, dt)
        if nbytes:
            check(cdll().thorcuda_sad_table_fetch(slot, bs, _p(a), nbytes))
        sync()
        out[bs] = a.reshape(height // bs, width // bs, n, n)
    return out


def sad_table(org, ref, search_range, centres=None, slot=0):
    sad_table_build([slot], [org], [ref], search_range, None if centres is None else [centres])
    return sad_table_fetch(slot, org.width, org.height, search_range)
