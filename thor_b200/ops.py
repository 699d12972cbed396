"""Host-side mirror of the reference's frame-level operators for the hot path, same names and
argument meaning, executing on the B200 through libthorcuda (no CPU fallback).

  deblock_frame_y / deblock_frame_uv   reference common/common_frame.c:46-321
  clpf_frame                           reference common/common_frame.c:485-557
  create_reference_frame, pad_yuv_frame reference common/common_frame.c:405-483
  loopfilter_frame                     the fused frame tail of enc/encode_frame.c:150-177 /
                                       dec/decode_frame.c:124-147
"""
import ctypes as C

import numpy as np

from .lib import (LF_CLPF, LF_DEBLOCK_UV, LF_DEBLOCK_Y, LF_MAKE_REF, COST_RESULT, COST_TASK, PRED_TASK, TB_TASK, _p, cdll,
                  check, sync)


def blockinfo_upload(deblock_data, width, height):
    """deblock_data: numpy array of the reference's deblock_data_t (44-byte records), raster order."""
    a = np.ascontiguousarray(deblock_data)
    assert a.dtype.itemsize == 44 and a.size == (width // 4) * (height // 4)
    check(cdll().thorcuda_blockinfo_upload(_p(a), width, height))
    sync()


def loopfilter_frame(rec, qp, stages, org=None, ref_out=None, sb_bits_in=None):
    """Fused deblock -> CLPF -> reference creation.  Returns the per-SB CLPF bits (0xFF: not signalled)."""
    nsb = (rec.width // 64) * (rec.height // 64)
    bits_out = np.full(max(nsb, 1), 0xFF, np.uint8)
    bin_ = None if sb_bits_in is None else np.ascontiguousarray(sb_bits_in, np.uint8)
    check(cdll().thorcuda_loopfilter_frame(rec.handle, -1 if org is None else org.handle,
                                           -1 if ref_out is None else ref_out.handle, int(qp), int(stages),
                                           _p(bin_), _p(bits_out)))
    sync()
    return bits_out[:nsb]


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


def predict_blocks(dst, refs, tasks, bipred_filter, planes=3):
    """tasks: numpy array of PRED_TASK.  get_inter_prediction_luma/chroma (+ bi-pred average) for a
    whole block list (reference common/inter_prediction.c:72-180, enc/encode_block.c:1534-1588)."""
    t = np.ascontiguousarray(tasks, PRED_TASK)
    check(cdll().thorcuda_predict_blocks(dst.handle, _handles(refs), len(refs), _p(t), t.size, int(bipred_filter), int(planes)))
    sync()


def encode_blocks(org, pred, rec, tasks, pool_entries, want_ssd=True):
    """Batched encode_and_reconstruct_block_inter (reference enc/encode_block.c:1469-1532).
    Returns (coeffq pool int16, cbp flags uint8, ssd uint32 or None)."""
    t = np.ascontiguousarray(tasks, TB_TASK)
    coeffq = np.zeros(pool_entries, np.int16)
    cbp = np.zeros(t.size, np.uint8)
    ssd = np.zeros(t.size, np.uint32) if want_ssd else None
    check(cdll().thorcuda_encode_blocks(org.handle, pred.handle, rec.handle, _p(t), t.size, pool_entries, _p(coeffq),
                                        _p(cbp), _p(ssd)))
    sync()
    return coeffq, cbp, ssd


def reconstruct_blocks(pred, rec, tasks, coeffq):
    """Batched decode_and_reconstruct_block_inter (reference dec/decode_block.c:90-120)."""
    t = np.ascontiguousarray(tasks, TB_TASK)
    cq = np.ascontiguousarray(coeffq, np.int16)
    check(cdll().thorcuda_reconstruct_blocks(pred.handle, rec.handle, _p(t), t.size, _p(cq), cq.size))
    sync()


def cost_batch(kind, org, refs, tasks):
    """Batched sad_calc / widesad_calc / sad_calc_fasthalf / sad_calc_fastquarter / ssd_calc
    (reference enc/encode_block.c:497-797).  Returns a COST_RESULT array."""
    t = np.ascontiguousarray(tasks, COST_TASK)
    out = np.zeros(t.size, COST_RESULT)
    check(cdll().thorcuda_cost_batch(int(kind), org.handle, _handles(refs), len(refs), _p(t), t.size, _p(out)))
    sync()
    return out


def sad_table(org, ref, search_range, centres=None):
    """Whole-frame integer-pel SAD tables for block sizes 8,16,32,64 (dict size -> ndarray
    [blocks_y, blocks_x, 2R+1, 2R+1])."""
    c = None if centres is None else np.ascontiguousarray(centres, np.int16)
    check(cdll().thorcuda_sad_table_build(org.handle, ref.handle, int(search_range), _p(c)))
    out = {}
    n = 2 * search_range + 1
    for bs in (8, 16, 32, 64):
        nbytes = cdll().thorcuda_sad_table_bytes(bs)
        dt = np.uint16 if bs <= 16 else np.uint32
        a = np.empty(nbytes // np.dtype(dt).itemsize, dt)
        check(cdll().thorcuda_sad_table_fetch(bs, _p(a), nbytes))
        sync()
        by, bx = (org.height - bs) // 8 + 1, (org.width - bs) // 8 + 1
        out[bs] = a.reshape(by, bx, n, n)
    return out
