"""Multi-GPU partitioning of the Thor pixel pipeline (host logic only).

Ranks never exchange pixels: the units of parallelism are independent *streams* (every config) and, for the RA*
configs (`-intra_period 64`), *closed-GOP segments* of one stream.  This module restates the part of the
reference's sequence loop that decides WHICH frames are coded in WHICH order (enc/mainenc.c:222-591) so that a
segment can be started in the middle of a stream with the state the serial run would have there, and gives the
join rule (ordered concatenation of per-frame chunks, enc/putbits.c:57-95).  Nothing here touches pixels.
"""
from dataclasses import dataclass

# coding order -> display order inside one dyadic sub-GOP (enc/mainenc.c:47-53)
_CODE_TO_DISPLAY = {1: [0], 2: [1, 0], 4: [3, 1, 0, 2], 8: [7, 3, 1, 5, 0, 2, 4, 6],
                    16: [15, 7, 3, 11, 1, 5, 9, 13, 0, 2, 4, 6, 8, 10, 12, 14]}


@dataclass(frozen=True)
class CodedFrame:
    frame_num: int            # display number (frame_info.frame_num)
    frame_type: str           # "I", "P" or "B"
    num_encoded_before: int   # num_encoded_frames when the frame is coded (drives num_ref, b_level: mainenc.c:250-289)
    last_intra_frame_num: int  # value used by the reference-stripping rule (mainenc.c:472-486)


def reorder_frame_offset(idx, sub_gop, dyadic=True):
    """enc/mainenc.c:63-71"""
    if dyadic and sub_gop > 1:
        return _CODE_TO_DISPLAY[sub_gop][idx] - sub_gop + 1
    return 0 if idx == 0 else idx - sub_gop


def coding_order(num_frames, num_reorder_pics=0, intra_period=0, dyadic=True):
    """Frames in the order the reference codes them (enc/mainenc.c:222-250, tail rule :586-590), `-skip 0`."""
    sub_gop = max(1, num_reorder_pics + 1)
    reorder = num_reorder_pics
    out = []
    encoded = 0
    last_intra = 0
    f0 = 0
    while f0 < num_frames:
        for k in range(sub_gop):
            fn = f0 + reorder_frame_offset(k, sub_gop, dyadic)
            if fn < 0:
                continue
            if reorder == 0:
                is_i = (encoded % intra_period == 0) if intra_period > 0 else encoded == 0
                ft = "I" if is_i else "P"
            else:
                is_i = (fn % intra_period == 0) if intra_period > 0 else fn == 0
                ft = "I" if is_i else ("P" if fn % sub_gop == 0 else "B")
            if ft == "I":
                last_intra = fn
            out.append(CodedFrame(fn, ft, encoded, last_intra))
            encoded += 1
        # the reference drops to one-frame sub-GOPs when a whole sub-GOP no longer fits (mainenc.c:586-590);
        # its for-loop then advances by the NEW sub_gop
        if f0 + sub_gop >= num_frames and sub_gop >= 2:
            sub_gop, reorder = 1, 0
        f0 += sub_gop
    return out


@dataclass(frozen=True)
class Segment:
    index: int
    first_intra: int          # display number of the I-frame that opens the segment (emitted by this segment)
    emit: tuple               # CodedFrames this segment emits, in coding order
    redundant: tuple          # CodedFrames it must ALSO encode but not emit (the next segment's opening I-frame)
    header: bool              # emits the sequence header (segment 0 only, mainenc.c:196-212)
    start_num_encoded: int    # num_encoded_frames to start from so that num_ref / coded_phase match the serial run


def ra_segments(num_frames, num_reorder_pics=7, intra_period=64, dyadic=True):
    """Closed-GOP segments of an RA stream.  Segment k emits I(k*P) and every frame coded between it and the next
    I-frame.  The leading B-frames of the next I-frame (display order before it, coded after it) belong to segment k
    and reference that I-frame, so segment k re-encodes it from the source (same qp: it depends only on the
    source and qp + dqpI, mainenc.c:254-256) without emitting it.  Frames after an I-frame never reference
    pictures before it (mainenc.c:472-486), so a segment needs nothing from its predecessor."""
    if intra_period <= 0:
        raise ValueError("only configurations with -intra_period have closed-GOP segments (SURVEY 0, fact 2)")
    order = coding_order(num_frames, num_reorder_pics, intra_period, dyadic)
    intra_pos = [i for i, f in enumerate(order) if f.frame_type == "I"]
    segs = []
    for k, p in enumerate(intra_pos):
        nxt = intra_pos[k + 1] if k + 1 < len(intra_pos) else None
        lo, hi = order[p].frame_num, (order[nxt].frame_num if nxt is not None else None)
        emit = [order[p]]
        for f in order[p + 1:]:
            if f.frame_type == "I":
                continue
            if f.frame_num > lo and (hi is None or f.frame_num < hi):
                emit.append(f)
        # the next I-frame is needed iff some emitted frame is coded after it (its leading B-frames)
        red = ()
        if nxt is not None and any(f.num_encoded_before > order[nxt].num_encoded_before for f in emit):
            red = (order[nxt],)
        segs.append(Segment(k, lo, tuple(emit), red, k == 0, order[p].num_encoded_before))
    return segs


def shard(items, rank, world):
    """Contiguous, balanced shards (segment k and k+1 share an I-frame: neighbours on one rank re-encode less)."""
    n = len(items)
    lo, hi = (n * rank) // world, (n * (rank + 1)) // world
    return list(items[lo:hi])


def join(per_rank_chunks):
    """per_rank_chunks: iterable of lists of (num_encoded_before, bytes).  The stream is the concatenation of the
    per-frame chunks in coding order (every frame ends byte-aligned: flush_all_bits per frame, mainenc.c:560)."""
    allc = sorted((c for chunks in per_rank_chunks for c in chunks), key=lambda c: c[0])
    keys = [c[0] for c in allc]
    if len(set(keys)) != len(keys):
        raise ValueError("a frame was emitted twice")
    return b"".join(c[1] for c in allc)
