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


# ---------------------------------------------------------------------------------------------------------------------
# The product side: run one stream as closed-GOP segments in separate encoder processes and join what they wrote.
# The encoder is integration/_build/Thorenc_gpu (or Thorenc_seg / Thorenc_emu in the CPU tests): the reference's main()
# with -seg_start / -seg_end (integration/enc_segment_hook.c).  No pixel is touched here.

def segment_ranges(num_frames, intra_period, parts):
    """Cut frames [0, num_frames) at intra frames into `parts` contiguous ranges [(a, b), ...]; b is None for the last
    range ("to the end").  Ranges hold whole intra periods and differ by at most one; fewer ranges come back when the
    stream has fewer intra periods than `parts`.  One process per range: a range of several periods re-encodes no
    intra frame inside it, only the one that ends it."""
    if intra_period <= 0:
        raise ValueError("only configurations with -intra_period have closed-GOP segments (SURVEY 0, fact 2)")
    periods = max(1, (num_frames + intra_period - 1) // intra_period)      # I-frames at 0, P, 2P, ... < num_frames
    parts = max(1, min(parts, periods))
    cuts = [(periods * k) // parts * intra_period for k in range(parts + 1)]
    return [(cuts[k], cuts[k + 1] if k + 1 < parts else None) for k in range(parts)]


def split_chunks(bitstream):
    """A .bit file as the list of its frame chunks, each WITH its 4-byte big-endian length prefix (flush_all_bits,
    enc/putbits.c:57-95; the first chunk of a stream also carries the sequence header)."""
    out, pos = [], 0
    while pos < len(bitstream):
        if pos + 4 > len(bitstream):
            raise ValueError("truncated frame length")
        n = int.from_bytes(bitstream[pos:pos + 4], "big")
        if pos + 4 + n > len(bitstream):
            raise ValueError("truncated frame chunk")
        out.append(bitstream[pos:pos + 4 + n])
        pos += 4 + n
    return out


def join_segment_streams(ranges, streams, num_frames, num_reorder_pics, intra_period, dyadic=True):
    """ranges[k] = (a, b) and streams[k] = the bytes process k wrote.  Returns the serial stream: every chunk goes to
    the position its frame has in the serial coding order (the leading B-frames of an intra frame are coded AFTER it
    but belong to the segment BEFORE it, so plain concatenation of the files is not the serial order)."""
    order = coding_order(num_frames, num_reorder_pics, intra_period, dyadic)
    slot = {}
    for (a, b), data in zip(ranges, streams):
        mine = [f for f in order if f.frame_num >= a and (b is None or f.frame_num < b)]
        chunks = split_chunks(data)
        if len(chunks) != len(mine):
            raise ValueError(f"segment [{a}, {b}) wrote {len(chunks)} frames, the coding order has {len(mine)}")
        for f, c in zip(mine, chunks):
            slot[f.num_encoded_before] = c
    if sorted(slot) != list(range(len(order))):
        raise ValueError("the segments do not cover the stream exactly once")
    return b"".join(slot[i] for i in range(len(order)))


def segment_args(a, b):
    return ["-seg_start", str(a)] + (["-seg_end", str(b)] if b is not None else [])


def encode_segments(exe, args, ranges, out_prefix, env=None, envs=None, max_parallel=None, timeout=3600, rec=True):
    """Start one encoder process per range (at most max_parallel at a time): `exe args -of <prefix>_k.bit -rf
    <prefix>_k_rec.yuv -seg_start a [-seg_end b]`.  envs[k] overrides env for process k (e.g. CUDA_VISIBLE_DEVICES);
    rec=False leaves the reconstruction files out.  Returns [(bit_path, rec_path)].  Raises if any process fails --
    there is no fallback."""
    import os
    import subprocess
    import time
    files = [(f"{out_prefix}_{k}.bit", f"{out_prefix}_{k}_rec.yuv") for k in range(len(ranges))]
    pending = list(range(len(ranges)))
    running = {}
    limit = max_parallel or len(ranges)
    t_end = time.monotonic() + timeout
    try:
        while pending or running:
            while pending and len(running) < limit:
                k = pending.pop(0)
                a, b = ranges[k]
                cmd = [exe] + list(args) + ["-of", files[k][0]] + (["-rf", files[k][1]] if rec else []) + segment_args(a, b)
                e = dict(os.environ if env is None else env)
                if envs is not None:
                    e.update(envs[k])
                log = open(f"{out_prefix}_{k}.log", "w")
                running[k] = (subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT, env=e), log)
            for k in list(running):
                p, log = running[k]
                rc = p.poll()
                if rc is None:
                    continue
                log.close()
                del running[k]
                if rc != 0:
                    raise RuntimeError(f"segment {ranges[k]} failed ({rc}): " + open(f"{out_prefix}_{k}.log").read()[-1500:])
            if time.monotonic() > t_end:
                raise TimeoutError("segment encoders did not finish")
            time.sleep(0.02)
    finally:
        for p, log in running.values():
            p.kill()
            p.wait()
            log.close()
    return files
