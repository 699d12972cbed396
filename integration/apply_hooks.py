#!/usr/bin/env python
"""Build-time substitution of the frame-tail hooks into a SCRATCH copy of the reference's dec/decode_frame.c
(INTEGRATION.md 2.2).  Nothing of the reference is stored in this repository: the script names the three call
sites it replaces and fails loudly if the reference no longer has them.

usage: apply_hooks.py <reference>/dec/decode_frame.c <out.c>"""
import sys

src = open(sys.argv[1]).read()


def once(text, needle):
    if text.count(needle) != 1:
        raise SystemExit(f"apply_hooks: expected exactly one {needle!r} in {sys.argv[1]}")
    return text.index(needle)


# 1. prototypes after the last include
inc = '#include "temporal_interp.h"\n'
i = once(src, inc) + len(inc)
src = (src[:i] + "void thorcuda_dec_tail(decoder_info_t *decoder_info, stream_t *stream, int qp);\n"
       "void thorcuda_dec_make_reference(yuv_frame_t *ref, yuv_frame_t *rec);\n" + src[i:])
# 2. deblock_frame_y / deblock_frame_uv / clpf_frame  ->  one fused call
a = once(src, "if (decoder_info->deblocking){")
b = once(src, "/* Sliding window operation")
for fn in ("deblock_frame_y(", "deblock_frame_uv(", "clpf_frame("):
    if fn not in src[a:b]:
        raise SystemExit(f"apply_hooks: {fn} not where expected")
src = src[:a] + "thorcuda_dec_tail(decoder_info, stream, qp);\n\n  " + src[b:]
# 3. create_reference_frame  ->  fetch the reference the same launch produced
c = "create_reference_frame(decoder_info->ref[0],decoder_info->rec);"
i = once(src, c)
src = src[:i] + "thorcuda_dec_make_reference(decoder_info->ref[0],decoder_info->rec);" + src[i + len(c):]
open(sys.argv[2], "w").write(src)
