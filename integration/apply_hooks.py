#!/usr/bin/env python
"""Build-time substitution of the frame-tail hooks into a SCRATCH copy of the reference's dec/decode_frame.c
(INTEGRATION.md 2.2).  Nothing of the reference is stored in this repository: the script names the three call
sites it replaces and fails loudly if the reference no longer has them.

usage: apply_hooks.py tail  <reference>/dec/decode_frame.c <out.c>     frame tail only
       apply_hooks.py recon <reference>/dec/decode_frame.c <out.c>     + device-resident reconstruction (2.3)
       apply_hooks.py block <reference>/dec/decode_block.c <out.c>     the block-level call sites of 2.3
       apply_hooks.py enc   <reference>/enc/encode_frame.c <out.c>     frame tail of the encoder (2.2)
       apply_hooks.py main  <reference>/dec/maindec.c <out.c>          output writes of the pipelined decoder (2.3)
       apply_hooks.py encgpu <reference>/enc/encode_frame.c <out.c>    the whole frame encoder on the device (2.7)
       apply_hooks.py encemu <reference>/enc/encode_frame.c <out.c>    superblock loop only (host emulation, tests)
       apply_hooks.py mainenc <reference>/enc/mainenc.c <out.c>        interpolated references of the encoder (2.7) + segments (2.8)
       apply_hooks.py mainseg <reference>/enc/mainenc.c <out.c>        closed-GOP segment encoding only (2.8)"""
import re
import sys

mode = sys.argv[1]
sys.argv = [sys.argv[0]] + sys.argv[2:]
src = open(sys.argv[1]).read()

if mode == "block":
    # prediction and reconstruction calls of decode_block() become task recorders with the same arguments
    n_l = len(re.findall(r"get_inter_prediction_luma\s*\(", src))
    n_c = len(re.findall(r"get_inter_prediction_chroma\s*\(", src))
    n_r = len(re.findall(r"decode_and_reconstruct_block_inter\(rec_", src))
    n_i = len(re.findall(r"decode_and_reconstruct_block_intra\(rec_", src))
    if (n_l, n_c, n_r, n_i) != (9, 18, 3, 3):
        raise SystemExit(f"apply_hooks: decode_block.c call sites changed: {(n_l, n_c, n_r, n_i)}")
    src = re.sub(r"get_inter_prediction_luma\s*\(", "hk_luma(", src)
    src = re.sub(r"get_inter_prediction_chroma\s*\(", "hk_chroma(", src)
    src = src.replace("decode_and_reconstruct_block_inter(rec_", "hk_recon_inter(rec_")
    src = src.replace("decode_and_reconstruct_block_intra(rec_", "hk_recon_intra(rec_")
    proto = ("void hk_luma(uint8_t *pblock, uint8_t *ref, int width, int height, int stride, int pstride, mv_t *mv, int sign, int bipred);\n"
             "void hk_chroma(uint8_t *pblock, uint8_t *ref, int width, int height, int stride, int pstride, mv_t *mv, int sign);\n"
             "void hk_recon_inter(uint8_t *rec, int stride, int size, int qp, uint8_t *pblock, int16_t *coeffq, int tb_split);\n"
             "void hk_recon_intra(uint8_t *rec, int stride, int size, int qp, uint8_t *pblock, int16_t *coeffq, int tb_split, int upright_available,\n"
             "                    int downleft_available, intra_mode_t intra_mode, int ypos, int xpos, int width, int comp);\n")
    anchor = "extern int chroma_qp[52];\n"
    if src.count(anchor) != 1:
        raise SystemExit("apply_hooks: anchor for the prototypes not found in decode_block.c")
    src = src.replace(anchor, anchor + proto)
    open(sys.argv[2], "w").write(src)
    raise SystemExit(0)

if mode == "main":
    # dec/maindec.c: the two write_yuv_frame calls go through the hook (a picture still in flight on the device is held
    # back), and the queue is drained before the statistics are printed
    call = "write_yuv_frame(&rec[op_rec_buffer_idx],width,height,outfile);"
    if src.count(call) != 2:
        raise SystemExit("apply_hooks: maindec.c no longer has the two write_yuv_frame call sites")
    src = src.replace(call, "thorcuda_dec_write_frame(&rec[op_rec_buffer_idx],width,height,outfile);")
    stats = "bit_count_t bit_count = decoder_info.bit_count;"
    if src.count(stats) != 1:
        raise SystemExit("apply_hooks: anchor after the output tail not found in maindec.c")
    src = src.replace(stats, "thorcuda_dec_finish();\n    " + stats)
    anchor = "int main(int argc, char** argv)"
    if src.count(anchor) != 1:
        raise SystemExit("apply_hooks: main() not found in maindec.c")
    src = src.replace(anchor, "void thorcuda_dec_write_frame(yuv_frame_t *frame, int width, int height, FILE *outfile);\n"
                              "void thorcuda_dec_finish(void);\n" + anchor)
    open(sys.argv[2], "w").write(src)
    raise SystemExit(0)

if mode == "enc":
    # enc/encode_frame.c: the same three call sites on the encoder side
    def once_e(text, needle):
        if text.count(needle) != 1:
            raise SystemExit(f"apply_hooks: expected exactly one {needle!r} in {sys.argv[1]}")
        return text.index(needle)
    inc = "extern int chroma_qp[52];\n"
    i = once_e(src, inc) + len(inc)
    src = (src[:i] + "void thorcuda_enc_tail(encoder_info_t *encoder_info, stream_t *stream, int qp);\n"
           "void thorcuda_enc_make_reference(yuv_frame_t *ref, yuv_frame_t *rec);\n" + src[i:])
    a = once_e(src, "if (encoder_info->params->deblocking){")
    b = once_e(src, "/* Sliding window operation")
    for fn in ("deblock_frame_y(", "deblock_frame_uv(", "clpf_frame(", "putbits(1, !sb_signal, stream);"):
        if fn not in src[a:b]:
            raise SystemExit(f"apply_hooks: {fn} not where expected")
    src = src[:a] + "thorcuda_enc_tail(encoder_info, stream, qp);\n\n  " + src[b:]
    c = "create_reference_frame(encoder_info->ref[0],encoder_info->rec);"
    i = once_e(src, c)
    src = src[:i] + "thorcuda_enc_make_reference(encoder_info->ref[0],encoder_info->rec);" + src[i + len(c):]
    open(sys.argv[2], "w").write(src)
    raise SystemExit(0)


if mode in ("encgpu", "encemu"):
    # enc/encode_frame.c: the superblock loop (:112-148) becomes one call; encgpu also takes the frame tail
    def once_g(text, needle):
        if text.count(needle) != 1:
            raise SystemExit(f"apply_hooks: expected exactly one {needle!r} in {sys.argv[1]}")
        return text.index(needle)
    inc = "extern int chroma_qp[52];\n"
    i = once_g(src, inc) + len(inc)
    src = (src[:i] + "void thorcuda_enc_superblocks(encoder_info_t *encoder_info);\n"
           "void thorcuda_enc_tail(encoder_info_t *encoder_info, stream_t *stream, int qp);\n"
           "void thorcuda_enc_make_reference(yuv_frame_t *ref, yuv_frame_t *rec);\n" + src[i:])
    a = once_g(src, "for (k=0;k<num_sb_ver;k++){")
    b = once_g(src, "if (encoder_info->params->deblocking){")
    body = src[a:b]
    for fn in ("process_block(encoder_info,MAX_BLOCK_SIZE,yposY,xposY,qp);", "frame_info->best_ref = -1;", "write_stream_pos(stream,&stream_pos_ref);"):
        if fn not in body:
            raise SystemExit(f"apply_hooks: {fn} not inside the superblock loop")
    src = src[:a] + "(void)k; (void)l; (void)num_sb_hor; (void)num_sb_ver;\n  thorcuda_enc_superblocks(encoder_info);\n\n  " + src[b:]
    if mode == "encgpu":
        a = once_g(src, "if (encoder_info->params->deblocking){")
        b = once_g(src, "/* Sliding window operation")
        for fn in ("deblock_frame_y(", "deblock_frame_uv(", "clpf_frame(", "putbits(1, !sb_signal, stream);"):
            if fn not in src[a:b]:
                raise SystemExit(f"apply_hooks: {fn} not where expected")
        src = src[:a] + "thorcuda_enc_tail(encoder_info, stream, qp);\n\n  " + src[b:]
        c = "create_reference_frame(encoder_info->ref[0],encoder_info->rec);"
        i = once_g(src, c)
        src = src[:i] + "thorcuda_enc_make_reference(encoder_info->ref[0],encoder_info->rec);" + src[i + len(c):]
    open(sys.argv[2], "w").write(src)
    raise SystemExit(0)

if mode in ("mainenc", "mainseg"):
    # enc/mainenc.c.  Both modes: closed-GOP segment encoding (-seg_start / -seg_end, integration/enc_segment_hook.c);
    # mainenc also puts interpolate_frames + pad_yuv_frame of the interpolated reference (:328-329, :385-386) on the device
    def sub_once(text, old, new):
        if text.count(old) != 1:
            raise SystemExit(f"apply_hooks: expected exactly one '{old}' in mainenc.c, found {text.count(old)}")
        return text.replace(old, new)
    fn = "encoder_info.frame_info.frame_num"
    pat = re.compile(r"interpolate_frames\(encoder_info\.interp_frames\[0\], ref1, ref2, ([^;]*?)\);\s*pad_yuv_frame\(encoder_info\.interp_frames\[0\]\);")
    m = pat.findall(src)
    if len(m) != 2:
        raise SystemExit(f"apply_hooks: expected two interpolate_frames call sites in mainenc.c, found {len(m)}")
    if mode == "mainenc":
        src = pat.sub(lambda mm: f"if (thor_seg_encodes({fn})) thorcuda_enc_interpolate(encoder_info.interp_frames[0], ref1, ref2, " + mm.group(1) + ");", src)
    else:
        src = pat.sub(lambda mm: f"if (thor_seg_encodes({fn})) {{ " + mm.group(0) + " }", src)
    src = sub_once(src, "init_use_simd();", "init_use_simd();\n  thor_seg_args(&argc, argv);")
    src = sub_once(src, "check_parameters(params);", "check_parameters(params);\n  thor_seg_check(params);")
    src = sub_once(src, "last_PorI_frame = -1;", "last_PorI_frame = -1;\n  last_frame_output = thor_seg_first_output();")
    src = sub_once(src, "encode_frame(&encoder_info);\n      rec_available[rec_buffer_idx]=1;",
                   f"if (thor_seg_encodes({fn})) encode_frame(&encoder_info); else thor_seg_skip_frame(&encoder_info);\n"
                   f"      rec_available[rec_buffer_idx]=thor_seg_emits({fn});")
    src = sub_once(src, "flush_all_bits(&stream, strfile);", f"thor_seg_flush(&stream, strfile, {fn});")
    src = sub_once(src, "read_yuv_frame(&orig,width,height,infile);", f"if (thor_seg_encodes({fn})) read_yuv_frame(&orig,width,height,infile);")
    src = sub_once(src, "if (params->snrcalc){", f"if (params->snrcalc && thor_seg_encodes({fn})){{")
    # the reference addresses the input with 32-bit arithmetic (:77, :222, :507): streams with enough segments for eight GPUs
    # pass 2 GiB at 1080p, so the three places go to 64 bits (same values below 2 GiB)
    src = sub_once(src, "uint32_t input_file_size;", "uint64_t input_file_size;")
    src = sub_once(src, "input_file_size = ftell(infile);", "input_file_size = (uint64_t)ftello(infile);")
    src = sub_once(src, "(frame_num0+1)*frame_size <= input_file_size;", "(uint64_t)(frame_num0+1)*frame_size <= input_file_size;")
    src = sub_once(src, "if (((frame_num0+sub_gop+1)*frame_size > input_file_size", "if (((uint64_t)(frame_num0+sub_gop+1)*frame_size > input_file_size")
    src = sub_once(src, "fseek(infile, frame_num*(frame_size+params->frame_headerlen)+params->file_headerlen+params->frame_headerlen, SEEK_SET);",
                   "fseeko(infile, (off_t)frame_num*(frame_size+params->frame_headerlen)+params->file_headerlen+params->frame_headerlen, SEEK_SET);")
    anchor = "int main(int argc, char **argv)"
    if src.count(anchor) != 1:
        anchor = "int main(int argc, char** argv)"
    if src.count(anchor) != 1:
        raise SystemExit("apply_hooks: main() not found in mainenc.c")
    proto = ("void thor_seg_args(int *argc, char **argv);\nvoid thor_seg_check(const enc_params *p);\nint thor_seg_encodes(int frame_num);\n"
             "int thor_seg_emits(int frame_num);\nint thor_seg_first_output(void);\nvoid thor_seg_skip_frame(encoder_info_t *e);\n"
             "void thor_seg_flush(stream_t *str, FILE *out, int frame_num);\n"
             "void thorcuda_enc_interpolate(yuv_frame_t *dst, yuv_frame_t *ref0, yuv_frame_t *ref1, int ratio, int pos);\n")
    src = "#define _FILE_OFFSET_BITS 64\n" + src.replace(anchor, proto + anchor)
    open(sys.argv[2], "w").write(src)
    raise SystemExit(0)

tail_call = "thorcuda_dec_tail(decoder_info, stream, qp);" if mode == "tail" else "thorcuda_dec_flush(decoder_info, stream, qp);"
ref_call = "thorcuda_dec_make_reference" if mode == "tail" else "thorcuda_dec_fetch_reference"


def once(text, needle):
    if text.count(needle) != 1:
        raise SystemExit(f"apply_hooks: expected exactly one {needle!r} in {sys.argv[1]}")
    return text.index(needle)


# 1. prototypes after the last include
inc = '#include "temporal_interp.h"\n'
i = once(src, inc) + len(inc)
src = (src[:i] + "void thorcuda_dec_tail(decoder_info_t *decoder_info, stream_t *stream, int qp);\n"
       "void thorcuda_dec_make_reference(yuv_frame_t *ref, yuv_frame_t *rec);\n"
       "void thorcuda_dec_frame_begin(decoder_info_t *decoder_info);\n"
       "void thorcuda_dec_flush(decoder_info_t *decoder_info, stream_t *stream, int qp);\n"
       "void thorcuda_dec_fetch_reference(yuv_frame_t *ref, yuv_frame_t *rec);\n"
       "void thorcuda_dec_interpolate(yuv_frame_t *dst, yuv_frame_t *ref0, yuv_frame_t *ref1, int ratio, int pos);\n" + src[i:])
if mode == "recon":
    loop = "for (k=0;k<num_sb_ver;k++){"
    i = once(src, loop)
    src = src[:i] + "thorcuda_dec_frame_begin(decoder_info);\n  " + src[i:]
    # interpolated reference: interpolate_frames + pad_yuv_frame on the device twins of the reference ring
    m = re.findall(r"interpolate_frames\(decoder_info->interp_frames\[0\], ref1, ref2, off1\+off2 , off2\);\s*"
                   r"pad_yuv_frame\(decoder_info->interp_frames\[0\]\);", src)
    if len(m) != 1:
        raise SystemExit("apply_hooks: interpolate_frames call site changed in decode_frame.c")
    src = src.replace(m[0], "thorcuda_dec_interpolate(decoder_info->interp_frames[0], ref1, ref2, off1+off2, off2);")
# 2. deblock_frame_y / deblock_frame_uv / clpf_frame  ->  one fused call
a = once(src, "if (decoder_info->deblocking){")
b = once(src, "/* Sliding window operation")
for fn in ("deblock_frame_y(", "deblock_frame_uv(", "clpf_frame("):
    if fn not in src[a:b]:
        raise SystemExit(f"apply_hooks: {fn} not where expected")
src = src[:a] + tail_call + "\n\n  " + src[b:]
# 3. create_reference_frame  ->  fetch the reference the same launch produced
c = "create_reference_frame(decoder_info->ref[0],decoder_info->rec);"
i = once(src, c)
src = src[:i] + ref_call + "(decoder_info->ref[0],decoder_info->rec);" + src[i + len(c):]
open(sys.argv[2], "w").write(src)
