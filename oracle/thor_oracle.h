/*
 * thor_oracle.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C99, scalar, single thread) of the Thor pixel pipeline that
 * libthorcuda implements on the GPU.  Every function cites the reference file:line
 * (paths relative to the awakecoding/thor tree) whose arithmetic it restates.
 * Parity status: PINNED -- tests/test_oracle_vs_ref.py checks every function here against
 * the unmodified reference compiled by oracle/Makefile into oracle/_ref/libthorref.so
 * (both its SIMD and its plain-C paths), and tests/golden/ holds vectors produced by that
 * build (tests/golden/make_golden.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg
 * may include, link or call this.  libthorcuda never does.
 */
#ifndef THOR_ORACLE_H
#define THOR_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Per-4x4-unit side information, laid out exactly like the reference's deblock_data_t
 * (common/types.h:127-135; 44 bytes, probed).  Block modes: common/types.h:76-83. */
typedef struct { int16_t x, y; } orc_mv_t;
typedef struct {
  int32_t mode;                 /* 0 SKIP 1 INTRA 2 INTER 3 BIPRED 4 MERGE */
  int32_t cbp_y, cbp_u, cbp_v;
  uint8_t size;                 /* coding block size 8..64 */
  uint8_t tb_split;
  int32_t pb_part;              /* 0 NONE 1 HOR 2 VER 3 QUAD */
  orc_mv_t mv0, mv1;
  uint32_t ref_idx0, ref_idx1, bipred_flag;
} orc_blockinfo_t;

/* ---- in-loop filters (common/common_frame.c) ---- */
void orc_deblock_frame_y(uint8_t *y, int stride, const orc_blockinfo_t *bi, int width, int height, int qp);
void orc_deblock_frame_uv(uint8_t *u, uint8_t *v, int stride_c, const orc_blockinfo_t *bi, int width, int height, int qpc);
int  orc_chroma_qp(int qp);
/* CLPF over all full 64x64 superblocks.  org==NULL: decoder mode, sb_bits_in[k*nsbx+l] gives the
 * signalled bit of each candidate SB.  org!=NULL: encoder mode, decision from SSDs, bits written to
 * sb_bits_out (0xff for SBs that signal nothing).  Returns number of signalled SBs. */
int  orc_clpf_frame(uint8_t *y, uint8_t *u, uint8_t *v, int stride_y, int stride_c,
                    const uint8_t *org_y, int org_stride, const orc_blockinfo_t *bi,
                    int width, int height, const uint8_t *sb_bits_in, uint8_t *sb_bits_out);
void orc_clpf_block(const uint8_t *src, uint8_t *dst, int sstride, int dstride, int x0, int y0, int size, int width, int height);
void orc_detect_clpf(const uint8_t *rec, const uint8_t *org, int x0, int y0, int width, int height, int so, int stride, int *sum0, int *sum1);
/* Replicate the visible picture into its borders; p points at the visible origin. */
void orc_pad_plane(uint8_t *p, int stride, int w, int h, int pad_hor, int pad_ver);

/* ---- inter prediction (common/inter_prediction.c:72-180) ---- */
void orc_inter_luma(uint8_t *pblock, const uint8_t *ref, int width, int height, int stride, int pstride, int mvx, int mvy, int sign, int bipred);
void orc_inter_chroma(uint8_t *pblock, const uint8_t *ref, int width, int height, int stride, int pstride, int mvx, int mvy, int sign);
void orc_average_trunc(uint8_t *dst, const uint8_t *a, const uint8_t *b, int n);             /* (a+b)>>1   */
void orc_block_avg(uint8_t *p, const uint8_t *r0, const uint8_t *r1, int sp, int s0, int s1, int w, int h); /* (a+b+1)>>1 */

/* ---- transforms, quantisation, residual/reconstruct ---- */
const int16_t *orc_dct_matrix(int size);        /* size 4,8,16,32: row-major size x size */
const int *orc_zigzag(int qsize);               /* qsize 4,8,16 */
void orc_transform(const int16_t *block, int16_t *coeff, int size, int fast);
void orc_inverse_transform(const int16_t *coeff, int16_t *block, int size);
int  orc_quantize(const int16_t *coeff, int16_t *coeffq, int qp, int size, int coeff_block_type);
void orc_dequantize(const int16_t *coeff, int16_t *rcoeff, int qp, int size);
void orc_get_residual(int16_t *block, const uint8_t *pblock, const uint8_t *orig, int size, int orig_stride);
void orc_reconstruct_block(const int16_t *block, const uint8_t *pblock, uint8_t *rec, int size, int stride);

/* ---- early-skip significance test (enc/encode_block.c:2481-2610) ---- */
int orc_early_skip_threshold(int qp, int size, double relative_threshold);
int orc_early_skip_luma(const uint8_t *orig, int orig_stride, int size, int qp, const uint8_t *pblock, float early_skip_threshold);
int orc_early_skip_chroma(const uint8_t *orig, int orig_stride, int size, int qp, const uint8_t *pblock, float early_skip_threshold);

/* ---- intra prediction (orc_intra.c; reference common/intra_prediction.c:39-389, common/common_block.c:110-128) ---- */
void orc_intra_neighbours(const uint8_t *rec, int stride, int x, int y, int cb_x, int size, int upright, int downleft,
                          uint8_t *left, uint8_t *top, uint8_t *top_left);
void orc_intra_predict(const uint8_t *left, const uint8_t *top, uint8_t top_left, int ypos, int xpos, int size, uint8_t *pblock, int mode);
int  orc_upright_available(int ypos, int xpos, int size, int width);
int  orc_downleft_available(int ypos, int xpos, int size, int height);

/* ---- temporal interpolation (orc_interp.c; reference common/temporal_interp.c:61-1053) ---- */
int  orc_interp_levels(int width, int height);
void orc_interp_counters(long *out6, int reset);
void orc_interp_field_size(int width, int height, int *bw, int *bh);
void orc_interpolate_frames(uint8_t *out_y, uint8_t *out_u, uint8_t *out_v, int out_stride_y, int out_stride_c,
                            const uint8_t *ref0_y, const uint8_t *ref0_u, const uint8_t *ref0_v,
                            const uint8_t *ref1_y, const uint8_t *ref1_u, const uint8_t *ref1_v,
                            int width, int height, int stride_y, int stride_c, int pad_y, int ratio, int pos,
                            int16_t *dbg_mv0, int16_t *dbg_mv1);

/* ---- motion-search costs (enc/encode_block.c:497-828) ---- */
unsigned orc_sad(const uint8_t *a, const uint8_t *b, int as, int bs, int w, int h);
unsigned orc_widesad(const uint8_t *a, const uint8_t *b, int as, int bs, int w, int h, int *x);
int      orc_ssd(const uint8_t *a, const uint8_t *b, int as, int bs, int w, int h);
unsigned orc_sad_fasthalf(const uint8_t *a, const uint8_t *b, int as, int bs, int w, int h, int *x, int *y);
unsigned orc_sad_fastquarter(const uint8_t *o, const uint8_t *r, int os, int rs, int w, int h, int *x, int *y);
void     orc_clip_mv(int16_t *mvx, int16_t *mvy, int ypos, int xpos, int fwidth, int fheight, int size, int sign);

/* ---- bit-cost model (enc/putvlc.c, enc/write_bits.c): lengths only, orc_bits.c ---- */
int orc_vlc_len(unsigned n, unsigned cn);                      /* quote_vlc; -1 where the reference aborts */
int orc_mv_bits(int mvx, int mvy, int mvpx, int mvpy);         /* write_mv */
int orc_find_code(int run, int level, int maxrun, int chroma, int eob);
int orc_coeff_bits(const int16_t *coeff, int stride, int size, int type);   /* write_coeff; 0 for an all-zero block */

/* write_block (enc/write_bits.c:364-650) on a flat int32 record: what the frame, the block's context and the decision
 * contribute to the syntax.  cbp_y/u/v are 1-bit flags, or 4-bit masks (sub-block 0 in bit 3) where tb_split is set. */
enum { ORC_BS_FRAME_TYPE = 0, ORC_BS_NUM_REF, ORC_BS_ENABLE_BIPRED, ORC_BS_INTERP_REF, ORC_BS_NUM_INTRA_MODES, ORC_BS_MAX_DELTA_QP,
       ORC_BS_SIZE, ORC_BS_CTX_INDEX, ORC_BS_CTX_CBP, ORC_BS_MAX_PB_PART, ORC_BS_MAX_TB_PART, ORC_BS_NUM_SKIP_VEC, ORC_BS_NUM_MERGE_VEC,
       ORC_BS_DELTA_QP, ORC_BS_MODE, ORC_BS_INTRA_MODE, ORC_BS_PB_PART, ORC_BS_TB_SPLIT, ORC_BS_SKIP_IDX, ORC_BS_REF_IDX0, ORC_BS_REF_IDX1,
       ORC_BS_CBP_Y, ORC_BS_CBP_U, ORC_BS_CBP_V, ORC_BS_MVP, ORC_BS_MV0 = ORC_BS_MVP + 2, ORC_BS_MV1 = ORC_BS_MV0 + 8,
       ORC_BS_FIELDS = ORC_BS_MV1 + 8 };
int orc_block_syntax_bits(const int32_t *d);     /* everything write_block emits except the write_coeff calls */
/* the whole of write_block: coeff_y is size x size (four contiguous (size/2)^2 sub-blocks under tb_split), coeff_u/v
 * (size/2)^2 likewise -- block_param_t's layout */
int orc_block_bits(const int32_t *d, const int16_t *coeff_y, const int16_t *coeff_u, const int16_t *coeff_v);

#ifdef __cplusplus
}
#endif
#endif
