/*
 * thorcuda.h -- C ABI of libthorcuda, the B200 (sm_100a) pixel pipeline for the Thor video codec.
 *
 * Two groups of entry points:
 *
 *  (1) thorcuda_*   : frame-level / batched calls.  This is the performance path: pictures live in
 *                     HBM, one launch processes a whole frame (or a whole list of blocks).
 *  (2) the reference's own kernel names (thorcuda_compat.h): per-block, synchronous, host pointers.
 *                     Same symbol names and signatures as awakecoding/thor's common_kernels.h,
 *                     enc_kernels.h and transform.h so Thorenc/Thordec link against libthorcuda
 *                     unchanged.  They exist for API fidelity and kernel-level parity tests.
 *
 * There is no CPU fallback anywhere: without an sm_100 device thorcuda_init() fails, and every
 * other call aborts the process the way the reference's fatalerror() does (common/global.h:38-44).
 *
 * All functions that return int return 0 on success, non-zero on error (thorcuda_last_error()).
 * The library is single-context, single-host-thread, like the reference (SURVEY 8b "Threading").
 */
#ifndef THORCUDA_H
#define THORCUDA_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ context ---- */
/* Replaces init_use_simd() (common/simd.h:134-140; called enc/mainenc.c:101, dec/maindec.c:108). */
int         thorcuda_init(int device);
void        thorcuda_shutdown(void);
const char *thorcuda_last_error(void);
int         thorcuda_sync(void);                   /* wait for all queued work                       */
void       *thorcuda_stream(void);                 /* the cudaStream_t all work is queued on         */
/* number of kernels this library has launched since init (bench.py "gpu_launches")                 */
uint64_t    thorcuda_launch_count(void);

/* pinned host memory for staging (replaces malloc in create_yuv_frame, common_frame.c:338-339)     */
void       *thorcuda_host_alloc(size_t bytes);
void        thorcuda_host_free(void *p);

/* ------------------------------------------------------------------ pictures ---- */
/* A picture = padded I420 planes resident in HBM with EXACTLY the reference geometry
 * (create_yuv_frame, common/common_frame.c:324-351): stride_y = (w+2*pad_y+15)&~15,
 * stride_c = (w/2+2*pad_c+15)&~15, (h+2*pad) rows, Y then U then V.  Reference pictures use
 * pad_y=96, pad_c=48 (enc/mainenc.c:165); orig/rec use 0.                                           */
typedef int thorcuda_pic;                          /* handle >= 0                                    */
thorcuda_pic thorcuda_pic_create(int width, int height, int pad_y, int pad_c);
void         thorcuda_pic_destroy(thorcuda_pic p); /* replaces close_yuv_frame (:353-357)             */
int          thorcuda_pic_stride(thorcuda_pic p, int plane);
/* copy the visible area host<->device (queued; host rows at the given strides, any alignment)      */
int thorcuda_pic_upload(thorcuda_pic p, const uint8_t *y, const uint8_t *u, const uint8_t *v, int stride_y, int stride_c);
int thorcuda_pic_download(thorcuda_pic p, uint8_t *y, uint8_t *u, uint8_t *v, int stride_y, int stride_c);
/* whole padded planes (borders included), for tests                                                 */
int thorcuda_pic_download_padded(thorcuda_pic p, uint8_t *y, uint8_t *u, uint8_t *v);
int thorcuda_pic_upload_padded(thorcuda_pic p, const uint8_t *y, const uint8_t *u, const uint8_t *v);
/* raw device pointer to the visible origin of a plane (for callers that own CUDA code / torch)     */
void *thorcuda_pic_devptr(thorcuda_pic p, int plane);

/* ---------------------------------------------------- in-loop filters, frame tail ---- */
/* Side information: the reference's deblock_data_t array (common/types.h:127-135, 44 bytes per
 * 4x4 luma unit, (h/4)*(w/4) entries, host memory).  Uploaded and packed to 2 bytes/unit on device. */
int thorcuda_blockinfo_upload(const void *deblock_data, int width, int height);

#define THORCUDA_LF_DEBLOCK_Y   1   /* deblock_frame_y   common/common_frame.c:46-241               */
#define THORCUDA_LF_DEBLOCK_UV  2   /* deblock_frame_uv  common/common_frame.c:243-321              */
#define THORCUDA_LF_CLPF        4   /* clpf_frame        common/common_frame.c:485-557              */
#define THORCUDA_LF_MAKE_REF    8   /* create_reference_frame common/common_frame.c:464-483         */
/* One fused pass over `rec` (unpadded, in-place semantics as in the reference) using the side
 * information of the last thorcuda_blockinfo_upload():
 *   deblock Y (qp) and UV (chroma_qp[qp]) -> CLPF -> optionally copy+pad into `ref_out`.
 * CLPF decision: org >= 0 (encoder, enc/encode_frame.c:50-63): per-SB bit = SSD(filtered) < SSD(unfiltered)
 * over coded non-bipred 8x8 luma blocks, written to sb_bits_out[] (0xFF = SB signals nothing);
 * org < 0 (decoder, dec/decode_frame.c:130-133): bits taken from sb_bits_in[] (one byte per full SB,
 * raster order, (h/64)*(w/64) entries; entries of non-candidate SBs are ignored).
 * sb_bits_* are host pointers; sb_bits_out is valid after thorcuda_sync().                           */
int thorcuda_loopfilter_frame(thorcuda_pic rec, thorcuda_pic org, thorcuda_pic ref_out, int qp, int stages,
                              const uint8_t *sb_bits_in, uint8_t *sb_bits_out);
/* pad_yuv_frame (common/common_frame.c:405-462) on a padded picture                                 */
int thorcuda_pad_frame(thorcuda_pic p);

/* ------------------------------------------------------------ inter prediction ---- */
/* One prediction block.  (x,y) luma position, (w,h) luma size (4..64, multiples of 4; chroma is
 * w/2 x h/2).  mv in quarter-pel luma units as stored by the reference (mv_t, common/types.h:105-109);
 * sign bit i negates mv i before use (get_inter_prediction_luma, common/inter_prediction.c:125-126).
 * ref1 < 0: uni-prediction.  ref1 >= 0: both predictions are formed and averaged (a+b)>>1
 * (truncating; enc/encode_block.c:1569-1588, dec/decode_block.c:272-283).                           */
typedef struct {
  uint16_t x, y;
  uint8_t  w, h;
  int8_t   ref0, ref1;      /* index into the refs[] array passed to the call                       */
  int16_t  mv0x, mv0y, mv1x, mv1y;
  uint8_t  sign0, sign1;
  uint16_t reserved;
} thorcuda_pred_task;       /* 20 bytes */

/* Predict all tasks from refs[] into picture dst (Y, U and V).  `bipred_filter` selects the luma
 * coefficient set exactly like the reference's `bipred` argument (the encoder CONFIG flag).
 * tasks is a host pointer.  planes: bit0 luma, bit1 chroma.                                         */
int thorcuda_predict_blocks(thorcuda_pic dst, const thorcuda_pic *refs, int nrefs,
                            const thorcuda_pred_task *tasks, int ntasks, int bipred_filter, int planes);

/* --------------------------------------- residual / transform / quant / reconstruct ---- */
/* One transform block of one plane.  size 4..64.  (x,y) in that plane's coordinates.               */
typedef struct {
  uint16_t x, y;
  uint8_t  size;
  uint8_t  plane;           /* 0 Y, 1 U, 2 V                                                         */
  uint8_t  qp;              /* already chroma-mapped for planes 1,2                                  */
  uint8_t  coeff_type;      /* bit0 chroma, bit1 intra  (quantize(), enc/encode_block.c:75)          */
  uint8_t  fast;            /* transform() `fast` flag (common/transform.c:273)                      */
  uint8_t  reserved[3];
  uint32_t coeff_offset;    /* int16 index into the coefficient pool (size*size entries per block)   */
} thorcuda_tb_task;         /* 16 bytes */

/* Encoder leg (encode_and_reconstruct_block_inter, enc/encode_block.c:1469-1532), per task:
 *   res = org - pred; coeff = transform(res); cbp = quantize(coeff) ; if cbp: rec = clip(pred +
 *   inverse_transform(dequantize(coeffq))) else rec = pred.
 * pred and rec may be the same picture (in place).  Quantised coefficients go to the device pool and
 * are copied to host_coeffq (if non-NULL); cbp flags (one byte per task) to host_cbp; SSD(org,rec)
 * per task (ssd_calc, enc/encode_block.c:782-797) to host_ssd (if non-NULL).                         */
int thorcuda_encode_blocks(thorcuda_pic org, thorcuda_pic pred, thorcuda_pic rec,
                           const thorcuda_tb_task *tasks, int ntasks, size_t pool_entries,
                           int16_t *host_coeffq, uint8_t *host_cbp, uint32_t *host_ssd);
/* Decoder leg (decode_and_reconstruct_block_inter, dec/decode_block.c:90-120):
 *   rec = clip(pred + inverse_transform(dequantize(coeffq))) for every task (cbp != 0 blocks only). */
int thorcuda_reconstruct_blocks(thorcuda_pic pred, thorcuda_pic rec, const thorcuda_tb_task *tasks, int ntasks,
                                const int16_t *host_coeffq, size_t pool_entries);

/* -------------------------------------------------------------- motion search ---- */
/* Whole-frame integer-pel cost table (north_star: "whole-frame per-block cost table").
 * For every 8x8-aligned block of `org` of size bs (8,16,32,64) and every displacement
 * (dx,dy) in [-range,range]^2 around that block's centre (cx,cy) (full-pel; centres[] = int16 pairs
 * per block in raster order, or NULL for (0,0)):  table = SAD(org block, ref at +d)  == sad_calc().
 * Table layout: [block][dy+range][dx+range], uint16 for bs<=16, uint32 for bs>=32.
 * Positions whose reference block would leave the padded picture are written as 0xFFFF/0xFFFFFFFF.
 * The table stays on the device; thorcuda_sad_table_fetch copies a level to the host.              */
int thorcuda_sad_table_build(thorcuda_pic org, thorcuda_pic ref, int range, const int16_t *centres8);
int thorcuda_sad_table_fetch(int bs, void *host_dst, size_t bytes);
size_t thorcuda_sad_table_bytes(int bs);
void *thorcuda_sad_table_devptr(int bs);

/* Generic batched costs: one task = one (block, candidate) pair.                                    */
typedef struct {
  uint16_t x, y;            /* luma position of the original block                                   */
  uint8_t  w, h;            /* prediction block size (4..64)                                         */
  int8_t   ref;             /* index into refs[]                                                     */
  uint8_t  hx, hy;          /* fastquarter only: incoming half-pel x,y + 2  (0,2,4)                  */
  uint8_t  reserved;
  int16_t  ix, iy;          /* full-pel displacement of the reference block                          */
  uint16_t pad;
} thorcuda_cost_task;       /* 16 bytes */
typedef struct { uint32_t cost; int16_t x, y; } thorcuda_cost_result;

#define THORCUDA_COST_SAD          0   /* sad_calc            enc/encode_block.c:740-755            */
#define THORCUDA_COST_WIDESAD      1   /* widesad_calc        enc/encode_block.c:757-780  (x out)   */
#define THORCUDA_COST_FASTHALF     2   /* sad_calc_fasthalf   enc/encode_block.c:497-605  (x,y out) */
#define THORCUDA_COST_FASTQUARTER  3   /* sad_calc_fastquarter enc/encode_block.c:609-738 (x,y out) */
#define THORCUDA_COST_SSD          4   /* ssd_calc            enc/encode_block.c:782-797            */
int thorcuda_cost_batch(int kind, thorcuda_pic org, const thorcuda_pic *refs, int nrefs,
                        const thorcuda_cost_task *tasks, int ntasks, thorcuda_cost_result *host_results);

#ifdef __cplusplus
}
#endif
#endif
