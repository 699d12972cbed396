/*
 * thorcuda.h -- C ABI of libthorcuda, the B200 (sm_100a) pixel pipeline for the Thor video codec.
 *
 * Two groups of entry points:
 *
 *  (1) thorcuda_*   : frame-level / batched calls.  This is the performance path: pictures live in
 *                     HBM, one launch processes whole frames (of one or many independent streams) or a
 *                     whole list of blocks.
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
 * Every call only QUEUES work on the library's stream; host output buffers are valid after
 * thorcuda_sync().  Host input buffers are consumed (copied to a pinned staging ring) before the
 * call returns, so the caller may reuse them immediately.
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
/* A marker behind everything queued SO FAR (library stream and the bulk fetches of its copy stream): wait for it without
 * waiting for what is queued later.  For callers that keep several independent groups of streams in flight: submit
 * group B while group A's results are still crossing PCIe, then wait for A alone.  At most 16 open fences; wait frees. */
int         thorcuda_fence_record(void);           /* returns the fence id, or -1                    */
int         thorcuda_fence_wait(int id);
void       *thorcuda_stream(void);                 /* the cudaStream_t all work is queued on         */
/* number of kernels this library has launched since init (bench.py "gpu_launches")                 */
uint64_t    thorcuda_launch_count(void);
/* abs-diffs per second of a register-only VABSDIFF4 loop on this device: the ceiling of the integer ALU pipe that bounds
 * the SAD kernels (the denominator of bench.py's roofline.alu).  iters <= 0: default length.  Synchronous.       */
double      thorcuda_calibrate_absdiff(int iters);
/* CUDA-event timing of everything queued between begin and end, on the library's stream (ms).
 * Nestable up to 64 open timers: begin returns a timer id, end(id) queues the stop event,
 * elapsed(id) synchronises on it.                                                                    */
int         thorcuda_timer_begin(void);
int         thorcuda_timer_end(int id);
float       thorcuda_timer_elapsed_ms(int id);
/* write `bytes` of device scratch (>= L2 size) to evict the L2 between timed iterations             */
int         thorcuda_flush_l2(size_t bytes);

/* pinned host memory for staging (replaces malloc in create_yuv_frame, common_frame.c:338-339)     */
void       *thorcuda_host_alloc(size_t bytes);
void        thorcuda_host_free(void *p);

/* ------------------------------------------------------------------ pictures ---- */
/* A picture = padded I420 planes resident in HBM with EXACTLY the reference geometry
 * (create_yuv_frame, common/common_frame.c:324-351): stride_y = (w+2*pad_y+15)&~15,
 * stride_c = (w/2+2*pad_c+15)&~15, (h+2*pad) rows, Y then U then V.  Reference pictures use
 * pad_y=96, pad_c=48 (enc/mainenc.c:165); orig/rec use 0.                                           */
typedef int thorcuda_pic;                          /* handle >= 0                                    */
#define THORCUDA_MAX_PICS 4096
thorcuda_pic thorcuda_pic_create(int width, int height, int pad_y, int pad_c);
void         thorcuda_pic_destroy(thorcuda_pic p); /* replaces close_yuv_frame (:353-357)             */
int          thorcuda_pic_stride(thorcuda_pic p, int plane);
/* copy the visible area host<->device (queued; host rows at the given strides, any alignment).
 * The HOST buffer must stay valid until thorcuda_sync() (it is the DMA source/target; use
 * thorcuda_host_alloc memory for full-speed transfers).                                              */
/* Page-lock an existing host allocation (e.g. the malloc'ed planes of the reference's create_yuv_frame) so that
 * uploads / downloads DMA directly instead of going through the staging ring.  unregister before free().       */
int  thorcuda_host_register(void *p, size_t bytes);
void thorcuda_host_unregister(void *p);
int thorcuda_pic_upload(thorcuda_pic p, const uint8_t *y, const uint8_t *u, const uint8_t *v, int stride_y, int stride_c);
int thorcuda_pic_download(thorcuda_pic p, uint8_t *y, uint8_t *u, uint8_t *v, int stride_y, int stride_c);
/* whole padded planes (borders included), for tests                                                 */
int thorcuda_pic_download_padded(thorcuda_pic p, uint8_t *y, uint8_t *u, uint8_t *v);
int thorcuda_pic_upload_padded(thorcuda_pic p, const uint8_t *y, const uint8_t *u, const uint8_t *v);
/* raw device pointer to the visible origin of a plane (for callers that own CUDA code / torch)     */
void *thorcuda_pic_devptr(thorcuda_pic p, int plane);
/* device-side copy of the visible area src -> dst (same size; paddings may differ)                  */
int thorcuda_pic_copy(thorcuda_pic dst, thorcuda_pic src);

/* ---------------------------------------------------- in-loop filters, frame tail ---- */
/* Side information: the reference's deblock_data_t array (common/types.h:127-135, 44 bytes per
 * 4x4 luma unit, (h/4)*(w/4) entries, host memory).  Uploaded, packed to 2 bytes/unit on the device
 * and attached to picture `rec` (it describes that picture's blocks).                               */
int thorcuda_blockinfo_upload(thorcuda_pic rec, const void *deblock_data);
/* The same, and additionally which superblocks clpf_frame asks its decision callback about (common/common_frame.c:507-521:
 * some 8x8 block of the SB is not bi-predicted and has coded coefficients): sb_signals[k * (w/64) + l], (h/64)*(w/64)
 * bytes, written before the call returns.  The decoder reads one CLPF bit from the stream per signalling superblock. */
int thorcuda_blockinfo_upload_signals(thorcuda_pic rec, const void *deblock_data, uint8_t *sb_signals);

#define THORCUDA_LF_DEBLOCK_Y   1   /* deblock_frame_y   common/common_frame.c:46-241               */
#define THORCUDA_LF_DEBLOCK_UV  2   /* deblock_frame_uv  common/common_frame.c:243-321              */
#define THORCUDA_LF_CLPF        4   /* clpf_frame        common/common_frame.c:485-557              */
#define THORCUDA_LF_MAKE_REF    8   /* create_reference_frame common/common_frame.c:464-483         */
/* One fused pass over `rec` (unpadded, in-place semantics as in the reference) using the side
 * information attached to it:
 *   deblock Y (qp) and UV (chroma_qp[qp]) -> CLPF -> optionally copy+pad into `ref_out`.
 * CLPF decision: org >= 0 (encoder, enc/encode_frame.c:50-63): per-SB bit = SSD(filtered) < SSD(unfiltered)
 * over coded non-bipred 8x8 luma blocks, written to sb_bits_out[] (0xFF = SB signals nothing);
 * org < 0 (decoder, dec/decode_frame.c:130-133): bits taken from sb_bits_in[] (one byte per full SB,
 * raster order, (h/64)*(w/64) entries; entries of non-candidate SBs are ignored).
 * sb_bits_* are host pointers; sb_bits_out is valid after thorcuda_sync().                           */
int thorcuda_loopfilter_frame(thorcuda_pic rec, thorcuda_pic org, thorcuda_pic ref_out, int qp, int stages,
                              const uint8_t *sb_bits_in, uint8_t *sb_bits_out);
/* The same for n frames of n independent streams in ONE launch (grid.z = frame): this is how a B200
 * is kept busy, a single 1080p frame is 3 MB.  Pictures may differ in size.                         */
typedef struct {
  thorcuda_pic   rec, org, ref_out;      /* org/ref_out: -1 when unused                              */
  int            qp, stages;
  const uint8_t *sb_bits_in;             /* host, may be NULL                                        */
  uint8_t       *sb_bits_out;            /* host, may be NULL                                        */
} thorcuda_lf_job;
int thorcuda_loopfilter_frames(int n, const thorcuda_lf_job *jobs);
/* pad_yuv_frame (common/common_frame.c:405-462) on a padded picture                                 */
int thorcuda_pad_frame(thorcuda_pic p);

/* ------------------------------------------------------------------ task lists ---- */
/* Block-level work is submitted as task lists.  A list is validated, expanded and made resident in
 * HBM once (thorcuda_list_create) and can then be executed any number of times; the *_blocks
 * convenience calls below create, run and release a list in one go.  Tasks name pictures by INDEX
 * into the pics[] table given at creation, so one list can span many frames / streams.             */
typedef int thorcuda_list;
#define THORCUDA_LIST_PRED 0            /* thorcuda_pred_task                                        */
#define THORCUDA_LIST_TB   1            /* thorcuda_tb_task                                          */
#define THORCUDA_LIST_COST 2            /* thorcuda_cost_task                                        */
/* Lifetime: a list is validated against the pictures it names (every vector against their padded extent -- the kernels never
 * clamp).  Destroying or recreating one of those pictures invalidates the list: every *_list call then fails with an error
 * instead of launching; destroy the list and create it again. */
thorcuda_list thorcuda_list_create(int kind, const thorcuda_pic *pics, int npics, const void *tasks, int ntasks);
void          thorcuda_list_destroy(thorcuda_list l);

/* ------------------------------------------------------------ inter prediction ---- */
/* One prediction block.  (x,y) luma position, (w,h) luma size (4..64, multiples of 4; chroma is
 * w/2 x h/2).  mv in quarter-pel luma units as stored by the reference (mv_t, common/types.h:105-109);
 * sign bit i negates mv i before use (get_inter_prediction_luma, common/inter_prediction.c:125-126).
 * ref1 < 0: uni-prediction.  ref1 >= 0: both predictions are formed and averaged (a+b)>>1
 * (truncating; enc/encode_block.c:1569-1588, dec/decode_block.c:272-283).                           */
typedef struct {
  uint16_t x, y;
  uint8_t  w, h;
  int8_t   ref0, ref1;      /* index into pics[]                                                    */
  int16_t  mv0x, mv0y, mv1x, mv1y;
  uint8_t  sign0, sign1;
  uint8_t  dst;             /* index into pics[]: picture the prediction is written to              */
  uint8_t  reserved;
} thorcuda_pred_task;       /* 20 bytes */

/* Predict all tasks (Y, U and V).  `bipred_filter` selects the luma coefficient set exactly like the
 * reference's `bipred` argument (the encoder CONFIG flag).  planes: bit0 luma, bit1 chroma.         */
int thorcuda_predict_list(thorcuda_list l, int bipred_filter, int planes);
int thorcuda_predict_blocks(const thorcuda_pic *pics, int npics, const thorcuda_pred_task *tasks, int ntasks,
                            int bipred_filter, int planes);

/* --------------------------------------- residual / transform / quant / reconstruct ---- */
/* One transform block of one plane.  size 4..64.  (x,y) in that plane's coordinates.
 * Coefficient pool layout: task t owns q*q int16 entries at coeff_offset, q = min(size,16), row-major
 * with stride q -- the low-frequency corner, the only part the reference ever produces or reads
 * (common/transform.c:269,309,320; enc/write_bits.c:114-135).                                       */
typedef struct {
  uint16_t x, y;
  uint8_t  size;
  uint8_t  plane;           /* 0 Y, 1 U, 2 V                                                         */
  uint8_t  qp;              /* already chroma-mapped for planes 1,2                                  */
  uint8_t  coeff_type;      /* bit0 chroma, bit1 intra  (quantize(), enc/encode_block.c:75)          */
  uint8_t  fast;            /* transform() `fast` flag (common/transform.c:273)                      */
  uint8_t  org, pred, rec;  /* indices into pics[] (decoder leg ignores org)                         */
  uint32_t coeff_offset;    /* int16 index into the coefficient pool                                 */
} thorcuda_tb_task;         /* 16 bytes */

/* Encoder leg (encode_and_reconstruct_block_inter, enc/encode_block.c:1469-1532), per task:
 *   res = org - pred; coeff = transform(res); cbp = quantize(coeff) ; if cbp: rec = clip(pred +
 *   inverse_transform(dequantize(coeffq))) else rec = pred.
 * pred and rec may be the same picture (in place).  Quantised coefficients go to the device pool and
 * are copied to host_coeffq (if non-NULL, pool_entries int16); cbp flags (one byte per task) to
 * host_cbp; SSD(org,rec) per task (ssd_calc, enc/encode_block.c:782-797) to host_ssd.               */
int thorcuda_encode_list(thorcuda_list l, int16_t *host_coeffq, uint8_t *host_cbp, uint32_t *host_ssd);
int thorcuda_encode_blocks(const thorcuda_pic *pics, int npics, const thorcuda_tb_task *tasks, int ntasks,
                           int16_t *host_coeffq, uint8_t *host_cbp, uint32_t *host_ssd);
/* Decoder leg (decode_and_reconstruct_block_inter, dec/decode_block.c:90-120):
 *   rec = clip(pred + inverse_transform(dequantize(coeffq))) for every task (cbp != 0 blocks only).
 * host_coeffq == NULL: use the pool contents already on the device (e.g. left by encode_list).      */
int thorcuda_reconstruct_list(thorcuda_list l, const int16_t *host_coeffq);
int thorcuda_reconstruct_blocks(const thorcuda_pic *pics, int npics, const thorcuda_tb_task *tasks, int ntasks,
                                const int16_t *host_coeffq);
size_t thorcuda_list_pool_entries(thorcuda_list l);   /* int16 entries of a TB list's pool            */

/* ----------------------------------------------------------- intra prediction ---- */
/* make_top_and_left + get_intra_prediction (common/intra_prediction.c:57-143, :363-389) for a list of blocks.
 * Neighbours are read from picture `src` around the block -- row above, column left, corner; 128 where the block
 * touches the top / left picture edge; one extra sample beyond `size` when flagged available, the last one then
 * repeated up to 2*size -- and the size x size prediction is written into picture `dst` at the block position.
 * x, y are the position of the predicted block in ITS plane: under tb_split each of the four sub-blocks is a task
 * with the availability the reference derives for it (:106-110):
 *     upright  = (j == 0 || (i == 0 && cb_upright_available)),  downleft = (j == 0 && (i == 0 || cb_downleft_available))
 * (a flag set on a block in the first row / column is legal and unused: those neighbours are 128).
 * cb_x is the plane x of the CODING block: the reference takes the corner sample only if that is > 0 (:83,:116 test
 * xpos, not xpos + j) and top[0] otherwise.  The decoder predicts chroma with the luma mode and the CB's flags
 * (dec/decode_block.c:234-239).
 * Tasks of ONE call must be independent: the caller batches by wavefront (a block may enter a call once the blocks
 * owning its neighbour samples are reconstructed -- thorcuda_reconstruct_blocks on the same picture).  Neighbours are
 * read before anything is written, so src == dst is allowed.                                                     */
typedef struct {
  uint16_t x, y;            /* position of the block in its plane                                    */
  uint16_t cb_x;            /* plane x of the coding block (= x unless tb_split)                     */
  uint8_t  size;            /* 4, 8, 16, 32, 64                                                      */
  uint8_t  plane;           /* 0 = Y, 1 = U, 2 = V                                                   */
  uint8_t  mode;            /* intra_mode_t (common/types.h:137-149); other values predict DC (:387) */
  uint8_t  flags;           /* bit0 up-right sample available, bit1 down-left sample available       */
  uint8_t  src, dst;        /* indices into pics[]                                                   */
} thorcuda_intra_task;      /* 12 bytes */
int thorcuda_intra_predict_blocks(const thorcuda_pic *pics, int npics, const thorcuda_intra_task *tasks, int ntasks);
/* A whole frame's intra blocks in ONE call (dec/decode_block.c:48-88 for every intra block of the frame): the caller
 * sorts the blocks into wavefront levels -- level(block) = 1 + the highest level among the blocks owning the samples it
 * reads, blocks reconstructed before the call are level 0 -- and passes, both in level order,
 *   tasks[intra_first[l] .. intra_first[l+1])   the predictions of level l, and
 *   tb[tb_first[l] .. tb_first[l+1])            the coded transform blocks of level l (decoder leg: pred = rec = the
 *                                                picture, coefficients in host_coeffq at their coeff_offset);
 * per level the predictions are written, then the coded blocks are reconstructed in place.  Tasks and coefficients
 * are uploaded once; the device work is two launches per level, queued on the library stream.                        */
int thorcuda_intra_reconstruct_levels(const thorcuda_pic *pics, int npics, const thorcuda_intra_task *tasks, const int *intra_first,
                                      const thorcuda_tb_task *tb, const int *tb_first, int nlevels, const int16_t *host_coeffq);

/* ------------------------------------------------------------ temporal interpolation ---- */
/* interpolate_frames (common/temporal_interp.c:972-1053): the picture between two references that B frames of the
 * RA / HDB configurations use as reference 0 (callers: enc/mainenc.c:328,385, dec/decode_frame.c:92-109).
 *   ref0, ref1  padded pictures of equal geometry (the codec's reference ring: border 96 / 48); the border must be
 *               valid (thorcuda_pad_frame / the fused frame tail), it is read wherever a vector leaves the picture
 *   ratio, pos  distance between the references and position of the new picture, the reference's arguments:
 *               weights wt0 = pos > ratio/2 ? pos : ratio - pos, wt1 = ratio - wt0; the picture roles swap when
 *               pos > ratio/2 (:120-126)
 *   dst         receives the visible picture; the caller pads it (thorcuda_pad_frame) like decode_frame.c:107.
 * Pictures smaller than 32 in either dimension have no pyramid level: the call does nothing, like the reference.
 * All work is queued on the library stream (3 + 2 launches per level).                                             */
int thorcuda_interpolate_frames(thorcuda_pic dst, thorcuda_pic ref0, thorcuda_pic ref1, int ratio, int pos);
int thorcuda_interp_levels(int width, int height);     /* pyramid levels: min(4, (int)(log2(min(w, h)) - 4)) (:977) */
/* inspection: the final full-resolution motion field of the last call (vectors into the farther picture, 1/8 pel,
 * x then y, one per 8x8 block, bw = 2*ceil(w/16) per row); mv1 may be NULL to query the size only. Synchronous.   */
int thorcuda_interp_field(int16_t *mv1, int *bw, int *bh);

/* ------------------------------------------------------------ bit-cost model ---- */
/* First slice of the device-side bit-cost model: for every task of a transform-block list the exact
 * number of bits write_coeff (enc/write_bits.c:110-253) emits for the block's quantised levels -- the pool contents
 * thorcuda_encode_list has just produced, or host_coeffq (pool_entries int16, uploaded first) when non-NULL.  The block
 * type is the task's coeff_type (bit0 chroma, bit1 intra), the transform size its size (the chroma code tables switch
 * at size <= 8, :196,:239).  A block without levels costs 0 (the reference does not call write_coeff for it).
 * host_bits (ntasks uint32, may be NULL) is valid after thorcuda_sync().  With it cost_calc's `nbits`
 * (enc/encode_block.c:1218-1228) of a trial encode is these counts plus the host's mode / vector / cbp bits. */
int thorcuda_coeff_bits_list(thorcuda_list l, const int16_t *host_coeffq, uint32_t *host_bits);
/* Second slice: the whole of write_block (enc/write_bits.c:364-650) as a length -- mode / reference (write_super_mode,
 * :268-362), delta qp, intra mode, partition, vector differences (write_mv), skip / merge index, cbp codes -- plus the
 * coefficient bits of the block's transform blocks.  One record per coded block (everything write_block reads):      */
typedef struct { int16_t x, y; } thorcuda_mv;               /* mv_t, common/types.h:105-109 */
typedef struct {
  /* frame_info / params */
  uint8_t frame_type;                 /* 0 I, 1 P, 2 B (frame_type_t, common/types.h:69-74)                       */
  uint8_t num_ref, enable_bipred, interp_ref, num_intra_modes, max_delta_qp;
  /* block_info: block_pos.size, block_context->index / ->cbp, partition limits, candidate counts, delta qp */
  uint8_t size, ctx_index; int8_t ctx_cbp;
  uint8_t max_num_pb_part, max_num_tb_part, num_skip_vec, num_merge_vec; int8_t delta_qp;
  /* block_param: mode 0 skip 1 intra 2 inter 3 bipred 4 merge (block_mode_t, common/types.h:76-83) */
  uint8_t mode, intra_mode, pb_part, tb_split, skip_idx, ref_idx0, ref_idx1;
  uint8_t cbp_y, cbp_u, cbp_v;        /* flags; under tb_split 4-bit masks, quadrant 0 in bit 3 (:589-591)        */
  thorcuda_mv mvp, mv0[4], mv1[4];    /* block_info->mvp, mv_arr0, mv_arr1                                        */
  uint32_t tb_first;                  /* first transform block of this block in the list, in write_block's order:
                                         Y U V | tb_split, size > 8: (Y U V) x 4 quadrants | size 8: Y0..Y3 U V;
                                         0xffffffff: syntax bits only                                              */
} thorcuda_block_syntax;              /* 64 bytes */
/* host_bits[i] = what write_block returns for block i.  `l` (a transform-block list on which thorcuda_coeff_bits_list
 * has run, or -1 for the syntax bits alone) supplies the coefficient bits; a transform block counts where the block's
 * cbp bit is set, as in the reference.  Records are validated on the host (error, no launch).  host_bits is valid
 * after thorcuda_sync().                                                                                           */
int thorcuda_block_bits(thorcuda_list l, const thorcuda_block_syntax *host_blocks, int nblocks, uint32_t *host_bits);
/* The syntax bits of the same records (what thorcuda_block_bits(-1, ...) returns) through an entry point that ALLOCATES NOTHING.
 * No counterpart in the reference: added for callers that cannot call cudaMalloc (CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  The device-side copy of the records and of the results lives in a
 * file-scope device array of the library, sized for THORCUDA_BLOCK_BITS_STATIC_MAX blocks (272 KiB, held by every GPU that loads the
 * library until exit); more blocks are an error (split the call).  The call makes no cudaMalloc / cudaFree / pinned allocation: host
 * buffers are copied directly.  It is synchronous (host_bits is valid on return).  Every call shares the one array, which keeps the
 * previous call's contents (nothing in it needs clearing: the entries a call reads are the ones it has just written); calls that
 * overlap on the GPU -- several host threads, or a caller driving several contexts -- race on it and must be serialised. */
#define THORCUDA_BLOCK_BITS_STATIC_MAX 4096
int thorcuda_block_bits_static(const thorcuda_block_syntax *host_blocks, int nblocks, uint32_t *host_bits);
/* host arithmetic (no device work): quote_vlc (enc/putvlc.c:133-229; -1 where the reference aborts) and the bits of
 * write_mv (enc/write_bits.c:50-69) = quote_mv_bits (enc/encode_block.c:799-814) */
int thorcuda_vlc_len(unsigned n, unsigned cn);
int thorcuda_mv_bits(int mvx, int mvy, int mvpx, int mvpy);

/* ------------------------------------------------------------ early-skip test ---- */
/* Significance test of a skip candidate (check_early_skip_block, enc/encode_block.c:2612-2741) for one
 * size x size luma sub-block (size = min(block size, 16): 16 or 8) and its two chroma sub-blocks, given the
 * candidate's prediction already formed in picture `pred` (thorcuda_predict_*):
 *   luma   (check_early_skip_sub_block  :2507-2543): residual -> 2x2 mean (sum+2)>>2 -> transform(size/2, fast=0)
 *          -> any |coefficient| > thr_y
 *   chroma (check_early_skip_sub_blockC :2545-2610): column sums of the residual (8 wide: each column, compared as
 *          int16 like the reference's SIMD; 4 wide: pairs of columns) > thr_c -- positive excursions only.
 * The thresholds are the reference's only floating-point step; thorcuda_early_skip_thresholds() evaluates them
 * on the HOST in double exactly as the reference does ((int)(rel * (double)(1<<shift2)/(double)scale)), with
 * rel = 0.5*early_skip_thr for luma, early_skip_thr for chroma at qp chroma_qp[qp_y] and log2 size fixed to 5.
 * flags[i]: bit0 luma significant, bit1 U, bit2 V (the block is skippable iff all sub-blocks return 0).       */
typedef struct {
  uint16_t x, y;            /* luma position of the sub-block                                        */
  uint8_t  size;            /* 8 or 16                                                               */
  uint8_t  org, pred;       /* indices into pics[]                                                   */
  uint8_t  reserved;
  int32_t  thr_y, thr_c;
} thorcuda_skip_task;       /* 16 bytes */
void thorcuda_early_skip_thresholds(int qp_y, int size, float early_skip_thr, int *thr_y, int *thr_c);
int  thorcuda_early_skip_batch(const thorcuda_pic *pics, int npics, const thorcuda_skip_task *tasks, int ntasks,
                               uint8_t *host_flags);

/* -------------------------------------------------------------- motion search ---- */
/* Whole-frame integer-pel cost tables (north_star: "whole-frame per-block cost table").
 * For every block of `org` of size bs in {8,16,32,64} on the bs-aligned grid (blocks that do not
 * lie completely inside the frame are omitted: (w/bs) x (h/bs) blocks, raster order) and every
 * full-pel displacement (dx,dy) in [-range, range-1]^2 (a 2*range x 2*range window, range a multiple of 4,
 * 4..32) around the superblock's search centre
 * (centres: int16 (cx,cy) pairs per 64x64 SB in raster order, host pointer, NULL = all (0,0)):
 *      table[block][dy+range][dx+range] = sad_calc(org block, ref + (cy+dy)*stride + cx+dx)
 * (rows of 2*range entries: for range 16 a block's table is 2 KiB and every row one 64-byte line).
 * uint16 for bs <= 16, uint32 for bs >= 32.  Rectangular partitions (HOR/VER/QUAD halves) are sums
 * of two entries of the next finer level -- SAD is additive -- so the host forms them itself.
 * The call fails if a window of an existing block would leave the padded reference picture.
 * Tables live in HBM in one of THORCUDA_SAD_SLOTS slots; fetch copies one level to the host on the library's
 * copy stream (ordered after everything queued before it, overlapping kernels queued after it; the data is valid
 * after thorcuda_sync(); the next build into the slot waits for it).
 * n (org,ref) pairs are processed in ONE launch.                                                     */
#define THORCUDA_SAD_SLOTS 64
/* Host helpers for forming table indices (pure host code, usable before thorcuda_init):
 *  thorcuda_clip_mv        = clip_mv (enc/encode_block.c:816-828): keeps the block within 80 px of the frame; the
 *                            right-edge test omits "+ size" and "/ 4" truncates toward zero, exactly as there.
 *                            mv[0] = x, mv[1] = y in quarter-pel (mv_t, common/types.h:105-109).
 *  thorcuda_sad_table_index: entry index inside one level for the integer-pel probe that motion_estimate makes
 *                            at mv (already clipped; the search walks ref + s*(mv>>2), s = sign ? -1 : 1), or -1 if
 *                            the probe lies outside the block's window (then use thorcuda_cost_batch).          */
void    thorcuda_clip_mv(int16_t mv[2], int ypos, int xpos, int fwidth, int fheight, int size, int sign);
int64_t thorcuda_sad_table_index(int fwidth, int bs, int xpos, int ypos, const int16_t mv[2], int sign,
                                 int centre_x, int centre_y, int range);
int    thorcuda_sad_table_build(int n, const int *slots, const thorcuda_pic *orgs, const thorcuda_pic *refs,
                                int range, const int16_t *const *centres);
size_t thorcuda_sad_table_bytes(int slot, int bs);
int    thorcuda_sad_table_fetch(int slot, int bs, void *host_dst, size_t bytes);
void  *thorcuda_sad_table_devptr(int slot, int bs);

/* Generic batched costs: one task = one (block, candidate) pair.                                    */
typedef struct {
  uint16_t x, y;            /* luma position of the original block                                   */
  uint8_t  w, h;            /* prediction block size (4..64)                                         */
  uint8_t  org, ref;        /* indices into pics[]                                                   */
  int8_t   hx, hy;          /* fastquarter only: incoming half-pel (*x,*y) in {-2,0,2}               */
  int16_t  ix, iy;          /* full-pel displacement of the reference block                          */
  uint16_t reserved;
} thorcuda_cost_task;       /* 16 bytes */
typedef struct { uint32_t cost; int16_t x, y; } thorcuda_cost_result;

#define THORCUDA_COST_SAD          0   /* sad_calc            enc/encode_block.c:740-755            */
#define THORCUDA_COST_WIDESAD      1   /* widesad_calc        enc/encode_block.c:757-780  (x out)   */
#define THORCUDA_COST_FASTHALF     2   /* sad_calc_fasthalf   enc/encode_block.c:497-605  (x,y out) */
#define THORCUDA_COST_FASTQUARTER  3   /* sad_calc_fastquarter enc/encode_block.c:609-738 (x,y out) */
#define THORCUDA_COST_SSD          4   /* ssd_calc            enc/encode_block.c:782-797            */
int thorcuda_cost_list(thorcuda_list l, int kind, thorcuda_cost_result *host_results);
int thorcuda_cost_batch(int kind, const thorcuda_pic *pics, int npics, const thorcuda_cost_task *tasks, int ntasks,
                        thorcuda_cost_result *host_results);

#ifdef __cplusplus
}
#endif
#endif
