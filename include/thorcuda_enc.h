/* Frame encoder of libthorcuda: plain-C records shared by the device code (enc_core.h), the library entry point
 * (encoder.cu, thorcuda_enc_frame) and the reference-side binding (integration/enc_gpu_hook.c).
 *
 * What the device decides per superblock is the reference's process_block() tree (enc/encode_block.c:2787-3033): for every
 * node of the 64/32/16/8 quad tree either "split", or one coded block with the parameters write_block()
 * (enc/write_bits.c:364-650) needs.  The host then only runs the reference's own write_super_mode / write_block on these
 * records (the entropy coder is out of scope). */
#ifndef THORCUDA_ENC_H
#define THORCUDA_ENC_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define THOR_ENC_MAX_REF 6           /* references a frame can address (max_num_ref <= 4, + the interpolated picture) */
#define THOR_ENC_NODES 85            /* 1 + 4 + 16 + 64 nodes of a superblock's quad tree                              */

enum { THOR_NODE_NONE = 0, THOR_NODE_SPLIT = 1, THOR_NODE_CODED = 2 };

/* One quad-tree node.  Node index = level offset {0, 1, 5, 21} + z-order index of the block inside the superblock. */
typedef struct {
  uint8_t state;                     /* THOR_NODE_*                                                                  */
  uint8_t mode;                      /* block_mode_t (common/types.h:76-83)                                          */
  uint8_t intra_mode;
  uint8_t skip_idx;
  uint8_t pb_part;
  int8_t  tb_param;                  /* block_param.tb_param: -1 zero block, 0, 1 split                              */
  int8_t  ref_idx0, ref_idx1;
  int8_t  dir;                       /* -1 intra, 0 uni, 2 bi (block_param.dir)                                      */
  uint8_t cbp_y, cbp_u, cbp_v;       /* as encode_block leaves them for write_block (4-bit masks under tb_split)      */
  int8_t  delta_qp;
  uint8_t num_skip_vec, num_merge_vec;   /* block_info.num_skip_vec / num_merge_vec (write_block reads them)         */
  int8_t  ctx_index;                 /* block_context->index                                                         */
  int8_t  ctx_cbp;                   /* block_context->cbp                                                           */
  uint8_t split_early;               /* a SPLIT node whose split code was written BEFORE the children (:2928-2935)   */
  uint8_t pad[2];
  int16_t mvp[2];                    /* block_info.mvp (x, y)                                                        */
  int16_t mv0[4][2];                 /* mv_arr0 (x, y)                                                               */
  int16_t mv1[4][2];
} thor_enc_node;                     /* 56 bytes */

/* Coefficients of a superblock in block_param_t's layout (common/types.h:164-166): a coded block of size N whose first
 * 8x8 unit has z-order index z owns y[64 z .. 64 z + N*N), u/v[16 z .. 16 z + N*N/4): compact, stride N (N/2).        */
typedef struct {
  int16_t y[4096];
  int16_t u[1024];
  int16_t v[1024];
} thor_enc_coef;                     /* 12288 bytes */

/* A plane triple as the reference's yuv_frame_t addresses it: pointers to the visible origin, strides in bytes. */
typedef struct {
  uint8_t *y, *u, *v;
  int32_t stride_y, stride_c;
  int32_t frame_num, reserved;
} thor_enc_pic;

/* Per-4x4-unit neighbour state the decisions read (the fields of deblock_data_t the search and the contexts use:
 * common/inter_prediction.c:182-600, common/common_block.c:158-178), 16 bytes.                                      */
typedef struct {
  int16_t mv0x, mv0y, mv1x, mv1y;
  uint8_t mode, size, cbp;           /* cbp: bit0 y, bit1 u, bit2 v                                                  */
  uint8_t flags;                     /* bit0 tb_split, bits 1-2 pb_part                                              */
  int8_t  ref_idx0, ref_idx1, dir;
  uint8_t pad;
} thor_enc_unit;

typedef struct {
  int32_t width, height;
  int32_t frame_type;                /* 0 I, 1 P, 2 B                                                                */
  int32_t qp;                        /* frame_info.qp                                                                */
  int32_t num_ref, interp_ref, num_intra_modes, frame_num;
  /* enc_params the block layer reads */
  int32_t encoder_speed, enable_bipred, enable_tb_split, enable_pb_split, intra_rdo, max_delta_qp, delta_qp_step,
          use_block_contexts;
  float   early_skip_thr;
  int32_t reserved0;
  double  lambda;                    /* frame_info.lambda (enc/encode_frame.c:94)                                    */
  thor_enc_pic orig, rec;            /* rec: written by the superblocks (unfiltered reconstruction)                  */
  thor_enc_pic ref[THOR_ENC_MAX_REF];/* picture of ref_idx r: ref_array[r] >= 0 ? ref[ref_array[r]] : interp_frames[0] */
  thor_enc_unit *units;              /* (height/4) * (width/4), zeroed by the caller before the frame               */
  thor_enc_node *nodes;              /* nsb * THOR_ENC_NODES                                                         */
  thor_enc_coef *coef;               /* nsb                                                                          */
  int32_t       *progress;           /* per superblock row: superblocks finished (wavefront hand-over), zeroed        */
} thor_enc_frame;

/* ------------------------------------------------------------------ entry point ---- */
/* The frame as the caller of the library describes it: pictures by handle (thorcuda.h), everything else as above. */
typedef struct {
  int32_t width, height, frame_type, qp, num_ref, interp_ref, num_intra_modes, frame_num;
  int32_t encoder_speed, enable_bipred, enable_tb_split, enable_pb_split, intra_rdo, max_delta_qp, delta_qp_step,
          use_block_contexts;
  float   early_skip_thr;
  int32_t reserved0;
  double  lambda;
  int32_t orig, rec;                    /* thorcuda_pic handles; rec is unpadded                                    */
  int32_t ref[THOR_ENC_MAX_REF];        /* padded reference pictures (96 / 48), borders valid                       */
  int32_t ref_frame_num[THOR_ENC_MAX_REF];  /* yuv_frame_t.frame_num of each: the sign of a vector depends on it    */
} thorcuda_enc_params;

/* Replaces the superblock loop of encode_frame (enc/encode_frame.c:112-148): every superblock's process_block
 * (enc/encode_block.c:2787-3033) with the searches, trial encodes and rate-distortion decisions it drives runs on the
 * device, superblock rows as a wavefront (a superblock starts when its left neighbour and the one above-right are done).
 * On return the work is QUEUED on the library stream:
 *   - picture `rec` holds the unfiltered reconstruction and carries the frame's side information (as after
 *     thorcuda_blockinfo_upload): thorcuda_loopfilter_frame can follow directly;
 *   - host_nodes (nsb * THOR_ENC_NODES) and host_coef (nsb), nsb = ceil(w/64) * ceil(h/64) superblocks in raster order,
 *     receive the decisions; they are valid after thorcuda_sync().  Page-locked memory (thorcuda_host_alloc) is DMA'd
 *     directly.
 * The reference's -sync 1 and -rdoq 1 (off in every shipped configuration) are not built: callers must refuse them
 * (integration/enc_gpu_hook.c does), there is no fallback. */
int thorcuda_enc_frame(const thorcuda_enc_params *p, thor_enc_node *host_nodes, thor_enc_coef *host_coef);

#ifdef __cplusplus
}
#endif
#endif
