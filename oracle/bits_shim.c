/* TEST INFRASTRUCTURE ONLY -- calls the UNMODIFIED reference's write_block (enc/write_bits.c:364-650) on a block
 * described by a flat int32 record, into a real stream_t, and returns what write_block returns (the number of bits it
 * emitted).  Compiled by oracle/Makefile against the reference's own headers into oracle/_ref/libthorbits_ref.so and
 * linked to oracle/_ref/libthorref.so; nothing of the reference is copied here.  Pins orc_block_bits (orc_bits.c).
 * Record layout: ORC_BS_* in thor_oracle.h. */
#include <stdlib.h>
#include <string.h>
#include "global.h"
#include "mainenc.h"
#include "putbits.h"
#include "write_bits.h"
#include "thor_oracle.h"

int ref_write_block_bits(const int32_t *d, const int16_t *cy, const int16_t *cu, const int16_t *cv) {
  static uint8_t buf[1 << 16];
  stream_t stream;
  memset(&stream, 0, sizeof stream);
  stream.bitstream = buf; stream.bytesize = sizeof buf; stream.bitrest = 32;
  enc_params params;
  memset(&params, 0, sizeof params);
  params.enable_bipred = d[ORC_BS_ENABLE_BIPRED]; params.max_delta_qp = d[ORC_BS_MAX_DELTA_QP];
  encoder_info_t *E = calloc(1, sizeof *E);
  block_info_t *B = calloc(1, sizeof *B);
  block_param_t *P = calloc(1, sizeof *P);
  block_context_t ctx;
  memset(&ctx, 0, sizeof ctx);
  E->params = &params; E->stream = &stream;
  E->frame_info.frame_type = (frame_type_t)d[ORC_BS_FRAME_TYPE]; E->frame_info.num_ref = d[ORC_BS_NUM_REF];
  E->frame_info.interp_ref = d[ORC_BS_INTERP_REF]; E->frame_info.num_intra_modes = d[ORC_BS_NUM_INTRA_MODES];
  const int size = d[ORC_BS_SIZE];
  B->block_pos.size = size; B->block_context = &ctx;
  ctx.index = (int8_t)d[ORC_BS_CTX_INDEX]; ctx.cbp = (int8_t)d[ORC_BS_CTX_CBP];
  B->max_num_pb_part = d[ORC_BS_MAX_PB_PART]; B->max_num_tb_part = d[ORC_BS_MAX_TB_PART];
  B->num_skip_vec = d[ORC_BS_NUM_SKIP_VEC]; B->num_merge_vec = d[ORC_BS_NUM_MERGE_VEC]; B->delta_qp = d[ORC_BS_DELTA_QP];
  B->mvp.x = (int16_t)d[ORC_BS_MVP]; B->mvp.y = (int16_t)d[ORC_BS_MVP + 1];
  P->mode = (block_mode_t)d[ORC_BS_MODE]; P->intra_mode = (intra_mode_t)d[ORC_BS_INTRA_MODE]; P->pb_part = d[ORC_BS_PB_PART];
  P->tb_split = d[ORC_BS_TB_SPLIT]; P->skip_idx = d[ORC_BS_SKIP_IDX]; P->ref_idx0 = d[ORC_BS_REF_IDX0]; P->ref_idx1 = d[ORC_BS_REF_IDX1];
  P->cbp.y = d[ORC_BS_CBP_Y]; P->cbp.u = d[ORC_BS_CBP_U]; P->cbp.v = d[ORC_BS_CBP_V];
  for (int i = 0; i < 4; i++) {
    P->mv_arr0[i].x = (int16_t)d[ORC_BS_MV0 + 2 * i]; P->mv_arr0[i].y = (int16_t)d[ORC_BS_MV0 + 2 * i + 1];
    P->mv_arr1[i].x = (int16_t)d[ORC_BS_MV1 + 2 * i]; P->mv_arr1[i].y = (int16_t)d[ORC_BS_MV1 + 2 * i + 1];
  }
  if (cy) memcpy(P->coeff_y, cy, (size_t)size * size * 2);
  if (cu) memcpy(P->coeff_u, cu, (size_t)size * size / 2);
  if (cv) memcpy(P->coeff_v, cv, (size_t)size * size / 2);
  const int bits = write_block(&stream, E, B, P);
  free(E); free(B); free(P);
  return bits;
}
