/* Reference-side binding of libthorcuda's FRAME ENCODER for Thorenc.
 *
 * encode_frame() (enc/encode_frame.c:65-192) keeps its frame header, calls
 *     thorcuda_enc_superblocks(encoder_info)          instead of the superblock loop (:112-148)
 *     thorcuda_enc_tail(encoder_info, stream, qp)     instead of deblock_frame_y / deblock_frame_uv / clpf_frame
 *     thorcuda_enc_make_reference(ref0, rec)          instead of create_reference_frame
 * (integration/apply_hooks.py substitutes them in a scratch copy at build time).  Every decision of process_block
 * (enc/encode_block.c:2787-3033) -- searches, trial encodes, rate-distortion comparisons, the quad tree -- is taken on the
 * device; what comes back per superblock is the decided tree (thor_enc_node) and the quantised coefficients
 * (thor_enc_coef).  This file then runs the reference's OWN entropy coder over that tree, in the reference's order:
 * write_super_mode / write_delta_qp / write_block (enc/write_bits.c), nothing re-implemented.
 *
 * Built twice from this source (integration/Makefile):
 *   Thorenc_gpu   the product: backend = thorcuda_enc_frame on the B200, pictures resident in HBM, reference ring by handle
 *   Thorenc_emu   -DTHOR_EMU, TEST INFRASTRUCTURE: backend = tests/hostemu (the device source compiled for the host, one
 *                 thread), pictures are the encoder's own; lets tests check the decision logic byte for byte without a GPU
 * Compiled against the reference's own headers where they lie. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "global.h"
#include "mainenc.h"
#include "putbits.h"
#include "write_bits.h"
#include "thorcuda_enc.h"
#ifndef THOR_EMU
#include "thorcuda.h"
#endif

void copy_deblock_data(encoder_info_t *encoder_info, block_info_t *block_info);   /* enc/encode_block.c:1947 */

static double now_ms(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; }
static double t_dev, t_bits, t_tail, t_first, t_last;
static int t_frames, t_on = -1;
static void report(void) {
  fprintf(stderr, "thorcuda frame encoder: %d frames; per frame: decisions %.2f ms, entropy coding %.2f ms, tail %.2f ms\n", t_frames,
          t_dev / (t_frames ? t_frames : 1), t_bits / (t_frames ? t_frames : 1), t_tail / (t_frames ? t_frames : 1));
#ifndef THOR_EMU
  {
    int thorcuda_enc_profile(unsigned long long *out32);
    static const char *nm[] = {"superblock", "wait", "load_org", "early_skip", "pred_inter", "pred_intra", "tb_chain", "bits", "ssd", "keep_best", "search_intra",
                               "sad_batch", "fasthalf", "fastquarter", "motion_estimate", "finalize", "neighbours", "bipred_search", "mode_decision"};
    unsigned long long pf[32];
    if (thorcuda_enc_profile(pf) == 0 && pf[0]) {
      fprintf(stderr, "thorcuda frame encoder: Mcycles of CTA time by scope (nested scopes overlap):");
      for (int i = 0; i < 19; i++) fprintf(stderr, " %s %.1f;", nm[i], pf[i] * 1e-6);
      fprintf(stderr, "\n");
    }
  }
#endif
  if (t_frames > 2) fprintf(stderr, "thorcuda frame encoder: steady state, %d frames after the second in %.3f s = %.3f frames/s\n", t_frames - 2,
                            (t_last - t_first) * 1e-3, (t_frames - 2) / ((t_last - t_first) * 1e-3));
}

static thor_enc_node *g_nodes;
static thor_enc_coef *g_coef;
static int g_nsb;

/* node addressing: the same arithmetic as thor_b200/csrc/enc/enc_core.h (zidx8 / node_index) */
static int zidx8(int ypos, int xpos) {
  const int by = (ypos & 63) >> 3, bx = (xpos & 63) >> 3;
  int z = 0;
  for (int b = 0; b < 3; b++) z |= (((bx >> b) & 1) << (2 * b)) | (((by >> b) & 1) << (2 * b + 1));
  return z;
}
static int node_index(int size, int ypos, int xpos) {
  const int z = zidx8(ypos, xpos);
  return size == 64 ? 0 : size == 32 ? 1 + (z >> 4) : size == 16 ? 5 + (z >> 2) : 21 + z;
}

/* The decided tree of one superblock -> bits, in process_block's order (enc/encode_block.c:2928-2946: the split code, then
 * the children top-left, bottom-left, top-right, bottom-right) and deblock_data[] as copy_deblock_data leaves it. */
static block_info_t g_bi;             /* holds a block_param_t (12 KB of coefficient arrays): static, not on the stack */
static void emit_node(encoder_info_t *ei, const thor_enc_node *nodes, const thor_enc_coef *coef, int size, int ypos, int xpos) {
  const int width = ei->width, height = ei->height;
  stream_t *stream = ei->stream;
  const frame_type_t frame_type = ei->frame_info.frame_type;
  if (ypos >= height || xpos >= width) return;
  const int encode_this_size = ypos + size <= height && xpos + size <= width;
  const int encode_smaller_size = size > MIN_BLOCK_SIZE * (encode_this_size && frame_type != I_FRAME && !ei->params->sync && ei->params->encoder_speed > 0 ? 2 : 1);
  if (encode_this_size == 0 && encode_smaller_size == 0) return;
  const thor_enc_node *nd = &nodes[node_index(size, ypos, xpos)];
  block_context_t ctx;
  block_info_t *bi = &g_bi;
  ctx.split = -1; ctx.mode = -1; ctx.size = -1;
  ctx.index = nd->ctx_index;
  ctx.cbp = nd->ctx_cbp;
  bi->block_context = &ctx;
  bi->block_pos.size = size;
  bi->block_pos.ypos = ypos;
  bi->block_pos.xpos = xpos;
  bi->block_pos.bwidth = min(size, width - xpos);
  bi->block_pos.bheight = min(size, height - ypos);
  bi->max_num_tb_part = (ei->params->enable_tb_split == 1) ? 2 : 1;
  bi->max_num_pb_part = ei->params->enable_pb_split ? 4 : 1;
  bi->delta_qp = nd->delta_qp;
  if (nd->state == THOR_NODE_SPLIT) {
    if (encode_this_size) {
      bi->block_param.mode = MODE_SKIP;
      write_super_mode(stream, ei, bi, &bi->block_param, 1);
    } else if (frame_type != I_FRAME) {
      putbits(1, 0, stream);
    }
    if (size == MAX_BLOCK_SIZE && ei->params->max_delta_qp) write_delta_qp(stream, bi->delta_qp);
    const int ns = size / 2;
    emit_node(ei, nodes, coef, ns, ypos, xpos);
    emit_node(ei, nodes, coef, ns, ypos + ns, xpos);
    emit_node(ei, nodes, coef, ns, ypos, xpos + ns);
    emit_node(ei, nodes, coef, ns, ypos + ns, xpos + ns);
    return;
  }
  if (nd->state != THOR_NODE_CODED) fatalerror("thorcuda frame encoder: undecided node in a superblock tree");
  block_param_t *bp = &bi->block_param;
  bp->mode = (block_mode_t)nd->mode;
  bp->intra_mode = (intra_mode_t)nd->intra_mode;
  bp->skip_idx = nd->skip_idx;
  bp->pb_part = nd->pb_part;
  bp->ref_idx0 = nd->ref_idx0;
  bp->ref_idx1 = nd->ref_idx1;
  bp->dir = nd->dir;
  bp->tb_param = nd->tb_param;
  bp->tb_split = nd->tb_param > 0 ? nd->tb_param : 0;
  bp->cbp.y = nd->cbp_y; bp->cbp.u = nd->cbp_u; bp->cbp.v = nd->cbp_v;
  for (int i = 0; i < 4; i++) {
    bp->mv_arr0[i].x = nd->mv0[i][0]; bp->mv_arr0[i].y = nd->mv0[i][1];
    bp->mv_arr1[i].x = nd->mv1[i][0]; bp->mv_arr1[i].y = nd->mv1[i][1];
  }
  bi->num_skip_vec = nd->num_skip_vec;
  bi->num_merge_vec = nd->num_merge_vec;
  bi->mvp.x = nd->mvp[0]; bi->mvp.y = nd->mvp[1];
  if (bp->mode != MODE_SKIP) {
    const int z = zidx8(ypos, xpos);
    if (nd->cbp_y) memcpy(bp->coeff_y, coef->y + 64 * z, (size_t)size * size * sizeof(int16_t));
    if (nd->cbp_u) memcpy(bp->coeff_u, coef->u + 16 * z, (size_t)size * size / 4 * sizeof(int16_t));
    if (nd->cbp_v) memcpy(bp->coeff_v, coef->v + 16 * z, (size_t)size * size / 4 * sizeof(int16_t));
  }
  write_block(stream, ei, bi, bp);
  if (bp->tb_split) bp->cbp.y = bp->cbp.u = bp->cbp.v = 1;            /* enc/encode_block.c:1781-1784 */
  copy_deblock_data(ei, bi);
}

static void emit_frame(encoder_info_t *ei) {
  const int nsbx = (ei->width + MAX_BLOCK_SIZE - 1) / MAX_BLOCK_SIZE, nsby = (ei->height + MAX_BLOCK_SIZE - 1) / MAX_BLOCK_SIZE;
  for (int k = 0; k < nsby; k++)
    for (int l = 0; l < nsbx; l++)
      emit_node(ei, g_nodes + (size_t)(k * nsbx + l) * THOR_ENC_NODES, g_coef + (k * nsbx + l), MAX_BLOCK_SIZE, k * MAX_BLOCK_SIZE, l * MAX_BLOCK_SIZE);
}

static void check_config(const encoder_info_t *ei) {
  const enc_params *p = ei->params;
  if (p->sync) fatalerror("thorcuda frame encoder: -sync 1 is not built");
  if (p->rdoq) fatalerror("thorcuda frame encoder: -rdoq 1 is not built");
  if (ei->frame_info.num_ref > THOR_ENC_MAX_REF) fatalerror("thorcuda frame encoder: too many reference pictures");
}

#ifdef THOR_EMU
/* ---------------------------------------------------------------- host emulation backend (test infrastructure) ---- */
void thor_emu_encode_frame(const thor_enc_frame *F);
static thor_enc_unit *g_units;

static thor_enc_pic host_pic(const yuv_frame_t *f) {
  thor_enc_pic p;
  p.y = f->y; p.u = f->u; p.v = f->v;
  p.stride_y = f->stride_y; p.stride_c = f->stride_c;
  p.frame_num = f->frame_num; p.reserved = 0;
  return p;
}

void thorcuda_enc_superblocks(encoder_info_t *ei) {
  const int nsb = ((ei->width + 63) / 64) * ((ei->height + 63) / 64);
  const size_t nunits = (size_t)(ei->height / 4) * (ei->width / 4);
  check_config(ei);
  if (!g_nodes) {
    g_nsb = nsb;
    g_nodes = (thor_enc_node *)calloc((size_t)nsb * THOR_ENC_NODES, sizeof(thor_enc_node));
    g_coef = (thor_enc_coef *)calloc(nsb, sizeof(thor_enc_coef));
    g_units = (thor_enc_unit *)calloc(nunits, sizeof(thor_enc_unit));
  }
  memset(g_units, 0, nunits * sizeof(thor_enc_unit));
  memset(g_nodes, 0, (size_t)nsb * THOR_ENC_NODES * sizeof(thor_enc_node));
  const enc_params *p = ei->params;
  const frame_info_t *fi = &ei->frame_info;
  thor_enc_frame F;
  memset(&F, 0, sizeof(F));
  F.width = ei->width; F.height = ei->height;
  F.frame_type = fi->frame_type; F.qp = fi->qp; F.num_ref = fi->num_ref; F.interp_ref = fi->interp_ref;
  F.num_intra_modes = fi->num_intra_modes; F.frame_num = fi->frame_num;
  F.encoder_speed = p->encoder_speed; F.enable_bipred = p->enable_bipred; F.enable_tb_split = p->enable_tb_split;
  F.enable_pb_split = p->enable_pb_split; F.intra_rdo = p->intra_rdo; F.max_delta_qp = p->max_delta_qp;
  F.delta_qp_step = p->delta_qp_step; F.use_block_contexts = p->use_block_contexts;
  F.early_skip_thr = p->early_skip_thr;
  F.lambda = fi->lambda;
  F.orig = host_pic(ei->orig);
  F.rec = host_pic(ei->rec);
  for (int r = 0; r < fi->num_ref; r++) F.ref[r] = host_pic(fi->ref_array[r] >= 0 ? ei->ref[fi->ref_array[r]] : ei->interp_frames[0]);
  F.units = g_units; F.nodes = g_nodes; F.coef = g_coef; F.progress = NULL;
  thor_emu_encode_frame(&F);
  emit_frame(ei);
}

#else
/* ------------------------------------------------------------------------------------------ device backend ---- */
static void need(int rc) { if (rc) fatalerror((char *)thorcuda_last_error()); }

static thorcuda_pic dev_rec = -1, dev_org = -1, dev_newref = -1, dev_interp = -1;
static int dev_timer = -1;            /* CUDA-event timer around a frame's device work (upload .. tail) */
static FILE *frame_log;               /* THORCUDA_HOOK_FRAMELOG: one line per frame: index, host clock (ms), device ms */
static double t_dev_events;
static uint8_t *sb_bits;
/* the encoder's reference ring by handle: the host pictures keep their identity (pointers rotate,
 * enc/encode_frame.c:168-177), each has a device twin created the first time it is seen */
static const yuv_frame_t *ring_host[MAX_REF_FRAMES + MAX_SKIP_FRAMES + 1];
static thorcuda_pic ring_dev[MAX_REF_FRAMES + MAX_SKIP_FRAMES + 1];
static int nring;
static int ring_slot(const yuv_frame_t *f) {
  for (int i = 0; i < nring; i++) if (ring_host[i] == f) return i;
  if (nring == (int)(sizeof(ring_dev) / sizeof(ring_dev[0]))) fatalerror("thorcuda frame encoder: reference ring overflow");
  ring_host[nring] = f;
  ring_dev[nring] = thorcuda_pic_create(f->width, f->height, PADDING_Y, PADDING_Y / 2);
  if (ring_dev[nring] < 0) fatalerror((char *)thorcuda_last_error());
  return nring++;
}

static const void *locked[80];
static int nlocked;
static void lock_frame(const yuv_frame_t *f) {
  for (int i = 0; i < nlocked; i++) if (locked[i] == f->y) return;
  if (nlocked == 80) return;
  locked[nlocked++] = f->y;
  uint8_t *y0 = f->y - f->pad_ver_y * f->stride_y - f->pad_hor_y, *u0 = f->u - f->pad_ver_c * f->stride_c - f->pad_hor_c;
  uint8_t *v0 = f->v - f->pad_ver_c * f->stride_c - f->pad_hor_c;
  const size_t csize = (size_t)(f->height / 2 + 2 * f->pad_ver_c) * f->stride_c;
  thorcuda_host_register(y0, (size_t)(f->height + 2 * f->pad_ver_y) * f->stride_y);
  thorcuda_host_register(u0, (size_t)(v0 - u0) + csize);          /* U and V share one allocation */
}

void thorcuda_enc_superblocks(encoder_info_t *ei) {
  const int nsb = ((ei->width + 63) / 64) * ((ei->height + 63) / 64);
  yuv_frame_t *rec = ei->rec, *org = ei->orig;
  check_config(ei);
  if (dev_rec < 0) {
    const char *dev = getenv("THORCUDA_DEVICE");
    need(thorcuda_init(dev ? atoi(dev) : 0));
    dev_rec = thorcuda_pic_create(rec->width, rec->height, 0, 0);
    dev_org = thorcuda_pic_create(rec->width, rec->height, 0, 0);
    dev_newref = thorcuda_pic_create(rec->width, rec->height, PADDING_Y, PADDING_Y / 2);
    if (dev_rec < 0 || dev_org < 0 || dev_newref < 0) fatalerror((char *)thorcuda_last_error());
    sb_bits = (uint8_t *)thorcuda_host_alloc((size_t)(rec->width / MAX_BLOCK_SIZE + 1) * (rec->height / MAX_BLOCK_SIZE + 1));
    g_nsb = nsb;
    g_nodes = (thor_enc_node *)thorcuda_host_alloc((size_t)nsb * THOR_ENC_NODES * sizeof(thor_enc_node));
    g_coef = (thor_enc_coef *)thorcuda_host_alloc((size_t)nsb * sizeof(thor_enc_coef));
    if (!sb_bits || !g_nodes || !g_coef) fatalerror("thorcuda frame encoder: pinned allocation failed");
  }
  /* registered after the CUDA runtime's own exit handler, so it runs before it */
  if (t_on < 0) { t_on = getenv("THORCUDA_HOOK_TIMING") != NULL; if (t_on) atexit(report); }
  const double t0 = now_ms();
  lock_frame(org); lock_frame(rec);
  dev_timer = thorcuda_timer_begin();
  need(thorcuda_pic_upload(dev_org, org->y, org->u, org->v, org->stride_y, org->stride_c));
  const enc_params *p = ei->params;
  const frame_info_t *fi = &ei->frame_info;
  thorcuda_enc_params P;
  memset(&P, 0, sizeof(P));
  P.width = ei->width; P.height = ei->height;
  P.frame_type = fi->frame_type; P.qp = fi->qp; P.num_ref = fi->num_ref; P.interp_ref = fi->interp_ref;
  P.num_intra_modes = fi->num_intra_modes; P.frame_num = fi->frame_num;
  P.encoder_speed = p->encoder_speed; P.enable_bipred = p->enable_bipred; P.enable_tb_split = p->enable_tb_split;
  P.enable_pb_split = p->enable_pb_split; P.intra_rdo = p->intra_rdo; P.max_delta_qp = p->max_delta_qp;
  P.delta_qp_step = p->delta_qp_step; P.use_block_contexts = p->use_block_contexts;
  P.early_skip_thr = p->early_skip_thr;
  P.lambda = fi->lambda;
  P.orig = dev_org; P.rec = dev_rec;
  for (int r = 0; r < fi->num_ref; r++) {
    const yuv_frame_t *f = fi->ref_array[r] >= 0 ? ei->ref[fi->ref_array[r]] : ei->interp_frames[0];
    P.ref[r] = fi->ref_array[r] >= 0 ? ring_dev[ring_slot(f)] : dev_interp;
    P.ref_frame_num[r] = f->frame_num;
  }
  need(thorcuda_enc_frame(&P, g_nodes, g_coef));
  need(thorcuda_sync());
  const double t1 = now_ms();
  emit_frame(ei);
  const double t2 = now_ms();
  t_dev += t1 - t0; t_bits += t2 - t1;
}

/* the frame tail on the device: the reconstruction and its side information are already there */
void thorcuda_enc_tail(encoder_info_t *ei, stream_t *stream, int qp) {
  yuv_frame_t *rec = ei->rec;
  const double t0 = now_ms();
  int stages = THORCUDA_LF_MAKE_REF;
  if (ei->params->deblocking) stages |= THORCUDA_LF_DEBLOCK_Y | THORCUDA_LF_DEBLOCK_UV;
  if (ei->params->clpf) stages |= THORCUDA_LF_CLPF;
  need(thorcuda_loopfilter_frame(dev_rec, dev_org, dev_newref, qp, stages, NULL, sb_bits));
  need(thorcuda_pic_download(dev_rec, rec->y, rec->u, rec->v, rec->stride_y, rec->stride_c));   /* -rf output, PSNR */
  if (dev_timer >= 0) thorcuda_timer_end(dev_timer);
  need(thorcuda_sync());
  if (dev_timer >= 0) { t_dev_events = thorcuda_timer_elapsed_ms(dev_timer); dev_timer = -1; }
  if (ei->params->clpf) {
    putbits(1, 1, stream);            /* frame-level enable */
    putbits(1, 0, stream);            /* "every superblock signals its own bit" */
    const int n = (rec->width / MAX_BLOCK_SIZE) * (rec->height / MAX_BLOCK_SIZE);
    for (int i = 0; i < n; i++)
      if (sb_bits[i] != 0xFF) putbits(1, sb_bits[i], stream);
  }
  t_tail += now_ms() - t0;
}

/* create_reference_frame: the picture the tail has just produced becomes the device twin of ref[0]; the twin it replaces is
 * the target of the next frame's tail.  Nothing is copied. */
void thorcuda_enc_make_reference(yuv_frame_t *ref, yuv_frame_t *rec) {
  ref->frame_num = rec->frame_num;
  const int s = ring_slot(ref);
  const thorcuda_pic old = ring_dev[s];
  ring_dev[s] = dev_newref;
  dev_newref = old;
  t_frames++;
  t_last = now_ms();
  if (t_frames == 2) t_first = t_last;
  if (!frame_log && getenv("THORCUDA_HOOK_FRAMELOG")) frame_log = fopen(getenv("THORCUDA_HOOK_FRAMELOG"), "w");
  if (frame_log) { fprintf(frame_log, "%d %.3f %.3f %llu\n", t_frames - 1, t_last, t_dev_events, (unsigned long long)thorcuda_launch_count()); fflush(frame_log); }
}

/* interpolate_frames + pad_yuv_frame (enc/mainenc.c:328-329, :385-386): thorcuda_interpolate_frames on the device twins of
 * the two references -- nothing crosses PCIe.  THORCUDA_ENC_HOST_INTERP=1 runs the reference's own interpolate_frames on
 * the host instead, fed with the two references fetched from the device (two padded pictures down, one up per B frame):
 * the cross-check used when an RA stream differs. */
void interpolate_frames(yuv_frame_t *new_frame, yuv_frame_t *ref0, yuv_frame_t *ref1, int ratio, int pos);
void pad_yuv_frame(yuv_frame_t *f);
static uint8_t *plane0(const yuv_frame_t *f, int pl) {
  return pl == 0 ? f->y - f->pad_ver_y * f->stride_y - f->pad_hor_y : (pl == 1 ? f->u : f->v) - f->pad_ver_c * f->stride_c - f->pad_hor_c;
}
void thorcuda_enc_interpolate(yuv_frame_t *dst, yuv_frame_t *ref0, yuv_frame_t *ref1, int ratio, int pos) {
  static int on_device = -1;
  if (on_device < 0) on_device = getenv("THORCUDA_ENC_HOST_INTERP") == NULL;
  if (dev_interp < 0) {
    dev_interp = thorcuda_pic_create(dst->width, dst->height, PADDING_Y, PADDING_Y / 2);
    if (dev_interp < 0) fatalerror((char *)thorcuda_last_error());
  }
  if (on_device) {
    need(thorcuda_interpolate_frames(dev_interp, ring_dev[ring_slot(ref0)], ring_dev[ring_slot(ref1)], ratio, pos));
    need(thorcuda_pad_frame(dev_interp));
    return;
  }
  yuv_frame_t *fr[2] = {ref0, ref1};
  for (int k = 0; k < 2; k++) {
    lock_frame(fr[k]);
    need(thorcuda_pic_download_padded(ring_dev[ring_slot(fr[k])], plane0(fr[k], 0), plane0(fr[k], 1), plane0(fr[k], 2)));
  }
  need(thorcuda_sync());
  interpolate_frames(dst, ref0, ref1, ratio, pos);
  pad_yuv_frame(dst);
  lock_frame(dst);
  need(thorcuda_pic_upload_padded(dev_interp, plane0(dst, 0), plane0(dst, 1), plane0(dst, 2)));
}
#endif
