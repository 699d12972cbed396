/* Reference-side binding of libthorcuda's fused frame tail for the Thor ENCODER.
 *
 * encode_frame() calls
 *     thorcuda_enc_tail(encoder_info, stream, qp)     instead of deblock_frame_y / deblock_frame_uv / clpf_frame
 *     thorcuda_enc_make_reference(ref0, rec)          instead of create_reference_frame
 * (integration/apply_hooks.py substitutes them in a scratch copy at build time).  The encoder-side CLPF decision
 * -- per superblock, SSD(filtered) < SSD(unfiltered) over the coded non-bi-predicted 8x8 luma blocks,
 * enc/encode_frame.c:50-63 -- is taken on the device against the original frame; the hook writes the resulting bits
 * into the stream in the order clpf_frame's callback would have (raster order of the signalling superblocks).
 * Compiled against the reference's own headers where they lie. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "global.h"
#include "mainenc.h"
#include "putbits.h"
#include "thorcuda.h"
#include <time.h>

/* THORCUDA_HOOK_TIMING=1: host milliseconds per phase, printed at exit */
static double t_ph[8], t_leave, t_host, t_mark2;
static int t_on = -1, t_frames;
static double now_ms(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; }
static void report(void) {
  static const char *nm[8] = {"pin", "upload rec+org", "blockinfo", "launch", "download rec", "sync", "clpf bits", "reference download+sync"};
  fprintf(stderr, "thorcuda encoder tail, %d frames, host ms per frame:", t_frames);
  for (int i = 0; i < 8; i++) fprintf(stderr, " %s %.3f;", nm[i], t_ph[i] / (t_frames ? t_frames : 1));
  fprintf(stderr, " host between hooks %.3f\n", t_host / (t_frames > 1 ? t_frames - 1 : 1));
  /* the same interval the marginal rate of two stock runs (2 and n frames) measures, but on this process's own clock:
   * CUDA start-up and tear-down (1.1-1.6 s, varying by tenths of a second) stay outside */
  if (t_frames > 2 && t_leave > t_mark2)
    fprintf(stderr, "thorcuda encoder tail: steady state, %d frames after the second in %.3f s = %.3f frames/s\n", t_frames - 2,
            (t_leave - t_mark2) * 1e-3, (t_frames - 2) / ((t_leave - t_mark2) * 1e-3));
}
#define PHASE(i) do { if (t_on) { const double t_ = now_ms(); t_ph[i] += t_ - t0_; t0_ = t_; } } while (0)

static thorcuda_pic dev_rec = -1, dev_org = -1, dev_ref = -1;
static uint8_t *sb_bits;

/* The encoder allocates a handful of frames once (create_yuv_frame, malloc) and reuses them for every picture: page-lock
 * each the first time it is seen, so that its transfers are DMA from / into the frame itself instead of staged copies
 * through pageable memory. */
static const void *locked[64];
static int nlocked;
static void lock_frame(const yuv_frame_t *f) {
  static int off = -1;
  if (off < 0) off = getenv("THORCUDA_HOOK_NOPIN") != NULL;
  if (off) return;
  for (int i = 0; i < nlocked; i++) if (locked[i] == f->y) return;
  if (nlocked == 64) return;
  locked[nlocked++] = f->y;
  uint8_t *y0 = f->y - f->pad_ver_y * f->stride_y - f->pad_hor_y, *u0 = f->u - f->pad_ver_c * f->stride_c - f->pad_hor_c;
  uint8_t *v0 = f->v - f->pad_ver_c * f->stride_c - f->pad_hor_c;
  const size_t csize = (size_t)(f->height / 2 + 2 * f->pad_ver_c) * f->stride_c;
  thorcuda_host_register(y0, (size_t)(f->height + 2 * f->pad_ver_y) * f->stride_y);
  thorcuda_host_register(u0, (size_t)(v0 - u0) + csize);          /* U and V share one allocation */
}

static void need(int rc) {
  if (rc) fatalerror((char *)thorcuda_last_error());
}

void thorcuda_enc_tail(encoder_info_t *ei, stream_t *stream, int qp) {
  yuv_frame_t *rec = ei->rec, *org = ei->orig;
  if (dev_rec < 0) {
    const char *dev = getenv("THORCUDA_DEVICE");
    need(thorcuda_init(dev ? atoi(dev) : 0));
    dev_rec = thorcuda_pic_create(rec->width, rec->height, 0, 0);
    dev_org = thorcuda_pic_create(rec->width, rec->height, 0, 0);
    dev_ref = thorcuda_pic_create(rec->width, rec->height, PADDING_Y, PADDING_Y / 2);
    if (dev_rec < 0 || dev_org < 0 || dev_ref < 0) fatalerror((char *)thorcuda_last_error());
    sb_bits = (uint8_t *)thorcuda_host_alloc((size_t)(rec->width / MAX_BLOCK_SIZE + 1) * (rec->height / MAX_BLOCK_SIZE + 1));
  }
  if (t_on < 0) { t_on = getenv("THORCUDA_HOOK_TIMING") != NULL; if (t_on) atexit(report); }
  double t0_ = t_on ? now_ms() : 0;
  if (t_on && t_leave > 0) t_host += t0_ - t_leave;
  t_frames++;
  lock_frame(rec); lock_frame(org);
  PHASE(0);
  need(thorcuda_pic_upload(dev_rec, rec->y, rec->u, rec->v, rec->stride_y, rec->stride_c));
  need(thorcuda_pic_upload(dev_org, org->y, org->u, org->v, org->stride_y, org->stride_c));
  PHASE(1);
  need(thorcuda_blockinfo_upload(dev_rec, ei->deblock_data));
  PHASE(2);
  int stages = THORCUDA_LF_MAKE_REF;
  if (ei->params->deblocking) stages |= THORCUDA_LF_DEBLOCK_Y | THORCUDA_LF_DEBLOCK_UV;
  if (ei->params->clpf) stages |= THORCUDA_LF_CLPF;
  need(thorcuda_loopfilter_frame(dev_rec, dev_org, dev_ref, qp, stages, NULL, sb_bits));
  PHASE(3);
  need(thorcuda_pic_download(dev_rec, rec->y, rec->u, rec->v, rec->stride_y, rec->stride_c));
  PHASE(4);
  need(thorcuda_sync());
  PHASE(5);
  if (ei->params->clpf) {
    putbits(1, 1, stream);            /* frame-level enable */
    putbits(1, 0, stream);            /* "every superblock signals its own bit" */
    const int n = (rec->width / MAX_BLOCK_SIZE) * (rec->height / MAX_BLOCK_SIZE);
    for (int i = 0; i < n; i++)
      if (sb_bits[i] != 0xFF) putbits(1, sb_bits[i], stream);
  }
  PHASE(6);
}

void thorcuda_enc_make_reference(yuv_frame_t *ref, yuv_frame_t *rec) {
  double t0_ = t_on > 0 ? now_ms() : 0;
  ref->frame_num = rec->frame_num;
  lock_frame(ref);
  PHASE(0);
  need(thorcuda_pic_download_padded(dev_ref, ref->y - ref->pad_ver_y * ref->stride_y - ref->pad_hor_y,
                                    ref->u - ref->pad_ver_c * ref->stride_c - ref->pad_hor_c,
                                    ref->v - ref->pad_ver_c * ref->stride_c - ref->pad_hor_c));
  need(thorcuda_sync());
  PHASE(7);
  if (t_on > 0) { t_leave = now_ms(); if (t_frames == 2) t_mark2 = t_leave; }
}
