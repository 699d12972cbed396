/* Reference-side binding of libthorcuda's fused frame tail for the Thor DECODER.
 *
 * This file is what a maintainer of awakecoding/thor adds next to dec/decode_frame.c; it is compiled against the
 * reference's own headers (common/, dec/) where they lie and linked with the otherwise unmodified decoder objects
 * (integration/Makefile).  decode_frame() calls
 *     thorcuda_dec_tail(decoder_info, stream, qp)        instead of deblock_frame_y / deblock_frame_uv / clpf_frame
 *     thorcuda_dec_make_reference(ref0, rec)             instead of create_reference_frame
 * (integration/apply_hooks.py makes those two substitutions in a scratch copy at build time).
 *
 * The reconstructed frame and the side information go to the device once, deblocking + CLPF + reference creation
 * run as ONE fused launch + border replication, and the filtered frame and the padded reference come back, because
 * the rest of this decoder (prediction, output) still reads them on the host. */
#define _POSIX_C_SOURCE 199309L
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "global.h"
#include "maindec.h"
#include "getbits.h"
#include "thorcuda.h"

/* THORCUDA_HOOK_TIMING=1: host milliseconds per phase, printed at exit */
static double t_phase[6];
static int t_frames, t_on = -1;
static double now_ms(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return 1e3 * ts.tv_sec + 1e-6 * ts.tv_nsec;
}
static void report(void) {
  static const char *name[6] = {"lock+upload", "blockinfo", "clpf bits", "launch", "download rec", "download ref + sync"};
  for (int i = 0; i < 6; i++) fprintf(stderr, "thorcuda hook: %-20s %8.3f ms/frame\n", name[i], t_phase[i] / (t_frames ? t_frames : 1));
}
#define MARK(i) do { if (t_on) { const double t_ = now_ms(); t_phase[i] += t_ - t0; t0 = t_; } } while (0)

static thorcuda_pic dev_rec = -1, dev_ref = -1;
static int dev_w, dev_h;
static uint8_t *sb_bits, *sb_sig;

/* page-lock a frame's planes the first time it is seen (the decoder allocates its frames once, with malloc) */
static const void *locked[128];
static int nlocked;
static void lock_frame(const yuv_frame_t *f) {
  /* worth it for long streams only: page-locking a 4 MB frame costs about as much as 3 staged transfers of it, and
   * the decoder walks through 33 reference slots before it reuses one.  THORCUDA_HOOK_PIN=1 turns it on. */
  static int pin = -1;
  if (pin < 0) pin = getenv("THORCUDA_HOOK_PIN") != NULL;
  if (!pin) return;
  for (int i = 0; i < nlocked; i++) if (locked[i] == f->y) return;
  if (nlocked == 128) return;
  locked[nlocked++] = f->y;
  uint8_t *y0 = f->y - f->pad_ver_y * f->stride_y - f->pad_hor_y, *u0 = f->u - f->pad_ver_c * f->stride_c - f->pad_hor_c;
  uint8_t *v0 = f->v - f->pad_ver_c * f->stride_c - f->pad_hor_c;
  const size_t csize = (size_t)(f->height / 2 + 2 * f->pad_ver_c) * f->stride_c;
  /* U and V live in one allocation (create_yuv_frame): one range */
  thorcuda_host_register(y0, (size_t)(f->height + 2 * f->pad_ver_y) * f->stride_y);
  thorcuda_host_register(u0, (size_t)(v0 - u0) + csize);
}

static void need(int rc) {
  if (rc) fatalerror((char *)thorcuda_last_error());
}

static void setup(const yuv_frame_t *rec) {
  if (dev_rec >= 0 && dev_w == rec->width && dev_h == rec->height) return;
  const char *dev = getenv("THORCUDA_DEVICE");
  need(thorcuda_init(dev ? atoi(dev) : 0));
  dev_w = rec->width; dev_h = rec->height;
  dev_rec = thorcuda_pic_create(dev_w, dev_h, 0, 0);
  dev_ref = thorcuda_pic_create(dev_w, dev_h, PADDING_Y, PADDING_Y / 2);
  if (dev_rec < 0 || dev_ref < 0) fatalerror((char *)thorcuda_last_error());
}

uint8_t *thorcuda_dec_out_compact;     /* set by dec_recon_hook.c: pinned w*h*3/2 target of the frame download */

/* blockinfo + CLPF bits + fused launch + download of the filtered frame, for a reconstruction that already sits in
 * device picture `rec_dev`; the new padded reference is written into `ref_out_dev` */
void thorcuda_dec_tail_device(decoder_info_t *di, stream_t *stream, int qp, thorcuda_pic rec_dev, thorcuda_pic ref_out_dev) {
  yuv_frame_t *rec = di->rec;
  double t0 = now_ms();
  if (!sb_bits) sb_bits = (uint8_t *)thorcuda_host_alloc((size_t)(rec->width / MAX_BLOCK_SIZE + 1) * (rec->height / MAX_BLOCK_SIZE + 1));
  if (!sb_sig) sb_sig = (uint8_t *)malloc((size_t)(rec->width / MAX_BLOCK_SIZE + 1) * (rec->height / MAX_BLOCK_SIZE + 1));
  need(thorcuda_blockinfo_upload_signals(rec_dev, di->deblock_data, sb_sig));   /* one pass over deblock_data for both */
  MARK(1);
  int stages = THORCUDA_LF_MAKE_REF;
  if (di->deblocking) stages |= THORCUDA_LF_DEBLOCK_Y | THORCUDA_LF_DEBLOCK_UV;
  /* the CLPF bits follow the last superblock in the stream: one frame-level enable, one "all superblocks" flag,
   * then one bit per signalling superblock in raster order */
  if (di->clpf && getbits(stream, 1)) {
    const int all = getbits(stream, 1);
    const int nsb = (rec->width / MAX_BLOCK_SIZE) * (rec->height / MAX_BLOCK_SIZE);
    for (int i = 0; i < nsb; i++) sb_bits[i] = sb_sig[i] ? (uint8_t)(all ? 1 : getbits(stream, 1)) : 0;
    stages |= THORCUDA_LF_CLPF;
  }
  MARK(2);
  need(thorcuda_loopfilter_frame(rec_dev, -1, ref_out_dev, qp, stages, sb_bits, NULL));
  MARK(3);
  /* the filtered frame is what gets written out and displayed; the pipelined decoder (dec_recon_hook.c) takes it in a
   * pinned compact buffer and copies it into `rec` once the frame is complete */
  if (thorcuda_dec_out_compact) {
    uint8_t *o = thorcuda_dec_out_compact;
    const size_t ny = (size_t)rec->width * rec->height;
    need(thorcuda_pic_download(rec_dev, o, o + ny, o + ny + ny / 4, rec->width, rec->width / 2));
  } else
    need(thorcuda_pic_download(rec_dev, rec->y, rec->u, rec->v, rec->stride_y, rec->stride_c));
  MARK(4);
}

void thorcuda_dec_tail(decoder_info_t *di, stream_t *stream, int qp) {
  yuv_frame_t *rec = di->rec;
  if (t_on < 0) { t_on = getenv("THORCUDA_HOOK_TIMING") != NULL; if (t_on) atexit(report); }
  setup(rec);
  double t0 = now_ms();
  lock_frame(rec);
  need(thorcuda_pic_upload(dev_rec, rec->y, rec->u, rec->v, rec->stride_y, rec->stride_c));
  MARK(0);
  thorcuda_dec_tail_device(di, stream, qp, dev_rec, dev_ref);
}

void thorcuda_dec_make_reference(yuv_frame_t *ref, yuv_frame_t *rec) {
  /* the padded reference was produced by the same launch; the host copy feeds this decoder's CPU prediction */
  ref->frame_num = rec->frame_num;
  double t0 = now_ms();
  lock_frame(ref);
  need(thorcuda_pic_download_padded(dev_ref, ref->y - ref->pad_ver_y * ref->stride_y - ref->pad_hor_y,
                                    ref->u - ref->pad_ver_c * ref->stride_c - ref->pad_hor_c,
                                    ref->v - ref->pad_ver_c * ref->stride_c - ref->pad_hor_c));
  need(thorcuda_sync());
  MARK(5);
  t_frames++;
}
