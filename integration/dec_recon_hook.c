/* Reference-side binding of libthorcuda's batched reconstruction for the Thor DECODER.
 *
 * With this file the decoder keeps only bitstream parsing on the host.  decode_block()'s calls of
 *     get_inter_prediction_luma / _chroma            -> hk_luma / hk_chroma          (record one prediction task)
 *     decode_and_reconstruct_block_inter / _intra    -> hk_recon_inter / _intra      (record transform-block tasks,
 *                                                                                     copy the parsed coefficients)
 * are substituted in a scratch copy at build time (integration/apply_hooks.py); nothing is computed while the
 * superblocks are parsed.  After the last superblock thorcuda_dec_flush() runs the frame on the device:
 *     inter prediction of every block (one launch)  ->  dequantise + inverse transform + reconstruct of every coded
 *     inter block (one launch)  ->  intra blocks in wavefront levels (a block joins a level once the blocks owning its
 *     neighbour samples are done; inter blocks are level 0)  ->  the fused frame tail of dec_tail_hook.c.
 * Reference pictures stay on the device (one device picture per slot of the decoder's reference ring), and so do the
 * interpolated references of B frames (interpolate_frames + pad_yuv_frame run on the ring's device twins); only the
 * filtered frame is copied back, because this decoder writes its output on the host.
 * Compiled against the reference's own headers where they lie. */
#define _POSIX_C_SOURCE 199309L
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <pthread.h>
#include "global.h"
#include "maindec.h"
#include "getbits.h"
#include "common_frame.h"
#include "thorcuda.h"

void thorcuda_dec_tail_device(decoder_info_t *di, stream_t *stream, int qp, thorcuda_pic dev_rec, thorcuda_pic dev_ref_out);
extern uint8_t *thorcuda_dec_out_compact;

/* THORCUDA_HOOK_TIMING=1: host milliseconds per phase and frame, printed at exit */
static double tp[6], t_mark, t_hk[5];
#define HK_T0 const double hk0_ = tp_on > 0 ? ms_now() : 0
#define HK_T1(i) do { if (tp_on > 0 && tp_seen >= tp_skip) t_hk[i] += ms_now() - hk0_; } while (0)
static double dev_ms, t_steady0, t_finish;       /* wall clock when the 37th frame starts / when the output is complete */
static int tp_frames, tp_on = -1, tp_levels, tp_intra, tp_pred, tp_tbs;
static double ms_now(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return 1e3 * ts.tv_sec + 1e-6 * ts.tv_nsec;
}
static int tp_seen;            /* frames seen; the first tp_skip (CUDA start-up, first use of every reference slot) are not counted */
static int tp_skip = 36;       /* THORCUDA_HOOK_STEADY_FROM overrides (short 2160p streams) */
static void lap(int i) {
  if (tp_on <= 0) return;
  const double t = ms_now();
  if (tp_seen >= tp_skip) tp[i] += t - t_mark;
  t_mark = t;
}
static void tp_report(void) {
  static const char *name[6] = {"parse + record", "inter submit + writes", "intra levels", "tail", "wait for previous frame + copy", "between frames"};
  const int n = tp_frames ? tp_frames : 1;
  for (int i = 0; i < 6; i++) fprintf(stderr, "thorcuda recon hook: %-18s %8.3f ms/frame\n", name[i], tp[i] / n);
  if (t_finish > t_steady0 && t_steady0 > 0 && tp_frames)
    fprintf(stderr, "thorcuda recon hook: steady state %.3f ms/frame wall clock = %.1f frames/s (frame %d .. output complete, %d frames)\n",
            (t_finish - t_steady0) / tp_frames, 1e3 * tp_frames / (t_finish - t_steady0), tp_skip + 1, tp_frames);
  fprintf(stderr, "thorcuda recon hook: inside hk_luma %.3f, hk_recon_inter %.3f, hk_recon_intra %.3f ms/frame\n", t_hk[0] / n, t_hk[1] / n, t_hk[2] / n);
  fprintf(stderr, "thorcuda recon hook: inside thorcuda_dec_write_frame %.3f, thorcuda_dec_interpolate %.3f ms/frame (both between frames)\n", t_hk[3] / n, t_hk[4] / n);
  fprintf(stderr, "thorcuda recon hook: device time of a frame's reconstruction (CUDA events, without interpolate_frames) %.3f ms/frame\n", dev_ms / n);
  fprintf(stderr, "thorcuda recon hook: per frame %d prediction tasks, %d transform blocks, %d intra blocks in %d levels\n",
          tp_pred / n, tp_tbs / n, tp_intra / n, tp_levels / n);
}

static int dry = -1;      /* THORCUDA_HOOK_DRY=1: record nothing, run nothing (measures the host's parsing alone; output is garbage) */
static decoder_info_t *g_di;
static void need(int rc) {
  if (rc) fatalerror((char *)thorcuda_last_error());
}

/* ---- device twins of the decoder's pictures ---- */
#define MAX_TWINS 64
static struct { const yuv_frame_t *host; thorcuda_pic dev; } twins[MAX_TWINS];
static int ntwins;
static thorcuda_pic dev_rec = -1;

static thorcuda_pic twin_of(const yuv_frame_t *f) {
  for (int i = 0; i < ntwins; i++) if (twins[i].host == f) return twins[i].dev;
  if (ntwins == MAX_TWINS) fatalerror("thorcuda hook: too many reference pictures");
  const thorcuda_pic p = thorcuda_pic_create(f->width, f->height, f->pad_hor_y, f->pad_hor_c);
  if (p < 0) fatalerror((char *)thorcuda_last_error());
  twins[ntwins].host = f; twins[ntwins].dev = p;
  ntwins++;
  return p;
}

/* ---- per-frame task lists ---- */
typedef struct { uint8_t plane, size, qp, level_known; uint16_t x, y; int coef; /* index into pool, -1: all zero */
                 int intra; /* index into intra[], -1: inter */ } Tb;
static thorcuda_pred_task *pred; static int npred, cpred;
static Tb *tbs; static int ntbs, ctbs;
static thorcuda_intra_task *intra; static int nintra, cintra;
static int16_t *pool; static size_t npool, cpool;
static thorcuda_pic pics[96]; static const yuv_frame_t *pic_host[96]; static int npics;
static int last_luma = -1;      /* prediction task the next luma call may complete to a bi-prediction */

static void *grow(void *p, int *cap, int need_n, size_t elem) {
  if (need_n <= *cap) return p;
  *cap = need_n * 2 + 1024;
  p = realloc(p, (size_t)*cap * elem);
  if (!p) fatalerror("thorcuda hook: out of memory");
  return p;
}

void thorcuda_dec_frame_begin(decoder_info_t *di) {
  g_di = di;
  if (dry < 0) dry = getenv("THORCUDA_HOOK_DRY") != NULL;
  if (tp_on < 0) { tp_on = getenv("THORCUDA_HOOK_TIMING") != NULL; if (tp_on) { atexit(tp_report); t_mark = ms_now(); if (getenv("THORCUDA_HOOK_STEADY_FROM")) tp_skip = atoi(getenv("THORCUDA_HOOK_STEADY_FROM")); } }
  lap(5);
  if (tp_on > 0 && tp_seen == tp_skip) t_steady0 = ms_now();
  if (dry) return;
  if (dev_rec < 0) {
    const char *dev = getenv("THORCUDA_DEVICE");
    need(thorcuda_init(dev ? atoi(dev) : 0));
    dev_rec = thorcuda_pic_create(di->width, di->height, 0, 0);
    if (dev_rec < 0) fatalerror((char *)thorcuda_last_error());
  }
  npred = ntbs = nintra = 0; npool = 0; last_luma = -1;
  npics = 1; pics[0] = dev_rec; pic_host[0] = NULL;
}

static int pic_index(const yuv_frame_t *f) {
  for (int i = 1; i < npics; i++) if (pic_host[i] == f) return i;
  if (npics == 96) fatalerror("thorcuda hook: too many pictures in one frame");
  pics[npics] = twin_of(f); pic_host[npics] = f;
  return npics++;
}

/* which picture does a sample pointer of decode_block() point into, and where */
static const yuv_frame_t *locate(const uint8_t *p, int *x, int *y) {
  const yuv_frame_t *cand[MAX_REF_FRAMES + 1];
  int n = 0;
  for (int i = 0; i < MAX_REF_FRAMES; i++) if (g_di->ref[i]) cand[n++] = g_di->ref[i];
  if (g_di->interp_frames[0]) cand[n++] = g_di->interp_frames[0];
  for (int i = 0; i < n; i++) {
    const yuv_frame_t *f = cand[i];
    if (p >= f->y && p < f->y + (size_t)f->height * f->stride_y) {
      const size_t off = (size_t)(p - f->y);
      *y = (int)(off / f->stride_y); *x = (int)(off % f->stride_y);
      return f;
    }
  }
  fatalerror("thorcuda hook: prediction source is not a reference picture");
  return NULL;
}

void hk_luma(uint8_t *pblock, uint8_t *ref, int width, int height, int stride, int pstride, mv_t *mv, int sign, int bipred) {
  int x, y;
  if (dry) return;
  HK_T0;
  const yuv_frame_t *f = locate(ref, &x, &y);
  const int ri = pic_index(f);
  if (last_luma >= 0) {
    thorcuda_pred_task *t = &pred[last_luma];
    if (t->x == x && t->y == y && t->w == width && t->h == height && t->ref1 < 0) {     /* second half of a bi-prediction */
      t->ref1 = (int8_t)ri; t->mv1x = mv->x; t->mv1y = mv->y; t->sign1 = (uint8_t)sign;
      last_luma = -1;
      HK_T1(0);
      return;
    }
  }
  pred = (thorcuda_pred_task *)grow(pred, &cpred, npred + 1, sizeof(*pred));
  thorcuda_pred_task *t = &pred[npred];
  memset(t, 0, sizeof(*t));
  t->x = (uint16_t)x; t->y = (uint16_t)y; t->w = (uint8_t)width; t->h = (uint8_t)height;
  t->ref0 = (int8_t)ri; t->ref1 = -1; t->mv0x = mv->x; t->mv0y = mv->y; t->sign0 = (uint8_t)sign; t->dst = 0;
  last_luma = npred++;
  HK_T1(0);
}

void hk_chroma(uint8_t *pblock, uint8_t *ref, int width, int height, int stride, int pstride, mv_t *mv, int sign) {
  /* the prediction task covers all three planes with the luma vector, as the decoder does */
}

static void where_in_rec(const uint8_t *p, int *plane, int *x, int *y) {
  const yuv_frame_t *r = g_di->rec;
  const uint8_t *base[3] = {r->y, r->u, r->v};
  for (int pl = 0; pl < 3; pl++) {
    const int stride = pl ? r->stride_c : r->stride_y, h = pl ? r->height / 2 : r->height;
    if (p >= base[pl] && p < base[pl] + (size_t)h * stride) {
      const size_t off = (size_t)(p - base[pl]);
      *plane = pl; *y = (int)(off / stride); *x = (int)(off % stride);
      return;
    }
  }
  fatalerror("thorcuda hook: reconstruction target is not the current picture");
}

/* one transform block: its parsed levels (n x n array, low q x q corner used) go to the frame's pool */
static int add_tb(int plane, int x, int y, int n, int qp, const int16_t *coeffq, int intra_index) {
  const int q = n < 16 ? n : 16;
  int any = 0;
  for (int i = 0; i < q && !any; i++)
    for (int j = 0; j < q; j++) if (coeffq[i * n + j]) { any = 1; break; }
  if (!any && intra_index < 0) return -1;             /* inter block without coefficients: the prediction is the result */
  tbs = (Tb *)grow(tbs, &ctbs, ntbs + 1, sizeof(*tbs));
  Tb *t = &tbs[ntbs];
  t->plane = (uint8_t)plane; t->size = (uint8_t)n; t->qp = (uint8_t)qp; t->x = (uint16_t)x; t->y = (uint16_t)y;
  t->intra = intra_index; t->coef = -1;
  if (any) {
    if (npool + (size_t)q * q > cpool) {
      cpool = (npool + (size_t)q * q) * 2 + 65536;
      pool = (int16_t *)realloc(pool, cpool * sizeof(int16_t));
      if (!pool) fatalerror("thorcuda hook: out of memory");
    }
    t->coef = (int)npool;
    for (int i = 0; i < q; i++) memcpy(pool + npool + (size_t)i * q, coeffq + i * n, (size_t)q * sizeof(int16_t));
    npool += (size_t)q * q;
  }
  return ntbs++;
}

void hk_recon_inter(uint8_t *rec, int stride, int size, int qp, uint8_t *pblock, int16_t *coeffq, int tb_split) {
  int plane, x, y;
  if (dry) return;
  HK_T0;
  where_in_rec(rec, &plane, &x, &y);
  last_luma = -1;
  if (!tb_split) { add_tb(plane, x, y, size, qp, coeffq, -1); HK_T1(1); return; }
  const int s2 = size / 2;
  for (int i = 0; i < size; i += s2)
    for (int j = 0; j < size; j += s2)
      add_tb(plane, x + j, y + i, s2, qp, coeffq + (2 * (i / s2) + (j / s2)) * s2 * s2, -1);
  HK_T1(1);
}

void hk_recon_intra(uint8_t *rec, int stride, int size, int qp, uint8_t *pblock, int16_t *coeffq, int tb_split, int upright_available,
                    int downleft_available, intra_mode_t intra_mode, int ypos, int xpos, int width, int comp) {
  last_luma = -1;
  if (dry) return;
  HK_T0;
  const int s = tb_split ? size / 2 : size;
  for (int i = 0; i < size; i += s)
    for (int j = 0; j < size; j += s) {
      intra = (thorcuda_intra_task *)grow(intra, &cintra, nintra + 1, sizeof(*intra));
      thorcuda_intra_task *t = &intra[nintra];
      memset(t, 0, sizeof(*t));
      t->x = (uint16_t)(xpos + j); t->y = (uint16_t)(ypos + i); t->cb_x = (uint16_t)xpos; t->size = (uint8_t)s; t->plane = (uint8_t)comp;
      t->mode = (uint8_t)intra_mode; t->src = t->dst = 0;
      /* availability of the extra sample per sub-block: common/intra_prediction.c:106-110 */
      const int up = tb_split ? (j == 0 || (i == 0 && upright_available)) : upright_available;
      const int dl = tb_split ? (j == 0 && (i == 0 || downleft_available)) : downleft_available;
      t->flags = (uint8_t)((up ? 1 : 0) | (dl ? 2 : 0));
      add_tb(comp, xpos + j, ypos + i, s, qp, tb_split ? coeffq + (2 * (i / s) + (j / s)) * s * s : coeffq, nintra);
      nintra++;
    }
  HK_T1(2);
}

/* ---- frame execution ---- */
static thorcuda_tb_task *lt; static int clt;
static int16_t *lpool; static size_t clpool;
static int *order, *lvl; static int corder, clvl_cap;
static int32_t *lmap[3]; static size_t lmap_n[3];

static void run_tbs(const int *idx, int n) {            /* dequantise + inverse + reconstruct of tbs[idx[..]] in place */
  lt = (thorcuda_tb_task *)grow(lt, &clt, n, sizeof(*lt));
  size_t need_pool = 0;
  int m = 0;
  for (int k = 0; k < n; k++) {
    const Tb *t = &tbs[idx ? idx[k] : k];
    if (t->coef < 0) continue;                          /* intra block without coefficients: prediction already in place */
    const int q = t->size < 16 ? t->size : 16;
    need_pool += ((size_t)q * q + 7) & ~(size_t)7;
    m++;
  }
  if (!m) return;
  if (need_pool > clpool) { clpool = need_pool * 2; lpool = (int16_t *)realloc(lpool, clpool * sizeof(int16_t)); }
  size_t off = 0;
  m = 0;
  for (int k = 0; k < n; k++) {
    const Tb *t = &tbs[idx ? idx[k] : k];
    if (t->coef < 0) continue;
    const int q = t->size < 16 ? t->size : 16;
    thorcuda_tb_task *d = &lt[m++];
    memset(d, 0, sizeof(*d));
    d->x = t->x; d->y = t->y; d->size = t->size; d->plane = t->plane; d->qp = t->qp; d->org = d->pred = d->rec = 0;
    d->coeff_offset = (uint32_t)off;
    memcpy(lpool + off, pool + t->coef, (size_t)q * q * sizeof(int16_t));
    off += ((size_t)q * q + 7) & ~(size_t)7;
  }
  need(thorcuda_reconstruct_blocks(pics, 1, lt, m, lpool));
}

/* ---- one frame in flight ----
 * The device works on frame n while the host parses frame n+1: thorcuda_dec_flush(n) queues everything, including the
 * download of the filtered frame into a pinned compact buffer, and returns without waiting.  The frame is completed
 * (wait for the event behind its download) at the start of flush(n+1); maindec's write_yuv_frame calls go
 * through thorcuda_dec_write_frame, which holds back a picture that is still in flight (and everything behind it, the
 * output stays in order) until then.  THORCUDA_HOOK_SERIAL=1 completes every frame at once (the old behaviour). */
static struct { yuv_frame_t *rec; int timer; } inflight;      /* timer: CUDA events around the frame's device work */

/* the filtered picture of each of the decoder's output slots lives in a pinned compact (w*h*3/2) buffer: the device
 * writes it there, the output thread writes the file from there; the decoder's own rec picture is never filled in
 * (nothing reads it any more: prediction, filters and reference creation are on the device) */
static struct { const yuv_frame_t *f; uint8_t *pin; } outs[MAX_REORDER_BUFFER + 2];
static int nouts;
static uint8_t *pin_of(const yuv_frame_t *f, int create) {
  for (int i = 0; i < nouts; i++) if (outs[i].f == f) return outs[i].pin;
  if (!create) return NULL;
  if (nouts == (int)(sizeof(outs) / sizeof(outs[0]))) fatalerror("thorcuda hook: more output pictures than the reorder buffer holds");
  outs[nouts].f = f;
  outs[nouts].pin = (uint8_t *)thorcuda_host_alloc((size_t)f->width * f->height * 3 / 2);
  if (!outs[nouts].pin) fatalerror("thorcuda hook: pinned allocation failed");
  return outs[nouts++].pin;
}
static void write_compact(const uint8_t *pin, int width, int height, FILE *file) {
  const size_t n = (size_t)width * height * 3 / 2;
  if (fwrite(pin, 1, n, file) != n) fatalerror("Error writing the decoded picture");
}
static struct { yuv_frame_t *f; int w, h; FILE *file; } wq[2 * MAX_REORDER_BUFFER + 4];
static int nwq, serial = -1;

static void output_wait_for(const yuv_frame_t *f);
static void complete_inflight(void) {
  yuv_frame_t *r = inflight.rec;
  if (!r) return;
  /* waits for this frame's last event only: work queued after it (the next frame's interpolation) keeps running */
  const float ms = thorcuda_timer_elapsed_ms(inflight.timer);
  if (ms < 0) fatalerror("thorcuda hook: waiting for the frame failed");
  if (tp_seen >= tp_skip) dev_ms += ms;
  inflight.rec = NULL;
}

/* ---- output thread (asynchronous write-back) ----
 * write_yuv_frame (2160 fwrite calls for a 1080p picture, 1.3 ms) runs on its own thread, in request order; the
 * decoder only waits for it when it is about to overwrite a picture that has not been written yet. */
#define OQ 64
static struct { yuv_frame_t *f; const uint8_t *pin; int w, h; FILE *file; } oq[OQ];
static unsigned oq_head, oq_tail;                   /* tail: next to write, head: next free; both only grow */
static pthread_mutex_t oq_mu = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t oq_cv = PTHREAD_COND_INITIALIZER;
static pthread_t oq_thread;
static int oq_started, oq_stop, sync_out = -1;

static void *writer_main(void *arg) {
  (void)arg;
  pthread_mutex_lock(&oq_mu);
  for (;;) {
    while (oq_tail == oq_head && !oq_stop) pthread_cond_wait(&oq_cv, &oq_mu);
    if (oq_tail == oq_head) break;
    const unsigned k = oq_tail % OQ;
    pthread_mutex_unlock(&oq_mu);
    if (oq[k].pin) write_compact(oq[k].pin, oq[k].w, oq[k].h, oq[k].file);
    else write_yuv_frame(oq[k].f, oq[k].w, oq[k].h, oq[k].file);
    pthread_mutex_lock(&oq_mu);
    oq_tail++;
    pthread_cond_broadcast(&oq_cv);
  }
  pthread_mutex_unlock(&oq_mu);
  return NULL;
}

static void output_frame(yuv_frame_t *f, int w, int h, FILE *file) {
  if (sync_out < 0) sync_out = getenv("THORCUDA_HOOK_SYNC_OUTPUT") != NULL;
  const uint8_t *pin = pin_of(f, 0);         /* NULL only in a dry run: the decoder's own (unfilled) picture is written */
  if (!file) return;                         /* write_yuv_frame is not called without an output file either (its fwrite would fault) */
  if (sync_out) { if (pin) write_compact(pin, w, h, file); else write_yuv_frame(f, w, h, file); return; }
  pthread_mutex_lock(&oq_mu);
  if (!oq_started) {
    if (pthread_create(&oq_thread, NULL, writer_main, NULL)) fatalerror("thorcuda hook: cannot start the output thread");
    oq_started = 1;
  }
  while (oq_head - oq_tail == OQ) pthread_cond_wait(&oq_cv, &oq_mu);
  const unsigned k = oq_head % OQ;
  oq[k].f = f; oq[k].pin = pin; oq[k].w = w; oq[k].h = h; oq[k].file = file;
  oq_head++;
  pthread_cond_broadcast(&oq_cv);
  pthread_mutex_unlock(&oq_mu);
}

/* before picture f is overwritten: every write request for it must be done */
static void output_wait_for(const yuv_frame_t *f) {
  if (!oq_started) return;
  pthread_mutex_lock(&oq_mu);
  for (;;) {
    int pending = 0;
    for (unsigned i = oq_tail; i != oq_head; i++) if (oq[i % OQ].f == f) pending = 1;
    if (!pending) break;
    pthread_cond_wait(&oq_cv, &oq_mu);
  }
  pthread_mutex_unlock(&oq_mu);
}

static void output_finish(void) {
  if (!oq_started) return;
  pthread_mutex_lock(&oq_mu);
  oq_stop = 1;
  pthread_cond_broadcast(&oq_cv);
  pthread_mutex_unlock(&oq_mu);
  pthread_join(oq_thread, NULL);
  oq_started = 0; oq_stop = 0;
}

static void drain_writes(int all) {
  int k = 0;
  for (; k < nwq; k++) {
    if (wq[k].f == inflight.rec) {
      if (!all) break;
      complete_inflight();
    }
    output_frame(wq[k].f, wq[k].w, wq[k].h, wq[k].file);
  }
  memmove(wq, wq + k, (size_t)(nwq - k) * sizeof(wq[0]));
  nwq -= k;
}

/* maindec.c's write_yuv_frame(&rec[i], width, height, outfile) */
void thorcuda_dec_write_frame(yuv_frame_t *frame, int width, int height, FILE *outfile) {
  HK_T0;
  if (nwq == 0 && frame != inflight.rec) { output_frame(frame, width, height, outfile); HK_T1(3); return; }
  if (nwq == (int)(sizeof(wq) / sizeof(wq[0]))) drain_writes(1);
  wq[nwq].f = frame; wq[nwq].w = width; wq[nwq].h = height; wq[nwq].file = outfile;
  nwq++;
  HK_T1(3);
}

/* before maindec.c prints its statistics and closes the output */
void thorcuda_dec_finish(void) {
  drain_writes(1); complete_inflight(); output_finish();
  if (tp_on > 0) t_finish = ms_now();            /* every picture is in the output file */
}

void thorcuda_dec_flush(decoder_info_t *di, stream_t *stream, int qp) {
  const yuv_frame_t *rec = di->rec;
  lap(0);
  if (!dry) { complete_inflight(); lap(4); }
  const int frame_timer = dry ? -1 : thorcuda_timer_begin();
  if (dry) { if (di->clpf && getbits(stream, 1)) fatalerror("dry run needs a stream without CLPF"); return; }
  if (tp_seen >= tp_skip) { tp_pred += npred; tp_tbs += ntbs; tp_intra += nintra; }
  /* 1. every inter block: prediction straight into the picture, then the coded ones are reconstructed in place */
  if (npred) need(thorcuda_predict_blocks(pics, npics, pred, npred, di->bipred, 3));
  int ninter = 0;
  order = (int *)grow(order, &corder, ntbs + 1, sizeof(int));
  for (int i = 0; i < ntbs; i++) if (tbs[i].intra < 0) order[ninter++] = i;
  if (ninter) run_tbs(order, ninter);
  lap(1);
  /* 2. intra blocks by wavefront level */
  if (nintra) {
    int maxl = 0;
    lvl = (int *)grow(lvl, &clvl_cap, ntbs + 1, sizeof(int));
    for (int pl = 0; pl < 3; pl++) {
      const size_t n = (size_t)((pl ? rec->width / 2 : rec->width) / 4 + 1) * ((pl ? rec->height / 2 : rec->height) / 4 + 1);
      if (n > lmap_n[pl]) { lmap[pl] = (int32_t *)realloc(lmap[pl], n * sizeof(int32_t)); lmap_n[pl] = n; }
      memset(lmap[pl], 0, n * sizeof(int32_t));
    }
    for (int i = 0; i < ntbs; i++) {
      const Tb *t = &tbs[i];
      lvl[i] = 0;
      if (t->intra < 0) continue;                       /* inter blocks are level 0 and their map entries stay 0 */
      const thorcuda_intra_task *it = &intra[t->intra];
      const int pl = t->plane, W = ((pl ? rec->width / 2 : rec->width) / 4 + 1), u = t->size / 4, ux = t->x / 4, uy = t->y / 4;
      const int32_t *M = lmap[pl];
      int dep = 0;
      if (t->y > 0) {
        const int x1 = ux + u + ((it->flags & 1) ? 1 : 0);
        for (int x = ux > 0 ? ux - 1 : 0; x < x1; x++) if (M[(uy - 1) * W + x] > dep) dep = M[(uy - 1) * W + x];
      }
      if (t->x > 0) {
        const int y1 = uy + u + ((it->flags & 2) ? 1 : 0);
        for (int y = uy > 0 ? uy - 1 : 0; y < y1; y++) if (M[y * W + ux - 1] > dep) dep = M[y * W + ux - 1];
      }
      lvl[i] = dep + 1;
      if (lvl[i] > maxl) maxl = lvl[i];
      for (int y = uy; y < uy + u; y++)
        for (int x = ux; x < ux + u; x++) lmap[pl][y * W + x] = lvl[i];
    }
    /* counting sort of the intra blocks by level */
    int *count = (int *)calloc((size_t)maxl + 2, sizeof(int));
    for (int i = 0; i < ntbs; i++) if (lvl[i]) count[lvl[i] + 1]++;
    for (int l = 1; l <= maxl + 1; l++) count[l] += count[l - 1];
    int *sorted = (int *)malloc((size_t)(nintra + 1) * sizeof(int)), *fill = (int *)malloc((size_t)(maxl + 2) * sizeof(int));
    memcpy(fill, count, (size_t)(maxl + 2) * sizeof(int));
    for (int i = 0; i < ntbs; i++) if (lvl[i]) sorted[fill[lvl[i]]++] = i;
    /* all levels in ONE call: predictions and coded transform blocks in level order, two launches per level */
    thorcuda_intra_task *batch = (thorcuda_intra_task *)malloc((size_t)(nintra + 1) * sizeof(*batch));
    int *ifirst = (int *)malloc((size_t)(maxl + 2) * sizeof(int)), *tfirst = (int *)malloc((size_t)(maxl + 2) * sizeof(int));
    if (tp_seen >= tp_skip) tp_levels += maxl;
    lt = (thorcuda_tb_task *)grow(lt, &clt, nintra + 1, sizeof(*lt));
    size_t need_pool = 0;
    for (int k = 0; k < nintra; k++) {
      const Tb *t = &tbs[sorted[k]];
      if (t->coef >= 0) { const int q = t->size < 16 ? t->size : 16; need_pool += ((size_t)q * q + 7) & ~(size_t)7; }
    }
    if (need_pool > clpool) { clpool = need_pool * 2; lpool = (int16_t *)realloc(lpool, clpool * sizeof(int16_t)); }
    size_t off = 0;
    int m = 0;
    for (int l = 1; l <= maxl; l++) {
      ifirst[l - 1] = count[l]; tfirst[l - 1] = m;
      for (int k = count[l]; k < count[l + 1]; k++) {
        const Tb *t = &tbs[sorted[k]];
        batch[k] = intra[t->intra];
        if (t->coef < 0) continue;                      /* no coefficients: the prediction is the reconstruction */
        const int q = t->size < 16 ? t->size : 16;
        thorcuda_tb_task *d = &lt[m++];
        memset(d, 0, sizeof(*d));
        d->x = t->x; d->y = t->y; d->size = t->size; d->plane = t->plane; d->qp = t->qp; d->org = d->pred = d->rec = 0;
        d->coeff_offset = (uint32_t)off;
        memcpy(lpool + off, pool + t->coef, (size_t)q * q * sizeof(int16_t));
        off += ((size_t)q * q + 7) & ~(size_t)7;
      }
    }
    ifirst[maxl] = nintra; tfirst[maxl] = m;
    need(thorcuda_intra_reconstruct_levels(pics, 1, batch, ifirst, lt, tfirst, maxl, lpool));
    free(batch); free(ifirst); free(tfirst); free(fill); free(sorted); free(count);
  }
  lap(2);
  /* 3. the frame tail, into the device twin of the ring slot the new reference will occupy */
  if (serial < 0) serial = getenv("THORCUDA_HOOK_SERIAL") != NULL;
  output_wait_for(di->rec);                  /* the slot's previous picture (32 frames ago) must have left for the file */
  thorcuda_dec_out_compact = pin_of(di->rec, 1);
  thorcuda_dec_tail_device(di, stream, qp, dev_rec, twin_of(di->ref[MAX_REF_FRAMES - 1]));
  need(thorcuda_timer_end(frame_timer));
  inflight.rec = di->rec;
  inflight.timer = frame_timer;
  lap(3);
  if (serial) complete_inflight();
  drain_writes(0);               /* pictures completed earlier go to the output file while the device works */
  lap(1);
}

/* interpolate_frames + pad_yuv_frame (dec/decode_frame.c:92-109) on the device twins: nothing travels */
void thorcuda_dec_interpolate(yuv_frame_t *dst, yuv_frame_t *ref0, yuv_frame_t *ref1, int ratio, int pos) {
  if (dry < 0) dry = getenv("THORCUDA_HOOK_DRY") != NULL;
  if (dry) return;
  HK_T0;
  need(thorcuda_interpolate_frames(twin_of(dst), twin_of(ref0), twin_of(ref1), ratio, pos));
  need(thorcuda_pad_frame(twin_of(dst)));
  HK_T1(4);
}

/* create_reference_frame: the padded reference already sits in the slot's device twin, which is all the decoder
 * reads from now on (prediction and interpolation are on the device); nothing to wait for here */
void thorcuda_dec_fetch_reference(yuv_frame_t *ref, yuv_frame_t *rec) {
  ref->frame_num = rec->frame_num;
  if (tp_seen++ >= tp_skip) tp_frames++;
}
