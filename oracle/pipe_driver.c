/*
 * pipe_driver.c -- TEST / BASELINE INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Runs one frame of the pixel-pipeline job that bench.py gives to libthorcuda, on the CPU, with
 *   -DPIPE_REF : the UNMODIFIED reference's own functions (linked from oracle/_ref/libthorref.so,
 *                headers included from where they lie under the reference tree), or
 *   (default)  : the oracle's C restatement (oracle/libthor_oracle.so).
 * The job, its task structs and its outputs are exactly the ones of include/thorcuda.h, so the
 * outputs can be compared byte for byte with the GPU's and the run can be timed beside it.
 * Used only by bench.py (cpu_baseline / --impl reference) and tests/.
 *
 * Picture index convention of the task lists: 0 = original, 1 = ref0, 2 = ref1, 3 = work picture
 * (prediction -> reconstruction -> filtered reconstruction).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "../include/thorcuda.h"

#ifdef PIPE_REF
#include "global.h"
#include "types.h"
#include "simd.h"
#include "common_block.h"
#include "common_frame.h"
#include "common_kernels.h"
#include "inter_prediction.h"
#include "transform.h"
#include "enc_kernels.h"
/* encoder-side functions with external linkage (enc/encode_block.c); declared here because the
 * reference's encode_block.h does not export them */
unsigned int sad_calc(uint8_t *a, uint8_t *b, int astride, int bstride, int width, int height);
int ssd_calc(uint8_t *a, uint8_t *b, int astride, int bstride, int width, int height);
unsigned int sad_calc_fasthalf(const uint8_t *a, const uint8_t *b, int astride, int bstride, int width, int height, int *x, int *y);
unsigned int sad_calc_fastquarter(const uint8_t *o, const uint8_t *r, int os, int rs, int width, int height, int *x, int *y);
int quantize(int16_t *coeff, int16_t *coeffq, int qp, int size, int coeff_block_type, int rdoq);
void get_residual(int16_t *block, uint8_t *pblock, uint8_t *orig, int size, int orig_stride);
void detect_clpf(const uint8_t *rec, const uint8_t *org, int x0, int y0, int width, int height, int so, int stride, int *sum0, int *sum1);
extern const int chroma_qp[52];
#else
#include "thor_oracle.h"
#endif

typedef struct {
  int32_t w, h, qp, range, bipred_filter, stages;
  int32_t band_y0, band_y1;        /* stage 1 only: luma rows [band_y0, band_y1) (multiples of 64) so threads can share a frame */
  uint8_t *org[3];                 /* unpadded, strides w, w/2, 16-byte aligned */
  uint8_t *ref[2][3];              /* visible origins of the two padded references */
  int32_t ref_stride_y, ref_stride_c;
  uint8_t *rec[3];                 /* work picture, strides w, w/2 */
  uint8_t *newref[3];              /* visible origins of the padded output reference (96/48) */
  int32_t newref_stride_y, newref_stride_c;
  const thorcuda_pred_task *pred; int32_t npred;
  const thorcuda_tb_task *tb; int32_t ntb;
  const thorcuda_cost_task *cost; int32_t ncost;
  const void *blockinfo;           /* deblock_data_t[(h/4)*(w/4)] */
  uint16_t *t8[2], *t16[2];
  uint32_t *t32[2], *t64[2];
  thorcuda_cost_result *half, *quarter;
  int16_t *coeffq; uint8_t *cbp; uint32_t *ssd;
  uint8_t *clpf_bits;
} pipe_job;

enum { ST_SAD = 1, ST_SUBPEL = 2, ST_PRED = 4, ST_TB = 8, ST_FILTER = 16 };

static void *amalloc(size_t n) { void *p = NULL; if (posix_memalign(&p, 64, n)) return NULL; return p; }

/* ---- thin adapters so that the job code below reads the same for both builds ---- */
#ifdef PIPE_REF
#define SAD(a, b, as, bs, w, h) sad_calc((uint8_t *)(a), (uint8_t *)(b), as, bs, w, h)
#define SSD(a, b, as, bs, w, h) ssd_calc((uint8_t *)(a), (uint8_t *)(b), as, bs, w, h)
#define FASTHALF(a, b, as, bs, w, h, x, y) ((use_simd && (w) > 4) ? sad_calc_fasthalf_simd(a, b, as, bs, w, h, x, y) : sad_calc_fasthalf(a, b, as, bs, w, h, x, y))
#define FASTQUARTER(a, b, as, bs, w, h, x, y) ((use_simd && (w) > 4) ? sad_calc_fastquarter_simd(a, b, as, bs, w, h, x, y) : sad_calc_fastquarter(a, b, as, bs, w, h, x, y))
static void LUMA(uint8_t *p, const uint8_t *ref, int w, int h, int stride, int ps, int mvx, int mvy, int sign, int bipred) {
  mv_t mv; mv.x = (int16_t)mvx; mv.y = (int16_t)mvy;
  get_inter_prediction_luma(p, (uint8_t *)ref, w, h, stride, ps, &mv, sign, bipred);
}
static void CHROMA(uint8_t *p, const uint8_t *ref, int w, int h, int stride, int ps, int mvx, int mvy, int sign) {
  mv_t mv; mv.x = (int16_t)mvx; mv.y = (int16_t)mvy;
  get_inter_prediction_chroma(p, (uint8_t *)ref, w, h, stride, ps, &mv, sign);
}
#define RESIDUAL(b, p, o, n, os) get_residual(b, p, (uint8_t *)(o), n, os)
#define TRANSFORM(b, c, n, f) transform(b, c, n, f)
#define QUANTIZE(c, q, qp, n, t) quantize(c, q, qp, n, t, 0)
#define DEQUANTIZE(q, r, qp, n) dequantize(q, r, qp, n)
#define INVERSE(r, b, n) inverse_transform(r, b, n)
#define RECONSTRUCT(b, p, r, n, s) reconstruct_block(b, p, r, n, s)
#else
#define SAD(a, b, as, bs, w, h) orc_sad(a, b, as, bs, w, h)
#define SSD(a, b, as, bs, w, h) orc_ssd(a, b, as, bs, w, h)
#define FASTHALF orc_sad_fasthalf
#define FASTQUARTER orc_sad_fastquarter
#define LUMA orc_inter_luma
#define CHROMA orc_inter_chroma
#define RESIDUAL orc_get_residual
#define TRANSFORM orc_transform
#define QUANTIZE orc_quantize
#define DEQUANTIZE orc_dequantize
#define INVERSE orc_inverse_transform
#define RECONSTRUCT orc_reconstruct_block
#endif

/* ---- stage 1: whole-frame SAD tables, one sad_calc per (block, displacement), as the reference's search would ---- */
static void stage_sad(const pipe_job *j) {
  const int R = j->range, n = 2 * R, W = j->w, H = j->h;   /* displacements -R .. R-1 (include/thorcuda.h) */
  for (int r = 0; r < 2; r++) {
    const uint8_t *ref = j->ref[r][0];
    const int rs = j->ref_stride_y;
    for (int lv = 0; lv < 4; lv++) {
      const int bs = 8 << lv, nbx = W / bs, nby = H / bs;
      const int by0 = j->band_y0 / bs, by1 = (j->band_y1 + bs - 1) / bs < nby ? (j->band_y1 + bs - 1) / bs : nby;
      for (int by = by0; by < by1; by++)
        for (int bx = 0; bx < nbx; bx++) {
          const uint8_t *a = j->org[0] + (size_t)by * bs * W + bx * bs;
          const uint8_t *b0 = ref + (ptrdiff_t)by * bs * rs + bx * bs;
          const size_t base = ((size_t)by * nbx + bx) * n * n;
          for (int dy = -R; dy < R; dy++)
            for (int dx = -R; dx < R; dx++) {
              const unsigned s = SAD(a, b0 + (ptrdiff_t)dy * rs + dx, W, rs, bs, bs);
              const size_t e = base + (size_t)(dy + R) * n + dx + R;
              if (lv == 0) j->t8[r][e] = (uint16_t)s;
              else if (lv == 1) j->t16[r][e] = (uint16_t)s;
              else if (lv == 2) j->t32[r][e] = s;
              else j->t64[r][e] = s;
            }
        }
    }
  }
}

/* ---- stage 2: bilinear half- and quarter-pel costs ---- */
static void stage_subpel(const pipe_job *j) {
  for (int i = 0; i < j->ncost; i++) {
    const thorcuda_cost_task *t = &j->cost[i];
    const uint8_t *a = j->org[0] + (size_t)t->y * j->w + t->x;
    const uint8_t *b = j->ref[t->ref - 1][0] + (ptrdiff_t)(t->y + t->iy) * j->ref_stride_y + t->x + t->ix;
    int x = 0, y = 0;
    j->half[i].cost = FASTHALF(a, b, j->w, j->ref_stride_y, t->w, t->h, &x, &y);
    j->half[i].x = (int16_t)x; j->half[i].y = (int16_t)y;
    x = t->hx; y = t->hy;
    j->quarter[i].cost = FASTQUARTER(a, b, j->w, j->ref_stride_y, t->w, t->h, &x, &y);
    j->quarter[i].x = (int16_t)x; j->quarter[i].y = (int16_t)y;
  }
}

/* ---- stage 3: inter prediction of every block into the work picture ---- */
static void stage_pred(const pipe_job *j) {
  uint8_t *p0 = amalloc(64 * 64), *p1 = amalloc(64 * 64);
  for (int i = 0; i < j->npred; i++) {
    const thorcuda_pred_task *t = &j->pred[i];
    for (int pl = 0; pl < 3; pl++) {
      const int w = pl ? t->w / 2 : t->w, h = pl ? t->h / 2 : t->h, x = pl ? t->x / 2 : t->x, y = pl ? t->y / 2 : t->y;
      const int rs = pl ? j->ref_stride_c : j->ref_stride_y, ds = pl ? j->w / 2 : j->w;
      for (int k = 0; k < 2; k++) {
        const int rix = k ? t->ref1 : t->ref0;
        if (rix < 0) break;
        const uint8_t *ref = j->ref[rix - 1][pl] + (ptrdiff_t)y * rs + x;
        const int mvx = k ? t->mv1x : t->mv0x, mvy = k ? t->mv1y : t->mv0y, sg = k ? t->sign1 : t->sign0;
        if (pl == 0) LUMA(k ? p1 : p0, ref, w, h, rs, w, mvx, mvy, sg, j->bipred_filter);
        else CHROMA(k ? p1 : p0, ref, w, h, rs, w, mvx, mvy, sg);
      }
      uint8_t *dst = j->rec[pl] + (size_t)y * ds + x;
      if (t->ref1 >= 0) {   /* enc/encode_block.c:1569-1588: truncating average */
        for (int r = 0; r < h; r++) for (int c = 0; c < w; c++) dst[(size_t)r * ds + c] = (uint8_t)((p0[r * w + c] + p1[r * w + c]) >> 1);
      } else {
        for (int r = 0; r < h; r++) memcpy(dst + (size_t)r * ds, p0 + r * w, w);
      }
    }
  }
  free(p0); free(p1);
}

/* ---- stage 4: encode_and_reconstruct_block_inter for every transform block, in place ---- */
static void stage_tb(const pipe_job *j) {
  int16_t *block = amalloc(64 * 64 * 2), *coeff = amalloc(64 * 64 * 2), *cq = amalloc(64 * 64 * 2), *rc = amalloc(64 * 64 * 2),
          *rb = amalloc(64 * 64 * 2);
  uint8_t *pb = amalloc(64 * 64);
  for (int i = 0; i < j->ntb; i++) {
    const thorcuda_tb_task *t = &j->tb[i];
    const int n = t->size, q = n < 16 ? n : 16, pl = t->plane, s = pl ? j->w / 2 : j->w;
    uint8_t *rec = j->rec[pl] + (size_t)t->y * s + t->x;
    const uint8_t *org = j->org[pl] + (size_t)t->y * s + t->x;
    for (int r = 0; r < n; r++) memcpy(pb + r * n, rec + (size_t)r * s, n);
    memset(coeff, 0, (size_t)n * n * 2); memset(cq, 0, (size_t)n * n * 2);
    RESIDUAL(block, pb, org, n, s);
    TRANSFORM(block, coeff, n, t->fast);
    const int cbp = QUANTIZE(coeff, cq, t->qp, n, t->coeff_type);
    if (cbp) {
      DEQUANTIZE(cq, rc, t->qp, n);
      INVERSE(rc, rb, n);
      RECONSTRUCT(rb, pb, rec, n, s);
    }
    for (int r = 0; r < q; r++) memcpy(j->coeffq + t->coeff_offset + r * q, cq + r * n, q * 2);
    j->cbp[i] = (uint8_t)(cbp != 0);
    j->ssd[i] = (uint32_t)SSD(org, rec, s, s, n, n);
  }
  free(block); free(coeff); free(cq); free(rc); free(rb); free(pb);
}

/* ---- stage 5: deblock + CLPF (encoder decision) + reference creation ---- */
#ifdef PIPE_REF
typedef struct { uint8_t *bits; int nsbx; } clpf_ctx;
/* same decision as the reference's static clpf_decision (enc/encode_frame.c:50-63); the bit goes to an
 * array instead of the bitstream */
static int decision(int k, int l, yuv_frame_t *rec, yuv_frame_t *org, const deblock_data_t *dd, int block_size, void *stream) {
  clpf_ctx *c = (clpf_ctx *)stream;
  int sum0 = 0, sum1 = 0;
  for (int m = 0; m < 64 / block_size; m++)
    for (int n = 0; n < 64 / block_size; n++) {
      const int xpos = l * 64 + n * block_size, ypos = k * 64 + m * block_size;
      const int index = (ypos / 4) * (rec->width / 4) + (xpos / 4);
      if (dd[index].cbp.y && dd[index].mode != MODE_BIPRED)
        (use_simd ? detect_clpf_simd : detect_clpf)(rec->y, org->y, xpos, ypos, rec->width, rec->height, org->stride_y, rec->stride_y, &sum0, &sum1);
    }
  c->bits[k * c->nsbx + l] = (uint8_t)(sum1 < sum0);
  return sum1 < sum0;
}
static void frame_of(yuv_frame_t *f, uint8_t *const p[3], int w, int h, int sy, int sc, int pady, int padc) {
  memset(f, 0, sizeof(*f));
  f->y = p[0]; f->u = p[1]; f->v = p[2];
  f->width = w; f->height = h; f->stride_y = sy; f->stride_c = sc;
  f->pad_hor_y = f->pad_ver_y = pady; f->pad_hor_c = f->pad_ver_c = padc;
}
#endif

static void stage_filter(const pipe_job *j) {
  const int W = j->w, H = j->h, nsbx = W / 64, nsb = nsbx * (H / 64);
  memset(j->clpf_bits, 0xFF, nsb > 0 ? nsb : 1);
#ifdef PIPE_REF
  yuv_frame_t rec, org, nref;
  frame_of(&rec, j->rec, W, H, W, W / 2, 0, 0);
  frame_of(&org, j->org, W, H, W, W / 2, 0, 0);
  frame_of(&nref, j->newref, W, H, j->newref_stride_y, j->newref_stride_c, 96, 48);
  deblock_frame_y(&rec, (deblock_data_t *)j->blockinfo, W, H, (uint8_t)j->qp);
  deblock_frame_uv(&rec, (deblock_data_t *)j->blockinfo, W, H, (uint8_t)chroma_qp[j->qp]);
  clpf_ctx c = { j->clpf_bits, nsbx };
  clpf_frame(&rec, &org, (const deblock_data_t *)j->blockinfo, &c, decision);
  create_reference_frame(&nref, &rec);
#else
  const orc_blockinfo_t *bi = (const orc_blockinfo_t *)j->blockinfo;
  orc_deblock_frame_y(j->rec[0], W, bi, W, H, j->qp);
  orc_deblock_frame_uv(j->rec[1], j->rec[2], W / 2, bi, W, H, orc_chroma_qp(j->qp));
  orc_clpf_frame(j->rec[0], j->rec[1], j->rec[2], W, W / 2, j->org[0], W, bi, W, H, NULL, j->clpf_bits);
  for (int pl = 0; pl < 3; pl++) {
    const int w = pl ? W / 2 : W, h = pl ? H / 2 : H, s = pl ? j->newref_stride_c : j->newref_stride_y;
    for (int r = 0; r < h; r++) memcpy(j->newref[pl] + (size_t)r * s, j->rec[pl] + (size_t)r * w, w);
    orc_pad_plane(j->newref[pl], s, w, h, pl ? 48 : 96, pl ? 48 : 96);
  }
#endif
}

void pipe_init(void) {
#ifdef PIPE_REF
  init_use_simd();                  /* the reference's own run-time switch (common/simd.h:134-140) */
#endif
}

int pipe_run(const pipe_job *j) {
  if (j->stages & ST_SAD) stage_sad(j);
  if (j->stages & ST_SUBPEL) stage_subpel(j);
  if (j->stages & ST_PRED) stage_pred(j);
  if (j->stages & ST_TB) stage_tb(j);
  if (j->stages & ST_FILTER) stage_filter(j);
  return 0;
}

int pipe_kind(void) {
#ifdef PIPE_REF
  return 1;
#else
  return 0;
#endif
}
