/* TEST INFRASTRUCTURE ONLY -- CPU oracle of Thor's temporal interpolation: the picture that
 * interpolate_frames builds between two references and that B frames then use as reference 0.  Restated from
 * reference common/temporal_interp.c:61-1053 (its USE_CHROMA is 0: motion is estimated on luma only).  Pinned against
 * the compiled reference in tests/test_oracle_vs_ref.py::test_interpolate_frames.  Never linked into libthorcuda.
 *
 * Shape of the algorithm:
 *   - a luma pyramid of up to 4 levels, each half the size of the one below (2x2 mean with the reference's double
 *     rounding), padded by 32;
 *   - per level, coarse to fine, one motion field on an 8x8 grid.  16x16 blocks are visited in raster order; each
 *     either accepts the "skip" vector (a median-like pick among its up-right / left / up neighbours) when the four
 *     8x8 SADs under it are small, or searches: candidates {0, the coarser level's vector x2, up-right, left, up},
 *     each refined by a +-1 pel cross (coarsest level: +-8 pel halving down to 1), cost = SAD between the two
 *     references displaced by (mv0, mv1) + lambda * distance to the neighbours' vectors;
 *   - a second pass on the 8x8 grid picks, per 8x8 block, the best of its own and its four neighbours' vectors;
 *   - the output is the rounding average of the two displaced references (8x8 luma / 4x4 chroma blocks).
 * Vectors are in 1/8 pel and point into pic[1]; pic[0]'s vector is the one scaled by -wt1/wt0. */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "thor_oracle.h"

typedef struct { int16_t x, y; } imv;

typedef struct {                  /* one pyramid level */
  int w, h, bw, bh;               /* luma size; field size in 8x8 units (always even) */
  const uint8_t *y[2];            /* pic[0], pic[1] luma origin */
  int stride, pad;
  imv *mv0, *mv1;                 /* [bh*bw] */
  uint8_t *skipped;               /* the reference's bgmap */
  uint8_t *own[2];                /* allocations of the down-scaled levels */
} level_t;

typedef struct { int wt0, wt1; } weights;

/* coverage counters for the tests: clamped SADs, skipped 16x16 blocks, blend branches (both / one / clamped), cross rounds */
static long g_cnt[6];
void orc_interp_counters(long *out, int reset) {
  if (out) memcpy(out, g_cnt, sizeof(g_cnt));
  if (reset) memset(g_cnt, 0, sizeof(g_cnt));
}

static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

/* temporal_interp.c:66-91; the result goes back into an int16 field like the reference's mv_t */
static int scale_component(int v, int numer, int denom) {
  if (denom == 0) return 0;
  int prod = v * numer;
  if (denom < 0) { denom = -denom; prod = -prod; }
  return prod >= 0 ? (prod + denom / 2) / denom : -((-prod + denom / 2) / denom);
}
static imv scale_vec(imv v, int numer, int denom) {
  imv r;
  if (numer == denom) return v;
  if (numer == -denom) { r.x = (int16_t)-v.x; r.y = (int16_t)-v.y; return r; }
  r.x = (int16_t)scale_component(v.x, numer, denom);
  r.y = (int16_t)scale_component(v.y, numer, denom);
  return r;
}
static imv to_pic0(imv v1, weights W) { return scale_vec(v1, -W.wt1, W.wt0); }

static int pel(int v) { return (v + 4) >> 3; }     /* 1/8 pel -> nearest full pel (:392, :448) */

/* SAD of the size x size blocks at (x0,y0) in pic[0] and (x1,y1) in pic[1]; both inside picture + pad, else the
 * per-sample clamped form (:443-523) */
static uint32_t pair_sad(const level_t *L, int x0, int y0, int x1, int y1, int size) {
  const int pad = L->pad, wP = L->w + pad, hP = L->h + pad, s = L->stride;
  uint32_t sum = 0;
  const int inside = x0 >= -pad && x0 + size <= wP && y0 >= -pad && y0 + size <= hP &&
                     x1 >= -pad && x1 + size <= wP && y1 >= -pad && y1 + size <= hP;
  if (!inside) g_cnt[0]++;
  for (int i = 0; i < size; i++)
    for (int j = 0; j < size; j++) {
      int ax = j + x0, ay = i + y0, bx = j + x1, by = i + y1;
      if (!inside) {
        ax = imin(wP - 1, imax(-pad, ax)); bx = imin(wP - 1, imax(-pad, bx));
        ay = imin(hP - 1, imax(-pad, ay)); by = imin(hP - 1, imax(-pad, by));
      }
      sum += (uint32_t)abs(L->y[1][by * s + bx] - L->y[0][ay * s + ax]);
    }
  return sum;
}

static uint32_t vec_sad(const level_t *L, int xs, int ys, imv v0, imv v1, int size) {
  return pair_sad(L, xs + pel(v0.x), ys + pel(v0.y), xs + pel(v1.x), ys + pel(v1.y), size);
}

/* the vector with the smallest summed L1 distance to the others; ties -> the LAST one (:761-782) */
static imv central_vec(const imv *v, int n) {
  int best = 0, best_cost = 0x3fffffff;
  for (int j = 0; j < n; j++) {
    int c = 0;
    for (int i = 0; i < n; i++) c += abs(v[i].x - v[j].x) + abs(v[i].y - v[j].y);
    if (c <= best_cost) { best = j; best_cost = c; }
  }
  return v[best];
}

static int push_unique(imv *list, int n, imv v) {
  for (int i = 0; i < n; i++)
    if (list[i].x == v.x && list[i].y == v.y) return n;
  list[n] = v;
  return n + 1;
}

static int l1(imv a, imv b) { return abs(a.x - b.x) + abs(a.y - b.y); }

/* smoothness term against the already decided neighbours (:366-385); lambda is 3000 (750 when guided), >> 7 */
static int field_cost(const level_t *L, imv v, int bx, int by, int lambda) {
  const imv *m = L->mv1;
  const int bw = L->bw;
  int d = 0;
  if (bx == 0 && by == 0) d = 0;
  else if (by > 0 && bx > 0 && bx < bw - 2)
    d = l1(v, m[(by - 2) * bw + bx + 2]) + l1(v, m[(by - 2) * bw + bx]) + l1(v, m[(by - 2) * bw + bx - 2]) + l1(v, m[by * bw + bx - 2]);
  else if (by == 0) d = l1(v, m[bx - 2]);
  else if (bx == 0) d = l1(v, m[(by - 2) * bw + 2]) + l1(v, m[(by - 2) * bw]);
  return (d * lambda) >> 7;
}

/* the four 8x8 SADs under the skip vector must all stay <= 8*64, and all must lie inside picture + pad (:525-581) */
static int skip_ok(const level_t *L, int xs, int ys, imv v0, imv v1) {
  const int pad = L->pad, wP = L->w + pad, hP = L->h + pad;
  for (int p = ys; p < ys + 16; p += 8)
    for (int q = xs; q < xs + 16; q += 8) {
      const int x0 = q + pel(v0.x), x1 = q + pel(v1.x), y0 = p + pel(v0.y), y1 = p + pel(v1.y);
      if (!(x0 >= -pad && x0 + 8 <= wP && y0 >= -pad && y0 + 8 <= hP && x1 >= -pad && x1 + 8 <= wP && y1 >= -pad && y1 + 8 <= hP)) return 0;
      if (pair_sad(L, x0, y0, x1, y1, 8) > 8 * 64) return 0;
    }
  return 1;
}

/* :650-725.  Every candidate is refined on its own by a cross search (only if it is not already much worse than the
 * best so far); the cheapest refined candidate wins, the first one on ties. */
static void search_block(level_t *L, weights W, const imv *cand, int ncand, int guided, int bx, int by) {
  const int xs = bx * 8, ys = by * 8, lambda = guided ? 750 : 3000;
  uint32_t best_cost = 0x3fffffff;
  imv best = cand[0];
  for (int c = 0; c < ncand; c++) {
    imv cur = cand[c];
    uint32_t cost = (uint32_t)field_cost(L, cur, bx, by, lambda);
    cost += vec_sad(L, xs, ys, to_pic0(cur, W), cur, 16);
    if (((4 + c) * cost) / 8 < best_cost) {
      int shift = guided ? 3 : 6, budget = guided ? 8 : 64;
      while (shift >= 3 && budget > 0) {
        const int off = 1 << shift;
        const imv centre = cur;
        int better = 0;
        for (int k = 0; k < 4; k++) {
          imv t = centre;
          if (k == 0) t.x = (int16_t)(centre.x - off);
          else if (k == 1) t.x = (int16_t)(centre.x + off);
          else if (k == 2) t.y = (int16_t)(centre.y - off);
          else t.y = (int16_t)(centre.y + off);
          uint32_t tc = (uint32_t)field_cost(L, t, bx, by, lambda);
          tc += vec_sad(L, xs, ys, to_pic0(t, W), t, 16);
          if (tc < cost) { cost = tc; cur = t; better = 1; }
        }
        if (!better) shift--;
        g_cnt[5]++;
        budget -= 4;
      }
    }
    if (cost < best_cost) { best_cost = cost; best = cur; }
  }
  L->mv1[by * L->bw + bx] = best;
  L->mv0[by * L->bw + bx] = to_pic0(best, W);
}

/* :852-918; guide = the coarser level's field (already doubled), or NULL on the coarsest level */
static void estimate_level(level_t *L, weights W, const imv *guide) {
  const int bw = L->bw, bh = L->bh;
  if (!guide) {
    memset(L->mv0, 0, sizeof(imv) * bw * bh);
    memset(L->mv1, 0, sizeof(imv) * bw * bh);
  }
  memset(L->skipped, 0, (size_t)bw * bh);
  for (int by = 0; by < bh; by += 2)
    for (int bx = 0; bx < bw; bx += 2) {
      const int pos = by * bw + bx;
      imv nb[3], skipv = {0, 0};
      int n = 0;
      if (by > 0 && bx < bw - 2) nb[n++] = L->mv1[(by - 2) * bw + bx + 2];
      if (bx > 0) nb[n++] = L->mv1[by * bw + bx - 2];
      if (by > 0) nb[n++] = L->mv1[(by - 2) * bw + bx];
      if (n) skipv = central_vec(nb, n);
      const imv skipv0 = to_pic0(skipv, W);
      if (skip_ok(L, bx * 8, by * 8, skipv0, skipv)) {
        L->skipped[pos] = 1;
        g_cnt[1]++;
        L->mv1[pos] = skipv;
        L->mv0[pos] = skipv0;
      } else {
        imv cand[8];
        const imv zero = {0, 0};
        int nc = push_unique(cand, 0, zero);
        if (guide) nc = push_unique(cand, nc, guide[pos]);       /* same weights on every level: no rescaling (:313-316) */
        if (by > 0 && bx < bw - 2) nc = push_unique(cand, nc, L->mv1[(by - 2) * bw + bx + 2]);
        if (bx > 0) nc = push_unique(cand, nc, L->mv1[by * bw + bx - 2]);
        if (by > 0) nc = push_unique(cand, nc, L->mv1[(by - 2) * bw + bx]);
        search_block(L, W, cand, nc, guide != NULL, bx, by);
      }
      for (int q = 0; q < 2; q++)
        for (int p = 0; p < 2; p++) {
          L->mv0[pos + q * bw + p] = L->mv0[pos];
          L->mv1[pos + q * bw + p] = L->mv1[pos];
          L->skipped[pos + q * bw + p] = L->skipped[pos];
        }
    }

  /* per 8x8 block: the best of its own and its neighbours' vectors by plain 8x8 SAD (:288-301, :727-759, :898-917).
   * Neighbour distance is 2 on odd ROWS (in both directions), 1 on even rows. */
  imv *n0 = (imv *)malloc(sizeof(imv) * bw * bh), *n1 = (imv *)malloc(sizeof(imv) * bw * bh);
  for (int by = 0; by < bh; by++)
    for (int bx = 0; bx < bw; bx++) {
      const int pos = by * bw + bx, off = (by & 1) ? 2 : 1;
      imv cand[5];
      int nc = push_unique(cand, 0, L->mv1[pos]);
      if (by - off >= 0) nc = push_unique(cand, nc, L->mv1[(by - off) * bw + bx]);
      if (by + off < bh) nc = push_unique(cand, nc, L->mv1[(by + off) * bw + bx]);
      if (bx - off >= 0) nc = push_unique(cand, nc, L->mv1[pos - off]);
      if (bx + off < bw) nc = push_unique(cand, nc, L->mv1[pos + off]);
      n0[pos] = L->mv0[pos];
      n1[pos] = L->mv1[pos];
      if (nc > 1) {
        uint32_t best = 0x3fffffff;
        for (int c = 0; c < nc; c++) {
          const imv v0 = to_pic0(cand[c], W);
          const uint32_t s = vec_sad(L, bx * 8, by * 8, v0, cand[c], 8);
          if (s < best) { best = s; n1[pos] = cand[c]; n0[pos] = v0; }
        }
      }
    }
  memcpy(L->mv0, n0, sizeof(imv) * bw * bh);
  memcpy(L->mv1, n1, sizeof(imv) * bw * bh);
  free(n0);
  free(n1);
}

/* 2x2 mean with the reference's double rounding, then edge replication by `pad` (:151-185 luma, common_frame.c:405-462) */
static void halve_luma(const uint8_t *src, int sstride, uint8_t *dst, int dstride, int wo, int ho, int pad) {
  for (int i = 0; i < ho; i++)
    for (int j = 0; j < wo; j++) {
      const uint8_t *a = src + (2 * i) * sstride + 2 * j, *b = a + sstride;
      dst[i * dstride + j] = (uint8_t)((((a[0] + b[0] + 1) >> 1) + ((a[1] + b[1] + 1) >> 1)) >> 1);
    }
  for (int i = 0; i < ho; i++) {
    memset(dst + i * dstride - pad, dst[i * dstride], pad);
    memset(dst + i * dstride + wo, dst[i * dstride + wo - 1], pad);
  }
  for (int i = 1; i <= pad; i++) {
    memcpy(dst - i * dstride - pad, dst - pad, wo + 2 * pad);
    memcpy(dst + (ho - 1 + i) * dstride - pad, dst + (ho - 1) * dstride - pad, wo + 2 * pad);
  }
}

/* one block of the output plane (:387-441).  (wP, hP, pad) are the plane's size + half a block and half a block. */
static void blend_block(const uint8_t *r0, const uint8_t *r1, int rstride, uint8_t *out, int ostride, int xs, int ys, imv v0, imv v1,
                        int wP, int hP, int pad, int size, int vis_w, int vis_h) {
  const int x0 = xs + pel(v0.x), y0 = ys + pel(v0.y), x1 = xs + pel(v1.x), y1 = ys + pel(v1.y);
  const int in0 = x0 >= -pad && x0 + size <= wP && y0 >= -pad && y0 + size <= hP;
  const int in1 = x1 >= -pad && x1 + size <= wP && y1 >= -pad && y1 + size <= hP;
  g_cnt[in0 && in1 ? 2 : (in0 || in1) ? 3 : 4]++;
  for (int i = 0; i < size; i++)
    for (int j = 0; j < size; j++) {
      /* the reference also writes the part of the last block row / column that hangs over the picture into the
       * padding, which the caller's pad_yuv_frame then overwrites: the visible picture is what counts */
      if (xs + j >= vis_w || ys + i >= vis_h) continue;
      int v;
      if (in0 && in1) v = (r0[(y0 + i) * rstride + x0 + j] + r1[(y1 + i) * rstride + x1 + j] + 1) >> 1;
      else if (in1) v = r1[(y1 + i) * rstride + x1 + j];
      else if (in0) v = r0[(y0 + i) * rstride + x0 + j];
      else {
        const int ax = imin(wP - 1, imax(-pad, j + x0)), bx = imin(wP - 1, imax(-pad, j + x1));
        const int ay = imin(hP - 1, imax(-pad, i + y0)), by = imin(hP - 1, imax(-pad, i + y1));
        v = (r0[ay * rstride + ax] + r1[by * rstride + bx] + 1) >> 1;
      }
      out[(ys + i) * ostride + xs + j] = (uint8_t)v;
    }
}

int orc_interp_levels(int width, int height) {
  const int m = width < height ? width : height;
  const int l = (int)(log10((double)m) / log10(2.0) - 4.0);       /* :977, same expression */
  return l < 4 ? l : 4;
}

void orc_interp_field_size(int width, int height, int *bw, int *bh) {
  *bw = 2 * ((width + 15) / 16);                                  /* :97-99 */
  *bh = 2 * ((height + 15) / 16);
}

/* interpolate_frames (:972-1053).  ref0/ref1: padded pictures of equal geometry (pad >= 4, and as large as the motion
 * the caller expects to stay "inside").  out: the visible picture only (the caller pads it afterwards).
 * dbg_mv0 / dbg_mv1 (may be NULL): the final level-0 field, bw*bh vectors each, for stage-wise comparisons. */
void orc_interpolate_frames(uint8_t *out_y, uint8_t *out_u, uint8_t *out_v, int out_stride_y, int out_stride_c,
                            const uint8_t *ref0_y, const uint8_t *ref0_u, const uint8_t *ref0_v,
                            const uint8_t *ref1_y, const uint8_t *ref1_u, const uint8_t *ref1_v,
                            int width, int height, int stride_y, int stride_c, int pad_y, int ratio, int pos,
                            int16_t *dbg_mv0, int16_t *dbg_mv1) {
  const int nlev = orc_interp_levels(width, height);
  if (nlev < 1) return;
  weights W;
  const int reversed = pos > ratio / 2;                           /* :120-126 */
  W.wt0 = reversed ? pos : ratio - pos;
  W.wt1 = ratio - W.wt0;
  level_t L[4];
  memset(L, 0, sizeof(L));
  for (int l = 0; l < nlev; l++) {
    level_t *q = &L[l];
    q->w = width >> l;
    q->h = height >> l;
    orc_interp_field_size(q->w, q->h, &q->bw, &q->bh);
    q->mv0 = (imv *)malloc(sizeof(imv) * q->bw * q->bh);
    q->mv1 = (imv *)malloc(sizeof(imv) * q->bw * q->bh);
    q->skipped = (uint8_t *)malloc((size_t)q->bw * q->bh);
    if (l == 0) {
      q->y[0] = reversed ? ref1_y : ref0_y;
      q->y[1] = reversed ? ref0_y : ref1_y;
      q->stride = stride_y;
      q->pad = pad_y;
    } else {
      q->pad = 32;
      q->stride = (q->w + 64 + 15) & ~15;
      for (int k = 0; k < 2; k++) {
        q->own[k] = (uint8_t *)malloc((size_t)q->stride * (q->h + 64));
        uint8_t *o = q->own[k] + 32 * q->stride + 32;
        halve_luma(L[l - 1].y[k], L[l - 1].stride, o, q->stride, q->w, q->h, 32);
        q->y[k] = o;
      }
    }
  }
  imv *guide = NULL;
  for (int l = nlev - 1; l >= 0; l--) {
    estimate_level(&L[l], W, guide);
    free(guide);
    guide = NULL;
    if (l > 0) {                                                  /* :247-271: the finer level's guide */
      const level_t *f = &L[l - 1];
      guide = (imv *)malloc(sizeof(imv) * f->bw * f->bh);
      for (int i = 0; i < f->bh; i++)
        for (int j = 0; j < f->bw; j++) {
          const imv c = L[l].mv1[(i / 2) * L[l].bw + j / 2];
          guide[i * f->bw + j].x = (int16_t)(c.x << 1);
          guide[i * f->bw + j].y = (int16_t)(c.y << 1);
        }
    }
  }
  /* :920-970 */
  const level_t *F = &L[0];
  const uint8_t *p0[3] = {F->y[0], reversed ? ref1_u : ref0_u, reversed ? ref1_v : ref0_v};
  const uint8_t *p1[3] = {F->y[1], reversed ? ref0_u : ref1_u, reversed ? ref0_v : ref1_v};
  uint8_t *o[3] = {out_y, out_u, out_v};
  const int wP = width + 4, hP = height + 4;
  for (int pl = 0; pl < 3; pl++)
    for (int by = 0; by < F->bh; by++)
      for (int bx = 0; bx < F->bw; bx++) {
        imv v1 = F->mv1[by * F->bw + bx], v0 = F->mv0[by * F->bw + bx];
        if (pl == 0) blend_block(p0[0], p1[0], stride_y, o[0], out_stride_y, bx * 8, by * 8, v0, v1, wP, hP, 4, 8, width, height);
        else {
          v1.x >>= 1;
          v1.y >>= 1;
          v0 = to_pic0(v1, W);
          blend_block(p0[pl], p1[pl], stride_c, o[pl], out_stride_c, bx * 4, by * 4, v0, v1, wP / 2, hP / 2, 2, 4, width / 2, height / 2);
        }
      }
  if (dbg_mv0) memcpy(dbg_mv0, F->mv0, sizeof(imv) * F->bw * F->bh);
  if (dbg_mv1) memcpy(dbg_mv1, F->mv1, sizeof(imv) * F->bw * F->bh);
  free(guide);
  for (int l = 0; l < nlev; l++) {
    free(L[l].mv0); free(L[l].mv1); free(L[l].skipped); free(L[l].own[0]); free(L[l].own[1]);
  }
}
