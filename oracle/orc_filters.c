/* TEST INFRASTRUCTURE (see thor_oracle.h).  In-loop filters: deblocking, CLPF, border padding. */
#include <stdlib.h>
#include <string.h>
#include "thor_oracle.h"

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
static inline int pix(int v) { return clampi(v, 0, 255); }

/* beta / tc thresholds per QP: reference common/common_frame.c:36-44.
 * beta: 0 below qp 16, then 6..18 in steps of 1, then steps of 2 up to 64.
 * tc has 56 entries in the reference (qp+4 for intra was planned); indices 0..51 are used. */
static int beta_of_qp(int qp) { return qp < 16 ? 0 : (qp <= 28 ? qp - 10 : 2 * qp - 38); }
static const uint8_t tc_of_qp[56] = {
  0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,1,1,1,1,1,1,1,1,1,2,2,2,2,3,3,3,3,4,4,4,5,5,6,6,7,8,9,9,
  10,10,11,11,12,12,13,13,14,14 };

/* common/common_block.c:78-83 */
int orc_chroma_qp(int qp) {
  static const uint8_t t[52] = {
    0,1,2,3,4,5,6,7,8,9,10,11,12,13,14,15,16,17,18,19,20,21,22,23,24,25,26,27,28,29,29,30,31,32,33,33,
    34,34,35,35,36,36,37,37,38,39,40,41,42,43,44,45 };
  return t[qp];
}

static int mv_is_large(const orc_blockinfo_t *b) {   /* NEW_MV_TEST, common_frame.c:119-120 */
  return abs(b->mv0.x) >= 4 || abs(b->mv0.y) >= 4 || abs(b->mv1.x) >= 4 || abs(b->mv1.y) >= 4;
}

/* One luma edge pass.  vertical=1: edges between columns (x = 8,16,..), samples step along x
 * (across = 1, along = stride).  vertical=0: edges between rows.  common_frame.c:72-153 / 155-240. */
static void luma_edge_pass(uint8_t *y, int stride, const orc_blockinfo_t *bi, int width, int height,
                           int beta, int tc, int vertical) {
  const int bw = width / 4;
  const int across = vertical ? 1 : stride;   /* step from p to q */
  const int along = vertical ? stride : 1;    /* step along the edge */
  for (int by = 0; by < height; by += 8) {
    for (int bx = 0; bx < width; bx += 8) {
      const int epos = vertical ? bx : by;    /* coordinate normal to the edge */
      if (epos == 0) continue;
      uint8_t *e = y + by * stride + bx;      /* q0 of line 0 of this 8-sample edge */
      /* activity on lines 2 and 5 (NEW_DEBLOCK_TEST, common_frame.c:77-86) */
      int d = 0;
      for (int ln = 2; ln <= 5; ln += 3) {
        const uint8_t *s = e + ln * along;
        d += abs(s[-2 * across] - s[-across]) + abs(s[across] - s[0]);
      }
      for (int seg = 0; seg < 8; seg += 4) {
        const int ux = (bx + (vertical ? 0 : seg)) / 4, uy = (by + (vertical ? seg : 0)) / 4;
        const orc_blockinfo_t *q = &bi[uy * bw + ux];
        const orc_blockinfo_t *p = vertical ? q - 1 : q - bw;
        int qsize = q->size;
        const int splits = q->tb_split || q->pb_part == 3 || q->pb_part == (vertical ? 2 : 1);
        if (splits && qsize > 8) qsize /= 2;
        const int on_boundary = (epos % qsize) == 0;
        const int strong = mv_is_large(p) || mv_is_large(q) || p->cbp_y || q->cbp_y || p->mode == 1 || q->mode == 1;
        if (!(d < beta && on_boundary && strong)) continue;
        for (int ln = seg; ln < seg + 4; ln++) {
          uint8_t *s = e + ln * along;
          const int p1 = s[-2 * across], p0 = s[-across], q0 = s[0], q1 = s[across];
          int delta = (18 * (q0 - p0) - 6 * (q1 - p1) + 16) >> 5;    /* NEW_DEBLOCK_FILTER :138 */
          delta = clampi(delta, -tc, tc);
          s[-2 * across] = (uint8_t)pix(p1 + delta / 2);             /* C division: toward zero */
          s[-across]     = (uint8_t)pix(p0 + delta);
          s[0]           = (uint8_t)pix(q0 - delta);
          s[across]      = (uint8_t)pix(q1 - delta / 2);
        }
      }
    }
  }
}

/* reference common/common_frame.c:46-241 */
void orc_deblock_frame_y(uint8_t *y, int stride, const orc_blockinfo_t *bi, int width, int height, int qp) {
  const int beta = beta_of_qp(qp), tc = tc_of_qp[qp];
  luma_edge_pass(y, stride, bi, width, height, beta, tc, 1);
  luma_edge_pass(y, stride, bi, width, height, beta, tc, 0);
}

/* Chroma: only edges on coding-block boundaries next to an intra block; 2 samples change.
 * reference common/common_frame.c:243-321 (side info of the first 4x4 unit of each 8x8 only). */
static void chroma_edge_pass(uint8_t *c, int stride, const orc_blockinfo_t *bi, int width, int height, int tc, int vertical) {
  const int bw = width / 4;
  const int across = vertical ? 1 : stride, along = vertical ? stride : 1;
  for (int by = 0; by < height; by += 8) {
    for (int bx = 0; bx < width; bx += 8) {
      const int epos = vertical ? bx : by;
      if (epos == 0) continue;
      const orc_blockinfo_t *q = &bi[(by / 4) * bw + bx / 4];
      const orc_blockinfo_t *p = vertical ? q - 1 : q - bw;
      if ((epos % q->size) != 0 || !(p->mode == 1 || q->mode == 1)) continue;
      uint8_t *e = c + (by / 2) * stride + bx / 2;
      for (int ln = 0; ln < 4; ln++) {
        uint8_t *s = e + ln * along;
        const int p1 = s[-2 * across], p0 = s[-across], q0 = s[0], q1 = s[across];
        int delta = clampi((4 * (q0 - p0) + (p1 - q1) + 4) >> 3, -tc, tc);
        s[-across] = (uint8_t)pix(p0 + delta);
        s[0]       = (uint8_t)pix(q0 - delta);
      }
    }
  }
}

void orc_deblock_frame_uv(uint8_t *u, uint8_t *v, int stride_c, const orc_blockinfo_t *bi, int width, int height, int qpc) {
  const int tc = tc_of_qp[qpc];
  for (int pl = 0; pl < 2; pl++) {
    uint8_t *c = pl ? v : u;
    chroma_edge_pass(c, stride_c, bi, width, height, tc, 1);
    chroma_edge_pass(c, stride_c, bi, width, height, tc, 0);
  }
}

/* CLPF sample rule: +1 if more than two of the 4-neighbours are larger, -1 if more than two are
 * smaller.  Neighbours outside [left..right]x[top..bottom] are replaced by the centre sample.
 * reference common/common_block.c:180-197. */
static int clpf_sample(const uint8_t *src, int stride, int x, int y, int left, int top, int right, int bottom) {
  const int X = src[y * stride + x];
  const int A = y == top ? X : src[(y - 1) * stride + x];
  const int B = x == left ? X : src[y * stride + x - 1];
  const int C = x == right ? X : src[y * stride + x + 1];
  const int D = y == bottom ? X : src[(y + 1) * stride + x];
  const int up = (A > X) + (B > X) + (C > X) + (D > X);
  const int dn = (A < X) + (B < X) + (C < X) + (D < X);
  return X + (up > 2) - (dn > 2);
}

/* dst is an SB-local buffer of pitch dstride; the window that bounds neighbour access is the
 * dstride-aligned square containing (x0,y0) (dstride = 64 luma / 32 chroma in clpf_frame). */
void orc_clpf_block(const uint8_t *src, uint8_t *dst, int sstride, int dstride, int x0, int y0, int size, int width, int height) {
  const int left = x0 & ~(dstride - 1), top = y0 & ~(dstride - 1);
  const int right = (left + dstride - 1 < width - 1) ? left + dstride - 1 : width - 1;
  const int bottom = (top + dstride - 1 < height - 1) ? top + dstride - 1 : height - 1;
  for (int y = y0; y < y0 + size; y++)
    for (int x = x0; x < x0 + size; x++)
      dst[(y - top) * dstride + (x - left)] = (uint8_t)clpf_sample(src, sstride, x, y, left, top, right, bottom);
}

/* reference enc/encode_block.c:3036-3057 (accumulates into *sum0,*sum1) */
void orc_detect_clpf(const uint8_t *rec, const uint8_t *org, int x0, int y0, int width, int height, int so, int stride, int *sum0, int *sum1) {
  const int left = x0 & ~63, top = y0 & ~63;
  const int right = (left + 63 < width - 1) ? left + 63 : width - 1;
  const int bottom = (top + 63 < height - 1) ? top + 63 : height - 1;
  for (int y = y0; y < y0 + 8; y++)
    for (int x = x0; x < x0 + 8; x++) {
      const int O = org[y * so + x], X = rec[y * stride + x];
      const int F = clpf_sample(rec, stride, x, y, left, top, right, bottom);
      *sum0 += (O - X) * (O - X);
      *sum1 += (O - F) * (O - F);
    }
}

/* reference common/common_frame.c:485-557 with the decision callbacks of enc/encode_frame.c:46-63
 * (encoder) and dec/decode_frame.c:130-133 (decoder) folded in. */
int orc_clpf_frame(uint8_t *y, uint8_t *u, uint8_t *v, int stride_y, int stride_c,
                   const uint8_t *org_y, int org_stride, const orc_blockinfo_t *bi,
                   int width, int height, const uint8_t *sb_bits_in, uint8_t *sb_bits_out) {
  const int nsbx = width / 64, nsby = height / 64, bw = width / 4;   /* floor: partial SBs untouched */
  int signalled = 0;
  uint8_t ty[64 * 64], tu[32 * 32], tv[32 * 32];
  for (int k = 0; k < nsby; k++)
    for (int l = 0; l < nsbx; l++) {
      int cand = 0;
      for (int m = 0; m < 8; m++)
        for (int n = 0; n < 8; n++) {
          const orc_blockinfo_t *b = &bi[(k * 16 + m * 2) * bw + l * 16 + n * 2];
          cand |= b->mode != 3 && (b->cbp_y || b->cbp_u || b->cbp_v);
        }
      if (sb_bits_out) sb_bits_out[k * nsbx + l] = 0xff;
      if (!cand) continue;
      int on;
      if (org_y) {
        int s0 = 0, s1 = 0;
        for (int m = 0; m < 8; m++)
          for (int n = 0; n < 8; n++) {
            const orc_blockinfo_t *b = &bi[(k * 16 + m * 2) * bw + l * 16 + n * 2];
            if (b->cbp_y && b->mode != 3)
              orc_detect_clpf(y, org_y, l * 64 + n * 8, k * 64 + m * 8, width, height, org_stride, stride_y, &s0, &s1);
          }
        on = s1 < s0;
      } else {
        on = sb_bits_in[k * nsbx + l] != 0;
      }
      if (sb_bits_out) sb_bits_out[k * nsbx + l] = (uint8_t)on;
      signalled++;
      if (!on) continue;
      for (int r = 0; r < 64; r++) memcpy(ty + r * 64, y + (k * 64 + r) * stride_y + l * 64, 64);
      for (int r = 0; r < 32; r++) {
        memcpy(tu + r * 32, u + (k * 32 + r) * stride_c + l * 32, 32);
        memcpy(tv + r * 32, v + (k * 32 + r) * stride_c + l * 32, 32);
      }
      for (int m = 0; m < 8; m++)
        for (int n = 0; n < 8; n++) {
          const int xpos = l * 64 + n * 8, ypos = k * 64 + m * 8;
          const orc_blockinfo_t *b = &bi[(ypos / 4) * bw + xpos / 4];
          if (b->mode == 3) continue;
          if (b->cbp_y) orc_clpf_block(y, ty, stride_y, 64, xpos, ypos, 8, width, height);
          if (b->cbp_u) orc_clpf_block(u, tu, stride_c, 32, xpos / 2, ypos / 2, 4, width / 2, height / 2);
          if (b->cbp_v) orc_clpf_block(v, tv, stride_c, 32, xpos / 2, ypos / 2, 4, width / 2, height / 2);
        }
      for (int r = 0; r < 64; r++) memcpy(y + (k * 64 + r) * stride_y + l * 64, ty + r * 64, 64);
      for (int r = 0; r < 32; r++) {
        memcpy(u + (k * 32 + r) * stride_c + l * 32, tu + r * 32, 32);
        memcpy(v + (k * 32 + r) * stride_c + l * 32, tv + r * 32, 32);
      }
    }
  return signalled;
}

/* reference common/common_frame.c:405-462 (one plane) */
void orc_pad_plane(uint8_t *p, int stride, int w, int h, int pad_hor, int pad_ver) {
  for (int r = 0; r < h; r++) {
    memset(p + r * stride - pad_hor, p[r * stride], pad_hor);
    memset(p + r * stride + w, p[r * stride + w - 1], pad_hor);
  }
  for (int r = 1; r <= pad_ver; r++) {
    memcpy(p - r * stride - pad_hor, p - pad_hor, w + 2 * pad_hor);
    memcpy(p + (h - 1 + r) * stride - pad_hor, p + (h - 1) * stride - pad_hor, w + 2 * pad_hor);
  }
}
