/* TEST INFRASTRUCTURE (see thor_oracle.h).  Motion-search cost functions. */
#include <stdlib.h>
#include "thor_oracle.h"

/* reference enc/encode_block.c:740-755 */
unsigned orc_sad(const uint8_t *a, const uint8_t *b, int as, int bs, int w, int h) {
  unsigned s = 0;
  for (int r = 0; r < h; r++) for (int c = 0; c < w; c++) s += (unsigned)abs(a[r * as + c] - b[r * bs + c]);
  return s;
}

/* reference enc/encode_block.c:757-780: five horizontal offsets, first minimum in the order
 * -3,-1,0,+1,+3 wins (the SIMD twin packs (sad<<3)|idx and takes the min: same order). */
unsigned orc_widesad(const uint8_t *a, const uint8_t *b, int as, int bs, int w, int h, int *x) {
  static const int off[5] = { -3, -1, 0, 1, 3 };
  unsigned best = 1u << 31; int bx = 0;
  for (int k = 0; k < 5; k++) {
    const unsigned s = orc_sad(a, b + off[k], as, bs, w, h);
    if (s < best) { best = s; bx = off[k]; }
  }
  *x = bx;
  return best;
}

/* reference enc/encode_block.c:782-797 */
int orc_ssd(const uint8_t *a, const uint8_t *b, int as, int bs, int w, int h) {
  int s = 0;
  for (int r = 0; r < h; r++) for (int c = 0; c < w; c++) { const int d = a[r * as + c] - b[r * bs + c]; s += d * d; }
  return s;
}

static inline int up(int a, int b) { return (a + b + 1) >> 1; }   /* rounding average  */
static inline int dn(int a, int b) { return (a + b) >> 1; }       /* truncating average */

/* reference enc/encode_block.c:497-605.  Eight bilinear "half-pel" candidates around the integer
 * position b.  With P(dy,dx) = b[dy*bs+dx] at the current sample, the four axis candidates are
 * rounding averages of two neighbours; each diagonal is a fixed tree of averages over the 4x4
 * neighbourhood whose rounding directions matter and are kept exactly. */
static int diag_half(const uint8_t *p, int bs, int sy, int sx) {
  /* sy,sx in {0,1}: candidate between rows (sy-1, sy) and columns (sx-1, sx) */
  const uint8_t *o = p + (sy - 1) * bs + (sx - 1);      /* top-left sample of the 2x2 core */
#define Q(r, c) o[(r) * bs + (c)]
  const int core = dn(up(Q(0,0), Q(0,1)), up(Q(1,0), Q(1,1)));
  const int vert = dn(up(Q(-1,0), Q(2,0)), up(Q(-1,1), Q(2,1)));
  const int hori = dn(up(Q(0,-1), Q(0,2)), up(Q(1,-1), Q(1,2)));
#undef Q
  /* the reference averages (vert, hori) in that operand order; dn() is symmetric */
  return dn(dn(vert, hori), core);
}

unsigned orc_sad_fasthalf(const uint8_t *a, const uint8_t *b, int as, int bs, int w, int h, int *x, int *y) {
  unsigned top = 0, down = 0, left = 0, right = 0, tl = 0, tr = 0, bl = 0, br = 0;
  for (int r = 0; r < h; r++)
    for (int c = 0; c < w; c++) {
      const uint8_t *p = b + r * bs + c;
      const int A = a[r * as + c];
      top   += (unsigned)abs(A - up(p[0], p[-bs]));
      down  += (unsigned)abs(A - up(p[0], p[bs]));
      left  += (unsigned)abs(A - up(p[-1], p[0]));
      right += (unsigned)abs(A - up(p[0], p[1]));
      tl += (unsigned)abs(A - diag_half(p, bs, 0, 0));
      tr += (unsigned)abs(A - diag_half(p, bs, 0, 1));
      bl += (unsigned)abs(A - diag_half(p, bs, 1, 0));
      br += (unsigned)abs(A - diag_half(p, bs, 1, 1));
    }
  /* strict '<' in this order; start from "top" = (0,-2) (:577-603) */
  unsigned best = top; int bx = 0, by = -2;
  if (down  < best) { best = down;  bx = 0;  by = 2; }
  if (right < best) { best = right; bx = 2;  by = 0; }
  if (left  < best) { best = left;  bx = -2; by = 0; }
  if (tl    < best) { best = tl;    bx = -2; by = -2; }
  if (tr    < best) { best = tr;    bx = 2;  by = -2; }
  if (br    < best) { best = br;    bx = 2;  by = 2; }
  if (bl    < best) { best = bl;    bx = -2; by = 2; }
  *x = bx; *y = by;
  return best;
}

/* reference enc/encode_block.c:609-738.  Eight quarter-pel candidates around the chosen half-pel
 * point; the incoming (*x,*y) only select which of four formula sets is used: "both" when the
 * bitwise AND of x and y is non-zero, else "x", else "y", else "none". */
unsigned orc_sad_fastquarter(const uint8_t *o, const uint8_t *r, int os, int rs, int w, int h, int *x, int *y) {
  const int sel = (*x & *y) ? 0 : (*x ? 1 : (*y ? 2 : 3));
  unsigned s[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };   /* tl top tr left right bl down br */
  for (int i = 0; i < h; i++, r += rs)
    for (int j = 0; j < w; j++) {
      const int O = o[i * os + j];
      const int a = r[j], b = r[j - rs], c = r[j - rs + 1], d = r[j + 1], e = r[j + rs + 1], f = r[j + rs],
                g = r[j + rs - 1], hh = r[j - 1];
      const int ad = up(a, d), de = up(d, e), af = up(a, f), fe = up(f, e), dc = up(d, c), ab = up(a, b),
                ah = up(a, hh), gf = up(g, f);
      int v[8];
      switch (sel) {
      case 0: v[0]=dn(ad,af); v[1]=dn(de,a);  v[2]=dn(ad,de); v[3]=dn(ad,f);  v[4]=dn(ad,e); v[5]=dn(af,fe); v[6]=dn(de,f); v[7]=dn(de,fe); break;
      case 1: v[0]=dn(ad,ab); v[1]=dn(dc,a);  v[2]=dn(ad,dc); v[3]=dn(ad,a);  v[4]=dn(ad,d); v[5]=dn(ad,af); v[6]=dn(af,d); v[7]=dn(ad,de); break;
      case 2: v[0]=dn(ah,af); v[1]=dn(af,a);  v[2]=dn(ad,af); v[3]=dn(gf,a);  v[4]=dn(ad,f); v[5]=dn(af,gf); v[6]=dn(af,f); v[7]=dn(af,fe); break;
      default:v[0]=dn(ah,ab); v[1]=dn(ab,a);  v[2]=dn(ad,ab); v[3]=dn(ah,a);  v[4]=dn(ad,a); v[5]=dn(ah,af); v[6]=dn(af,a); v[7]=dn(af,ad); break;
      }
      for (int k = 0; k < 8; k++) s[k] += (unsigned)abs(O - v[k]);
    }
  /* start from top = (0,-1); order tl,tr,left,right,bl,down,br; note tl/tr keep besty (:703-734) */
  unsigned best = s[1]; int bx = 0, by = -1;
  if (s[0] < best) { best = s[0]; bx = -1; }
  if (s[2] < best) { best = s[2]; bx = 1; }
  if (s[3] < best) { best = s[3]; bx = -1; by = 0; }
  if (s[4] < best) { best = s[4]; bx = 1;  by = 0; }
  if (s[5] < best) { best = s[5]; bx = -1; by = 1; }
  if (s[6] < best) { best = s[6]; bx = 0;  by = 1; }
  if (s[7] < best) { best = s[7]; bx = 1;  by = 1; }
  *x = bx; *y = by;
  return best;
}

/* reference enc/encode_block.c:816-828 (note: the right-edge test omits "+ size"; C '/' truncates) */
void orc_clip_mv(int16_t *mvx, int16_t *mvy, int ypos, int xpos, int fwidth, int fheight, int size, int sign) {
  const int ext = 96 - 16;
  int vy = sign ? -*mvy : *mvy, vx = sign ? -*mvx : *mvx;
  if (ypos + vy / 4 < -ext) vy = 4 * (-ext - ypos);
  if (ypos + vy / 4 + size > fheight + ext) vy = 4 * (fheight + ext - ypos - size);
  if (xpos + vx / 4 < -ext) vx = 4 * (-ext - xpos);
  if (xpos + vx / 4 > fwidth + ext) vx = 4 * (fwidth + ext - xpos - size);
  *mvy = (int16_t)(sign ? -vy : vy);
  *mvx = (int16_t)(sign ? -vx : vx);
}
