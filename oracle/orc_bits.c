/* TEST INFRASTRUCTURE ONLY -- CPU oracle of the bit-cost model, first slice: the exact number of bits
 * the reference's entropy coder spends on a block's quantised coefficients and on a motion vector difference.
 * Restates   enc/putvlc.c:133-229   (quote_vlc: code lengths of the 14 VLC tables),
 *            enc/write_bits.c:50-69  (write_mv), :71-108 (find_code), :110-253 (write_coeff)
 * as a pure length computation (nothing is written anywhere).  Pinned against the compiled reference by
 * tests/test_oracle_vs_ref.py::test_coeff_bits_* (write_coeff into a real stream_t, bit position before / after).
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may use this file. */
#include <stdlib.h>
#include "thor_oracle.h"

static int ilog2u(unsigned v) { int n = 0; while (v >>= 1) n++; return n; }

/* length of code number cn in table n; -1 where the reference calls fatalerror (putvlc.c:177,225) */
int orc_vlc_len(unsigned n, unsigned cn) {
  if (n <= 5) {                                   /* :141-157  Golomb-like prefix, exp-Golomb escape */
    const unsigned lim = 6u << n;
    if ((int)cn < (int)lim) return 1 + (int)n + (int)(cn >> n);
    return (6 - (int)n) + 1 + 2 * ilog2u(cn - lim + (1u << n));
  }
  if (n <= 7) return 1 + ((int)n - 4) + (int)(cn >> (n - 4));        /* :158-163 */
  if (n == 8) return cn == 0 ? 1 : cn <= 2 ? 2 : -1;                 /* :164-181 */
  if (n == 9) return cn == 0 ? 3 : cn <= 2 ? 4 : cn < 11 ? 5 : 5 + (int)((cn + 5) >> 4);   /* :182-209 */
  if (n == 10) return 1 + 2 * ilog2u(cn + 1);                        /* :210-213  exp-Golomb */
  if (n == 11) return cn < 2 ? (int)cn + 1 : (int)(cn / 2) + 3;      /* :214-217 */
  if (n == 12) return cn + 1 < 4 ? (int)cn + 1 : 4;                  /* :218-221 */
  if (n == 13) return cn + 1 < 6 ? (int)cn + 1 : 6;                  /* :222-225 */
  return -1;
}

/* write_mv (write_bits.c:50-69): two exp-Golomb codes of the signed differences */
int orc_mv_bits(int mvx, int mvy, int mvpx, int mvpy) {
  const int dx = mvx - mvpx, dy = mvy - mvpy;
  return orc_vlc_len(10, (unsigned)(2 * abs(dx) - (dx < 0))) + orc_vlc_len(10, (unsigned)(2 * abs(dy) - (dy < 0)));
}

/* find_code (write_bits.c:71-108): code number of the joint event (run, level > 1), or of the end of block */
int orc_find_code(int run, int level, int maxrun, int chroma, int eob) {
  const int m = maxrun > 4 ? maxrun : 4;
  const int index = run + (level > 1 ? m + 1 : 0);
  if (eob) return chroma ? 0 : 2;
  if (!chroma && index < 2) return index;
  if (index <= 4) return index + 1;
  if (index <= m) return index + 3;
  if (index == m + 1) return 6;
  if (index == m + 2) return 7;
  return index + 1;
}

/* bits of one run-mode / end-of-block code number (write_bits.c:196-206, :238-249) */
static int event_bits(int cn, int chroma, int size, int eob) {
  if (chroma && size <= 8) return orc_vlc_len(eob ? 0 : 10, (unsigned)cn);
  return cn == 0 ? 2 : orc_vlc_len(2, (unsigned)cn + 1);
}

/* write_coeff (write_bits.c:110-253) as a length: coeff = the block's quantised levels, row stride `stride` (the
 * reference passes stride == size; libthorcuda's pool uses stride == min(size,16)), type bit0 chroma, bit1 intra.
 * An all-zero block costs 0 bits here (the reference never calls write_coeff for one: cbp is 0, :146-147). */
int orc_coeff_bits(const int16_t *coeff, int stride, int size, int type) {
  const int q = size < 16 ? size : 16, n = q * q, chroma = type & 1;
  const int *zz = orc_zigzag(q);
  int16_t s[256];
  int last = -1, bits = 0;
  for (int i = 0; i < q; i++)
    for (int j = 0; j < q; j++) s[zz[i * q + j]] = coeff[i * stride + j];
  for (int p = 0; p < n; p++) if (s[p]) last = p;
  if (last < 0) return 0;
  if (chroma) {                                                 /* :147-158 */
    if (last == 0 && abs(s[0]) == 1) return 2;
    bits = 1;
  }
  int adaptive = ((type >> 1) & 1) && !chroma;                   /* table of the next level-mode symbol */
  int level_mode = 1, pos = 0, run = 0, maxrun = 0;
  while (pos <= last) {
    const int level = abs(s[pos]);
    if (level_mode) {                                            /* :166-178 */
      bits += orc_vlc_len((unsigned)adaptive, (unsigned)level) + (level > 0);
      if (!chroma) adaptive = level > 3;
      pos++;
      if (level == 0) { level_mode = 0; run = 0; maxrun = n - pos - 1; }
    } else if (level == 0) {                                     /* :184-190 */
      run++;
      pos++;
    } else {                                                     /* :191-218 */
      bits += event_bits(orc_find_code(run, level, maxrun, chroma, 0), chroma, size, 0);
      bits += level > 1 ? orc_vlc_len(0, (unsigned)(2 * (level - 2) + (s[pos] < 0))) : 1;
      pos++;
      level_mode = level > 1;
      run = 0;
      maxrun = n - pos - 1;
    }
  }
  if (pos < n && level_mode) {                                   /* :222-234  one more (zero) level before the EOB */
    bits += orc_vlc_len((unsigned)adaptive, 0);
    pos++;
  }
  if (pos < n) bits += event_bits(orc_find_code(0, 0, 0, chroma, 1), chroma, size, 1);   /* :237-250 */
  return bits;
}
