/* TEST INFRASTRUCTURE (see thor_oracle.h).  Integer DCT, quantisation, residual / reconstruct. */
#include <stdlib.h>
#include <string.h>
#include "thor_oracle.h"

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
static int ilog2(int v) { int n = 0; while (v >>= 1) n++; return n; }

/* The HEVC-style integer DCT basis (reference common/transform.c:41-245, g1..g4mat_hevc) is
 * M_N[k][n] = C[ (2n+1) * k * 32/N ] with C the 32-point cosine ladder below extended by the
 * symmetries of cos(pi*a/64); orc_dct_matrix() expands it, tests compare with the reference tables. */
static const int8_t COS32[33] = { 64, 90,90,90,89,88,87,85,83,82,80,78,75,73,70,67,64,61,57,54,50,46,43,38,36,31,25,22,18,13,9,4, 0 };
static int16_t MAT[4][32 * 32];
static int zz[3][256];
static int tables_ready;

static void build_tables(void) {
  for (int t = 0; t < 4; t++) {
    const int N = 4 << t;
    for (int k = 0; k < N; k++)
      for (int n = 0; n < N; n++) {
        int a = ((2 * n + 1) * k * (32 / N)) & 127;
        if (a > 64) a = 128 - a;
        MAT[t][k * N + n] = (int16_t)(a > 32 ? -COS32[64 - a] : COS32[a]);
      }
  }
  for (int t = 0; t < 3; t++) {               /* classic zig-zag; reference common/common_block.c:38-73 */
    const int q = 4 << t;
    int idx = 0;
    for (int s = 0; s <= 2 * q - 2; s++)
      for (int m = 0; m <= s; m++) {
        const int i = (s & 1) ? m : s - m, j = s - i;
        if (i < q && j < q) zz[t][i * q + j] = idx++;
      }
  }
  tables_ready = 1;
}

const int16_t *orc_dct_matrix(int size) { if (!tables_ready) build_tables(); return MAT[ilog2(size) - 2]; }
const int *orc_zigzag(int qsize) { if (!tables_ready) build_tables(); return zz[ilog2(qsize) - 2]; }

/* reference common/transform.c:249-330.  Only the low min(N,16)^2 coefficients are produced; the
 * rest of coeff[] is left untouched (the reference leaves it uninitialised). */
void orc_transform(const int16_t *block, int16_t *coeff, int size, int fast) {
  const int out_stride = size, keep = size < 16 ? size : 16;
  int16_t pooled[32 * 32], mid[16][32];
  const int16_t *in = block;
  int n = size, sh1 = ilog2(size), sh2 = ilog2(size) + 5;
  if (size > 16 && fast) {                 /* box-sum down to 16x16, 16-point transform (:273-293) */
    const int f = size / 16;
    for (int i = 0; i < 16; i++)
      for (int j = 0; j < 16; j++) {
        int s = 0;
        for (int a = 0; a < f; a++) for (int b = 0; b < f; b++) s += block[(i * f + a) * size + j * f + b];
        pooled[i * 16 + j] = (int16_t)s;
      }
    in = pooled; n = 16; sh1 = (size == 64) ? 8 : 6; sh2 = 9;
  } else if (size == 64) {                 /* 2x2-sum down to 32x32, 32-point transform (:295-307) */
    for (int i = 0; i < 32; i++)
      for (int j = 0; j < 32; j++)
        pooled[i * 32 + j] = (int16_t)(block[(2 * i) * 64 + 2 * j] + block[(2 * i + 1) * 64 + 2 * j] +
                                       block[(2 * i) * 64 + 2 * j + 1] + block[(2 * i + 1) * 64 + 2 * j + 1]);
    in = pooled; n = 32; sh1 = 7; sh2 = 10;
  }
  const int16_t *M = orc_dct_matrix(n);
  for (int k = 0; k < keep; k++)           /* rows of the block -> horizontal frequency k */
    for (int r = 0; r < n; r++) {
      int acc = 0;
      for (int c = 0; c < n; c++) acc += M[k * n + c] * in[r * n + c];
      mid[k][r] = (int16_t)((acc + (1 << (sh1 - 1))) >> sh1);
    }
  for (int v = 0; v < keep; v++)           /* vertical frequency v */
    for (int k = 0; k < keep; k++) {
      int acc = 0;
      for (int r = 0; r < n; r++) acc += M[v * n + r] * mid[k][r];
      coeff[v * out_stride + k] = (int16_t)((acc + (1 << (sh2 - 1))) >> sh2);
    }
}

/* reference common/transform.c:432-518.  The even/odd butterflies there are an exact integer
 * factorisation of the matrix product over the first min(N,16) coefficient rows/columns. */
static void inverse_core(const int16_t *coeff, int cstride, int16_t *block, int n) {
  const int keep = n < 16 ? n : 16;
  const int16_t *M = orc_dct_matrix(n);
  int16_t mid[16][32];
  for (int k = 0; k < keep; k++)           /* column k of coeff -> spatial column vector */
    for (int r = 0; r < n; r++) {
      int acc = 0;
      for (int v = 0; v < keep; v++) acc += M[v * n + r] * coeff[v * cstride + k];
      mid[k][r] = (int16_t)clampi((acc + 64) >> 7, -32768, 32767);
    }
  for (int r = 0; r < n; r++)
    for (int c = 0; c < n; c++) {
      int acc = 0;
      for (int k = 0; k < keep; k++) acc += M[k * n + c] * mid[k][r];
      block[r * n + c] = (int16_t)clampi((acc + 2048) >> 12, -32768, 32767);
    }
}

void orc_inverse_transform(const int16_t *coeff, int16_t *block, int size) {
  if (size < 64) { inverse_core(coeff, size, block, size); return; }
  int16_t half[32 * 32];                   /* 64: 32-point inverse of the top-left, 2x replicate (:494-516) */
  inverse_core(coeff, 64, half, 32);
  for (int r = 0; r < 64; r++)
    for (int c = 0; c < 64; c++) block[r * 64 + c] = half[(r >> 1) * 32 + (c >> 1)];
}

/* reference enc/encode_block.c:75-178 (rdoq == 0, the only setting any shipped config uses). */
int orc_quantize(const int16_t *coeff, int16_t *coeffq, int qp, int size, int coeff_block_type) {
  static const uint16_t QSCALE[6] = { 26214, 23302, 20560, 18396, 16384, 14564 };   /* common_block.c:97 */
  static const uint16_t DQSCALE[6] = { 40, 45, 51, 57, 64, 72 };                     /* common_block.c:98 */
  const int intra = (coeff_block_type >> 1) & 1, chroma = coeff_block_type & 1;
  const int lg = ilog2(size), q = size < 16 ? size : 16, nq = q * q;
  const int scale = QSCALE[qp % 6], shift = 21 - lg + qp / 6;
  const int *scan = orc_zigzag(q);
  int lin[256], lev[256];
  for (int i = 0; i < q; i++) for (int j = 0; j < q; j++) lin[scan[i * q + j]] = coeff[i * size + j];
  memset(lev, 0, sizeof(int) * nq);

  /* last significant position under the (smaller) trailing-coefficient offset */
  const int tail_off = (intra ? 38 : -26) << (shift - 8);
  int last = -1;
  for (int p = nq - 1; p >= 0; p--)
    if ((abs(abs(lin[p]) * scale + tail_off) >> shift) != 0) { last = p; break; }

  int any = 0;
  const int off_small = (intra ? 102 : 51) << (shift - 8), off_big = (intra ? 115 : 90) << (shift - 8);
  for (int p = 0; p <= last; p++) {
    const int mag = scale * abs(lin[p]);
    const int off = ((mag >> shift) == 0 || chroma) ? off_small : off_big;
    const int l = (mag + off) >> shift;
    lev[p] = lin[p] < 0 ? -l : l;
    any |= (l != 0);
  }

  if (any) {   /* "RDOQ light": isolated level>1 after two zeros is traded for a +-1 (:135-168) */
    const int N = chroma ? last + 1 : nq;
    const int thr = (73 * DQSCALE[qp % 6] << (qp / 6)) >> (4 + lg);
    for (int p = 2; p < N; p++) {
      int ok = 1;
      if (p > 2 && abs(lev[p - 3]) > 1) ok = 0;
      if (p > 3 && abs(lev[p - 4]) > 1 && abs(lev[p - 3]) > 0) ok = 0;
      if (p == 2 && (!chroma || last >= 6)) ok = 0;
      if (!(ok && lev[p - 2] == 0 && lev[p - 1] == 0 && abs(lev[p]) > 1)) continue;
      const int k1 = abs(lin[p]), k2 = abs(lin[p - 1]), k3 = abs(lin[p - 2]);
      if (k1 + (k2 > k3 ? k2 : k3) < thr) lev[p] = lin[p] < 0 ? -1 : 1;
      else if (k2 > k3) lev[p - 1] = lin[p - 1] < 0 ? -1 : 1;
      else lev[p - 2] = lin[p - 2] < 0 ? -1 : 1;
    }
  }
  for (int i = 0; i < q; i++) for (int j = 0; j < q; j++) coeffq[i * size + j] = (int16_t)lev[scan[i * q + j]];
  return any;
}

/* reference common/common_block.c:132-146 (all size*size entries are scaled) */
void orc_dequantize(const int16_t *coeff, int16_t *rcoeff, int qp, int size) {
  static const uint16_t DQSCALE[6] = { 40, 45, 51, 57, 64, 72 };
  const int rshift = ilog2(size) - 1, add = 1 << (rshift - 1);
  for (int i = 0; i < size * size; i++)
    rcoeff[i] = (int16_t)((((int)coeff[i] * DQSCALE[qp % 6] << (qp / 6)) + add) >> rshift);
}

/* reference enc/encode_block.c:484-493 */
void orc_get_residual(int16_t *block, const uint8_t *pblock, const uint8_t *orig, int size, int orig_stride) {
  for (int r = 0; r < size; r++)
    for (int c = 0; c < size; c++) block[r * size + c] = (int16_t)(orig[r * orig_stride + c] - pblock[r * size + c]);
}

/* reference common/common_block.c:148-156 */
void orc_reconstruct_block(const int16_t *block, const uint8_t *pblock, uint8_t *rec, int size, int stride) {
  for (int r = 0; r < size; r++)
    for (int c = 0; c < size; c++) rec[r * stride + c] = (uint8_t)clampi(block[r * size + c] + pblock[r * size + c], 0, 255);
}

/* ---- early-skip significance test (reference enc/encode_block.c:2481-2610) ---- */

/* check_early_skip_transform_coeff :2481-2505 -- the threshold is the only floating-point step (double) */
int orc_early_skip_threshold(int qp, int size, double relative_threshold) {
  static const uint16_t QSCALE[6] = { 26214, 23302, 20560, 18396, 16384, 14564 };
  const int shift2 = 21 - ilog2(size) + qp / 6;
  const double first_quantizer_level = (double)(1 << shift2) / (double)QSCALE[qp % 6];
  return (int)(relative_threshold * first_quantizer_level);
}

/* check_early_skip_sub_block :2507-2543: residual -> 2x2 mean -> (N/2)-point transform -> any |coef| > threshold */
int orc_early_skip_luma(const uint8_t *orig, int orig_stride, int size, int qp, const uint8_t *pblock, float early_skip_threshold) {
  int16_t block[16 * 16], tmp[8 * 8], coeff[16 * 16];
  orc_get_residual(block, pblock, orig, size, orig_stride);
  int n = size, thr;
  if (size > 4) {
    n = size / 2;
    for (int i = 0; i < n; i++)
      for (int j = 0; j < n; j++)
        tmp[i * n + j] = (int16_t)((block[(2 * i) * size + 2 * j] + block[(2 * i) * size + 2 * j + 1] + block[(2 * i + 1) * size + 2 * j] +
                                    block[(2 * i + 1) * size + 2 * j + 1] + 2) >> 2);
    orc_transform(tmp, coeff, n, 0);
    thr = orc_early_skip_threshold(qp, n, 0.5 * early_skip_threshold);
  } else {
    orc_transform(block, coeff, n, 0);
    thr = orc_early_skip_threshold(qp, n, early_skip_threshold);
  }
  for (int i = 0; i < n * n; i++)
    if (abs(coeff[i]) > thr) return 1;
  return 0;
}

/* check_early_skip_sub_blockC :2545-2610: column sums of the residual (8 wide: every column; 4 wide: pairs of
 * columns) against a threshold derived with a FIXED log2 size of 5; only positive excursions count */
int orc_early_skip_chroma(const uint8_t *orig, int orig_stride, int size, int qp, const uint8_t *pblock, float early_skip_threshold) {
  static const uint16_t QSCALE[6] = { 26214, 23302, 20560, 18396, 16384, 14564 };
  const int shift2 = 21 - 5 + qp / 6;
  const double first_quantizer_level = (double)(1 << shift2) / (double)QSCALE[qp % 6];
  const int threshold = (int)(early_skip_threshold * first_quantizer_level);
  int16_t block[8 * 8];
  orc_get_residual(block, pblock, orig, size, orig_stride);
  if (size == 8) {
    for (int j = 0; j < 8; j++) {
      int sum = 0;
      for (int i = 0; i < 8; i++) sum += block[i * 8 + j];
      if ((int16_t)sum > (int16_t)threshold) return 1;      /* v128_cmpgt_s16 against v128_dup_16(threshold) */
    }
    return 0;
  }
  for (int j = 0; j < 4; j += 2) {
    int sum = 0;
    for (int i = 0; i < 4; i++) sum += block[i * 4 + j] + block[i * 4 + j + 1];
    if (sum > threshold) return 1;
  }
  return 0;
}
