/* TEST INFRASTRUCTURE (see thor_oracle.h).  Sub-pel inter prediction and block averaging. */
#include <string.h>
#include "thor_oracle.h"

static inline int pix(int v) { return v < 0 ? 0 : (v > 255 ? 255 : v); }

/* 6-tap quarter-pel luma kernels, phase 0..3; "bi" set is used whenever the encoder was configured
 * with bi-prediction enabled (the flag is the config flag, not the block mode).
 * reference common/inter_prediction.c:47-59. */
static const int8_t LUMA_BI[4][6]  = { {0,0,64,0,0,0}, {2,-10,59,17,-5,1}, {1,-8,39,39,-8,1}, {1,-5,17,59,-10,2} };
static const int8_t LUMA_UNI[4][6] = { {0,0,64,0,0,0}, {1,-7,55,19,-5,1},  {1,-7,38,38,-7,1}, {1,-5,19,55,-7,1} };
/* 4-tap eighth-pel chroma kernels: reference common/inter_prediction.c:61-70 */
static const int8_t CHROMA[8][4] = { {0,64,0,0}, {-2,58,10,-2}, {-4,54,16,-2}, {-4,44,28,-4},
                                     {-4,36,36,-4}, {-4,28,44,-4}, {-2,16,54,-4}, {-2,10,58,-2} };

/* reference common/inter_prediction.c:120-180.  `ref` points at the block origin in a padded
 * reference plane; the motion vector is in quarter-pel units and negated when sign!=0. */
void orc_inter_luma(uint8_t *pblock, const uint8_t *ref, int width, int height, int stride, int pstride,
                    int mvx, int mvy, int sign, int bipred) {
  if (sign) { mvx = -mvx; mvy = -mvy; }
  const int fx = mvx & 3, fy = mvy & 3;
  const uint8_t *src = ref + (mvy >> 2) * stride + (mvx >> 2);
  if (!fx && !fy) {
    for (int r = 0; r < height; r++) memcpy(pblock + r * pstride, src + r * stride, width);
    return;
  }
  if (fx == 2 && fy == 2) {
    /* centre position: 12-tap diamond-ish low-pass, weights [0 1 1 0; 1 2 2 1; 1 2 2 1; 0 1 1 0]/16 */
    static const int8_t W[4][4] = { {0,1,1,0}, {1,2,2,1}, {1,2,2,1}, {0,1,1,0} };
    for (int r = 0; r < height; r++)
      for (int c = 0; c < width; c++) {
        int acc = 0;
        for (int dy = 0; dy < 4; dy++)
          for (int dx = 0; dx < 4; dx++) acc += W[dy][dx] * src[(r + dy - 1) * stride + c + dx - 1];
        pblock[r * pstride + c] = (uint8_t)pix((acc + 8) >> 4);
      }
    return;
  }
  const int8_t (*K)[6] = bipred ? LUMA_BI : LUMA_UNI;
  /* vertical 6-tap first at full precision, then horizontal, single rounding (sum+2048)>>12 */
  for (int r = 0; r < height; r++)
    for (int c = 0; c < width; c++) {
      int acc = 0;
      for (int t = 0; t < 6; t++) {
        int col = 0;
        for (int m = 0; m < 6; m++) col += K[fy][m] * src[(r + m - 2) * stride + c + t - 2];
        acc += K[fx][t] * col;
      }
      pblock[r * pstride + c] = (uint8_t)pix((acc + 2048) >> 12);
    }
}

/* reference common/inter_prediction.c:72-118: the luma vector is reused at eighth-pel precision. */
void orc_inter_chroma(uint8_t *pblock, const uint8_t *ref, int width, int height, int stride, int pstride,
                      int mvx, int mvy, int sign) {
  if (sign) { mvx = -mvx; mvy = -mvy; }
  const int fx = mvx & 7, fy = mvy & 7;
  const uint8_t *src = ref + (mvy >> 3) * stride + (mvx >> 3);
  if (!fx && !fy) {
    for (int r = 0; r < height; r++) memcpy(pblock + r * pstride, src + r * stride, width);
    return;
  }
  for (int r = 0; r < height; r++)
    for (int c = 0; c < width; c++) {
      int acc = 0;
      for (int m = 0; m < 4; m++) {
        int row = 0;
        for (int t = 0; t < 4; t++) row += CHROMA[fx][t] * src[(r + m - 1) * stride + c + t - 1];
        acc += CHROMA[fy][m] * row;
      }
      pblock[r * pstride + c] = (uint8_t)pix((acc + 2048) >> 12);
    }
}

/* bi-prediction average, truncating: reference enc/encode_block.c:1569-1588, dec/decode_block.c:272-283 */
void orc_average_trunc(uint8_t *dst, const uint8_t *a, const uint8_t *b, int n) {
  for (int i = 0; i < n; i++) dst[i] = (uint8_t)((a[i] + b[i]) >> 1);
}

/* rounding average used by temporal interpolation: reference common/common_kernels.c:34-74 */
void orc_block_avg(uint8_t *p, const uint8_t *r0, const uint8_t *r1, int sp, int s0, int s1, int w, int h) {
  for (int r = 0; r < h; r++)
    for (int c = 0; c < w; c++) p[r * sp + c] = (uint8_t)((r0[r * s0 + c] + r1[r * s1 + c] + 1) >> 1);
}
