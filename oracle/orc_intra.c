/* TEST INFRASTRUCTURE ONLY -- CPU oracle of the Thor intra prediction: the neighbour arrays
 * (make_top_and_left) and the ten prediction modes (get_intra_prediction), restated from
 * reference common/intra_prediction.c:39-389.  Pinned against the compiled reference in
 * tests/test_oracle_vs_ref.py::test_intra_*.  Never linked into libthorcuda. */
#include <string.h>
#include "thor_oracle.h"

static uint8_t clip8(int v) { return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }
static int clampi(int v, int lo, int hi) { return v < lo ? lo : v > hi ? hi : v; }

/* Neighbours of the size x size block whose top-left sample is rec[0] (plane pointer already offset), in picture
 * terms.  (x, y): position of THIS block in its plane; cb_x: x of the coding block it belongs to (differs only
 * under tb_split).  intra_prediction.c:57-143 expressed once for both of its branches:
 *   - row above / column left come from the picture (under tb_split the "rblock" pointer is the same picture);
 *   - 128 fill when the block touches the top (y == 0) or left (x == 0) picture edge;
 *   - one extra sample when up-right / down-left is available, then replicate the last one up to 2*size;
 *   - top_left: the real corner only if the CODING BLOCK is not at the left edge (:83,:116 test xpos, not xpos+j),
 *     else top[0]; at the top edge it becomes left[0]. */
void orc_intra_neighbours(const uint8_t *rec, int stride, int x, int y, int cb_x, int size, int upright, int downleft,
                          uint8_t *left, uint8_t *top, uint8_t *top_left) {
  const int toplen = upright ? size + 1 : size, leftlen = downleft ? size + 1 : size;
  int k;
  if (y == 0) {
    memset(top, 128, 2 * size);
    *top_left = 128;
  } else {
    for (k = 0; k < toplen; k++) top[k] = rec[-stride + k];
    for (k = size; k < 2 * size; k++) top[k] = top[toplen - 1];
    *top_left = cb_x > 0 ? rec[-stride - 1] : top[0];
  }
  if (x == 0) {
    memset(left, 128, 2 * size);
  } else {
    for (k = 0; k < leftlen; k++) left[k] = rec[k * stride - 1];
    for (k = size; k < 2 * size; k++) left[k] = left[leftlen - 1];
  }
  if (y == 0) *top_left = left[0];
}

/* [1 2 1]/4 with the end samples repeated (:39-48) */
static void smooth121(const uint8_t *in, uint8_t *out, int len) {
  for (int j = 0; j < len; j++)
    out[j] = (uint8_t)((in[clampi(j - 1, 0, len - 1)] + 2 * in[j] + in[clampi(j + 1, 0, len - 1)] + 2) >> 2);
}

/* get_intra_prediction :363-389 with the mode functions :145-361; pblock is size x size, stride size */
void orc_intra_predict(const uint8_t *left, const uint8_t *top, uint8_t top_left, int ypos, int xpos, int size, uint8_t *pblock, int mode) {
  uint8_t tf[128], lf[128], tlf;
  int i, j;
  switch (mode) {
    case 2: /* MODE_HOR :161 */
      for (i = 0; i < size; i++) memset(pblock + i * size, left[i], size);
      return;
    case 3: /* MODE_VER :172 */
      for (i = 0; i < size; i++) memcpy(pblock + i * size, top, size);
      return;
    case 1: { /* MODE_PLANAR :182-214: [1 2 2 2 1] smoothing with repeated ends, plane sum, C division then clip */
      int16_t tF[64], lF[64];
      for (j = 0; j < size; j++) {
        const int a = clampi(j - 2, 0, size - 1), b = clampi(j - 1, 0, size - 1), c = clampi(j + 1, 0, size - 1), d = clampi(j + 2, 0, size - 1);
        tF[j] = (int16_t)(top[a] + 2 * top[b] + 2 * top[j] + 2 * top[c] + top[d]);
        lF[j] = (int16_t)(left[a] + 2 * left[b] + 2 * left[j] + 2 * left[c] + left[d]);
      }
      const int tlF = left[1] + 2 * left[0] + 2 * top_left + 2 * top[0] + top[1];
      for (i = 0; i < size; i++)
        for (j = 0; j < size; j++) pblock[i * size + j] = clip8((lF[i] + tF[j] - tlF + 4) / 8);
      return;
    }
    case 4: case 7: case 8: /* UPLEFT :216, UPUPLEFT :281, UPLEFTLEFT :312: smoothed over `size` samples */
      smooth121(left, lf, size);
      smooth121(top, tf, size);
      tlf = (uint8_t)((2 * top_left + left[0] + top[0] + 2) >> 2);
      for (i = 0; i < size; i++)
        for (j = 0; j < size; j++) {
          int v;
          if (mode == 4) {
            const int d = i - j;
            v = d > 0 ? lf[d - 1] : d == 0 ? tlf : tf[-d - 1];
          } else if (mode == 7) {
            const int d = i - 2 * j;
            if (d > 1) v = lf[d - 2];
            else if (d == 1) v = tlf;
            else if (d == 0) v = (tlf + tf[0]) >> 1;
            else v = (d & 1) ? tf[(-d) / 2] : (tf[(-d) / 2] + tf[(-d) / 2 - 1]) >> 1;
          } else {
            const int d = 2 * i - j;
            if (d < -1) v = tf[-d - 2];
            else if (d == -1) v = tlf;
            else if (d == 0) v = (tlf + lf[0]) >> 1;
            else v = (d & 1) ? lf[d / 2] : (lf[d / 2] + lf[d / 2 - 1]) >> 1;
          }
          pblock[i * size + j] = (uint8_t)v;
        }
      return;
    case 5: case 6: /* UPRIGHT :243, UPUPRIGHT :260: top smoothed over 2*size samples */
      smooth121(top, tf, 2 * size);
      for (i = 0; i < size; i++)
        for (j = 0; j < size; j++) {
          if (mode == 5) pblock[i * size + j] = tf[i + j + 1];
          else {
            const int d = i + 2 * j;
            pblock[i * size + j] = (uint8_t)((d & 1) ? tf[(d + 1) / 2] : (tf[d / 2] + tf[d / 2 + 1]) >> 1);
          }
        }
      return;
    case 9: /* DOWNLEFTLEFT :343: left smoothed over 2*size samples */
      smooth121(left, lf, 2 * size);
      for (i = 0; i < size; i++)
        for (j = 0; j < size; j++) {
          const int d = 2 * i + j;
          pblock[i * size + j] = (uint8_t)((d & 1) ? lf[(d + 1) / 2] : (lf[d / 2] + lf[d / 2 + 1]) >> 1);
        }
      return;
    default: { /* MODE_DC :145 (and any unknown mode, :387): the available side counts twice at a picture edge (:367) */
      const uint8_t *a = xpos != 0 ? left : top, *b = ypos != 0 ? top : left;
      int sum = 0;
      for (j = 0; j < size; j++) sum += a[j] + b[j];
      memset(pblock, (sum + size) / (2 * size), (size_t)size * size);
      return;
    }
  }
}

/* reference common/common_block.c:110-128 */
int orc_upright_available(int ypos, int xpos, int size, int width) {
  int a = ypos > 0 && xpos + size < width;
  if (size == 32 && ypos % 64 == 32) a = 0;
  if (size == 16 && (ypos % 32 == 16 || (ypos % 64 == 32 && xpos % 32 == 16))) a = 0;
  if (size == 8 && (ypos % 16 == 8 || (ypos % 32 == 16 && xpos % 16 == 8) || (ypos % 64 == 32 && xpos % 32 == 24))) a = 0;
  return a;
}
int orc_downleft_available(int ypos, int xpos, int size, int height) {
  int a = xpos > 0 && ypos + size < height;
  if (size == 64) a = 0;
  if (size == 32 && ypos % 64 == 32) a = 0;
  if (size == 16 && (ypos % 64 == 48 || (ypos % 64 == 16 && xpos % 32 == 16))) a = 0;
  if (size == 8 && (ypos % 64 == 56 || (ypos % 16 == 8 && xpos % 16 == 8) || (ypos % 64 == 24 && xpos % 32 == 16))) a = 0;
  return a;
}
