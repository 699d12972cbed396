// libthorcuda: batched sub-pel inter prediction (luma 6-tap quarter-pel, chroma 4-tap eighth-pel,
// bi-prediction average) for a whole block list in one launch.
//
// Restates reference common/inter_prediction.c:72-180 (get_inter_prediction_chroma/_luma) and the
// truncating bi-pred average of enc/encode_block.c:1569-1588 / dec/decode_block.c:272-283.
//
// Work decomposition: the host splits every prediction block into tiles of at most 16x16 luma samples;
// one warp owns one tile (luma + both chroma planes).  The reference window of the tile is staged in
// shared memory with aligned 32-bit loads, the separable filter runs vertical-then-horizontal exactly
// like the reference (single rounding at the end), results leave as 32-bit words.
#include "internal.cuh"
#include "lists.cuh"

namespace thor {

// reference common/inter_prediction.c:47-70
__constant__ int8_t c_luma_bi[4][6] = {{0, 0, 64, 0, 0, 0}, {2, -10, 59, 17, -5, 1}, {1, -8, 39, 39, -8, 1}, {1, -5, 17, 59, -10, 2}};
__constant__ int8_t c_luma_uni[4][6] = {{0, 0, 64, 0, 0, 0}, {1, -7, 55, 19, -5, 1}, {1, -7, 38, 38, -7, 1}, {1, -5, 19, 55, -7, 1}};
__constant__ int8_t c_chroma[8][4] = {{0, 64, 0, 0}, {-2, 58, 10, -2}, {-4, 54, 16, -2}, {-4, 44, 28, -4},
                                      {-4, 36, 36, -4}, {-4, 28, 44, -4}, {-2, 16, 54, -4}, {-2, 10, 58, -2}};

constexpr int PW = 8;                 // warps per CTA
constexpr int WIN_PITCH = 24;         // bytes per staged window row (21 + up to 3 alignment bytes)
constexpr int WIN_ROWS = 21;
constexpr int TMP_PITCH = 22;         // int16 per row of the vertical-pass result

struct PredParams {
  const PicView *pics;                // device picture table, indexed by handle
  const uint16_t *map;                // list-local picture index -> handle
  const thorcuda_pred_task *tasks;
  const PredTile *tiles;
  int ntiles;
  int bipred_filter;
  int planes;
};

// Stage rows [0,rows) x bytes [0,cols) starting at src (any alignment) into win (pitch WIN_PITCH);
// returns the byte offset of src inside the first staged word.
__device__ __forceinline__ int stage_window(uint8_t *win, const uint8_t *src, int stride, int rows, int cols, int lane) {
  const int o = (int)((uintptr_t)src & 3);
  const uint8_t *base = src - o;
  const int words = (cols + o + 3) >> 2;          // <= 6
  for (int i = lane; i < rows * words; i += 32) {
    const int r = i / words, wd = i - r * words;
    reinterpret_cast<uint32_t *>(win + r * WIN_PITCH)[wd] = *reinterpret_cast<const uint32_t *>(base + (size_t)r * stride + 4 * wd);
  }
  return o;
}

// One luma tile from one reference.  out[2]: each lane owns up to two groups of 4 horizontally adjacent
// samples (group g = lane + 32*k, row = g / (tw/4), col = 4*(g % (tw/4))).
__device__ __forceinline__ void luma_tile(const PicView &ref, int x, int y, int tw, int th, int mvx, int mvy, int bipred,
                                          uint8_t *win, int16_t *tmp, int lane, uint32_t out[2]) {
  const int fx = mvx & 3, fy = mvy & 3;
  const uint8_t *src = ref.p[0] + (ptrdiff_t)(y + (mvy >> 2)) * ref.stride[0] + (x + (mvx >> 2));
  const int gpr = tw >> 2, ngroups = gpr * th;
  if (!fx && !fy) {
    // integer position: straight copy (memcpy in the reference)
    const int o = stage_window(win, src, ref.stride[0], th, tw, lane);
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 2; k++) {
      const int g = lane + 32 * k;
      if (g < ngroups) {
        const uint8_t *p = win + (g / gpr) * WIN_PITCH + o + 4 * (g % gpr);
        out[k] = p[0] | (p[1] << 8) | (p[2] << 16) | (p[3] << 24);
      }
    }
    __syncwarp();
    return;
  }
  // window rows -2..th+2, cols -2..tw+2
  const int o = stage_window(win, src - 2 * ref.stride[0] - 2, ref.stride[0], th + 5, tw + 5, lane);
  __syncwarp();
  if (fx == 2 && fy == 2) {
    // centre position: [0 1 1 0; 1 2 2 1; 1 2 2 1; 0 1 1 0], (sum+8)>>4   (inter_prediction.c:146-157)
#pragma unroll
    for (int k = 0; k < 2; k++) {
      const int g = lane + 32 * k;
      if (g < ngroups) {
        const int r = g / gpr, c = 4 * (g % gpr);
        uint32_t w = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const uint8_t *p = win + (r + 1) * WIN_PITCH + o + c + j + 1;      // sample (r-1, c+j-1) of the block
          const int s = p[1] + p[2] + p[WIN_PITCH] + 2 * p[WIN_PITCH + 1] + 2 * p[WIN_PITCH + 2] + p[WIN_PITCH + 3] +
                        p[2 * WIN_PITCH] + 2 * p[2 * WIN_PITCH + 1] + 2 * p[2 * WIN_PITCH + 2] + p[2 * WIN_PITCH + 3] +
                        p[3 * WIN_PITCH + 1] + p[3 * WIN_PITCH + 2];
          w |= (uint32_t)clip255((s + 8) >> 4) << (8 * j);
        }
        out[k] = w;
      }
    }
    __syncwarp();
    return;
  }
  const int8_t *kv = bipred ? c_luma_bi[fy] : c_luma_uni[fy];
  const int8_t *kh = bipred ? c_luma_bi[fx] : c_luma_uni[fx];
  const int v0 = kv[0], v1 = kv[1], v2 = kv[2], v3 = kv[3], v4 = kv[4], v5 = kv[5];
  const int h0 = kh[0], h1 = kh[1], h2 = kh[2], h3 = kh[3], h4 = kh[4], h5 = kh[5];
  // vertical pass: item = (column c in 0..tw+4, group of 4 rows); |sum| <= 94*255 fits int16
  const int ncols = tw + 5, rgroups = th >> 2;
  for (int i = lane; i < ncols * rgroups; i += 32) {
    const int c = i % ncols, r0 = (i / ncols) * 4;
    const uint8_t *p = win + r0 * WIN_PITCH + o + c;
    int a0 = p[0], a1 = p[WIN_PITCH], a2 = p[2 * WIN_PITCH], a3 = p[3 * WIN_PITCH], a4 = p[4 * WIN_PITCH];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int a5 = p[(j + 5) * WIN_PITCH];
      tmp[(r0 + j) * TMP_PITCH + c] = (int16_t)(v0 * a0 + v1 * a1 + v2 * a2 + v3 * a3 + v4 * a4 + v5 * a5);
      a0 = a1; a1 = a2; a2 = a3; a3 = a4; a4 = a5;
    }
  }
  __syncwarp();
  // horizontal pass
#pragma unroll
  for (int k = 0; k < 2; k++) {
    const int g = lane + 32 * k;
    if (g < ngroups) {
      const int r = g / gpr, c = 4 * (g % gpr);
      const int16_t *t = tmp + r * TMP_PITCH + c;
      int a0 = t[0], a1 = t[1], a2 = t[2], a3 = t[3], a4 = t[4];
      uint32_t w = 0;
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int a5 = t[j + 5];
        const int s = h0 * a0 + h1 * a1 + h2 * a2 + h3 * a3 + h4 * a4 + h5 * a5;
        w |= (uint32_t)clip255((s + 2048) >> 12) << (8 * j);
        a0 = a1; a1 = a2; a2 = a3; a3 = a4; a4 = a5;
      }
      out[k] = w;
    }
  }
  __syncwarp();
}

// One chroma tile (cw x ch <= 8x8) of one plane from one reference; each lane owns up to 2 samples
// (sample s = lane + 32*k, row = s / cw, col = s % cw).
__device__ __forceinline__ void chroma_tile(const uint8_t *plane, int stride, int cx, int cy, int cw, int ch, int mvx, int mvy,
                                            uint8_t *win, int16_t *tmp, int lane, int out[2]) {
  const int fx = mvx & 7, fy = mvy & 7;
  const uint8_t *src = plane + (ptrdiff_t)(cy + (mvy >> 3)) * stride + (cx + (mvx >> 3));
  const int n = cw * ch;
  if (!fx && !fy) {
    const int o = stage_window(win, src, stride, ch, cw, lane);
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 2; k++) {
      const int s = lane + 32 * k;
      if (s < n) out[k] = win[(s / cw) * WIN_PITCH + o + (s % cw)];
    }
    __syncwarp();
    return;
  }
  const int o = stage_window(win, src - stride - 1, stride, ch + 3, cw + 3, lane);
  __syncwarp();
  const int8_t *kh = c_chroma[fx], *kv = c_chroma[fy];
  // horizontal pass over rows -1..ch+1 (inter_prediction.c:97-106); |sum| <= 72*255 fits int16
  for (int i = lane; i < (ch + 3) * cw; i += 32) {
    const int r = i / cw, c = i % cw;
    const uint8_t *p = win + r * WIN_PITCH + o + c;
    tmp[r * TMP_PITCH + c] = (int16_t)(kh[0] * p[0] + kh[1] * p[1] + kh[2] * p[2] + kh[3] * p[3]);
  }
  __syncwarp();
#pragma unroll
  for (int k = 0; k < 2; k++) {
    const int s = lane + 32 * k;
    if (s < n) {
      const int16_t *t = tmp + (s / cw) * TMP_PITCH + (s % cw);
      const int acc = kv[0] * t[0] + kv[1] * t[TMP_PITCH] + kv[2] * t[2 * TMP_PITCH] + kv[3] * t[3 * TMP_PITCH];
      out[k] = clip255((acc + 2048) >> 12);
    }
  }
  __syncwarp();
}

__device__ __forceinline__ uint32_t avg_trunc4(uint32_t a, uint32_t b) {   // per-byte (a+b)>>1
  return (a & b) + (((a ^ b) & 0xfefefefeu) >> 1);
}

__global__ void __launch_bounds__(PW * 32) k_predict(const PredParams P) {
  __shared__ __align__(16) uint8_t sWin[PW][WIN_ROWS * WIN_PITCH];
  __shared__ __align__(16) int16_t sTmp[PW][16 * TMP_PITCH];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile = blockIdx.x * PW + warp;
  if (tile >= P.ntiles) return;
  const PredTile tl = P.tiles[tile];
  const thorcuda_pred_task t = P.tasks[tl.task];
  const int x = t.x + 16 * tl.tx, y = t.y + 16 * tl.ty;
  const int tw = min(16, t.w - 16 * tl.tx), th = min(16, t.h - 16 * tl.ty);
  uint8_t *win = sWin[warp];
  int16_t *tmp = sTmp[warp];
  const int mv0x = t.sign0 ? -t.mv0x : t.mv0x, mv0y = t.sign0 ? -t.mv0y : t.mv0y;
  const int mv1x = t.sign1 ? -t.mv1x : t.mv1x, mv1y = t.sign1 ? -t.mv1y : t.mv1y;
  const bool bi = t.ref1 >= 0;
  const PicView dst = P.pics[P.map[t.dst]];
  const PicView r0 = P.pics[P.map[t.ref0]];
  const PicView r1 = P.pics[P.map[bi ? t.ref1 : t.ref0]];

  if (P.planes & 1) {
    uint32_t a[2] = {0, 0}, b[2] = {0, 0};
    luma_tile(r0, x, y, tw, th, mv0x, mv0y, P.bipred_filter, win, tmp, lane, a);
    if (bi) {
      luma_tile(r1, x, y, tw, th, mv1x, mv1y, P.bipred_filter, win, tmp, lane, b);
      a[0] = avg_trunc4(a[0], b[0]);
      a[1] = avg_trunc4(a[1], b[1]);
    }
    const int gpr = tw >> 2;
#pragma unroll
    for (int k = 0; k < 2; k++) {
      const int g = lane + 32 * k;
      if (g < gpr * th)
        *reinterpret_cast<uint32_t *>(dst.p[0] + (size_t)(y + g / gpr) * dst.stride[0] + x + 4 * (g % gpr)) = a[k];
    }
  }
  if (P.planes & 2) {
    const int cx = x >> 1, cy = y >> 1, cw = tw >> 1, ch = th >> 1;
    for (int pl = 1; pl <= 2; pl++) {
      int a[2] = {0, 0}, b[2] = {0, 0};
      chroma_tile(r0.p[pl], r0.stride[pl], cx, cy, cw, ch, mv0x, mv0y, win, tmp, lane, a);
      if (bi) {
        chroma_tile(r1.p[pl], r1.stride[pl], cx, cy, cw, ch, mv1x, mv1y, win, tmp, lane, b);
        a[0] = (a[0] + b[0]) >> 1;
        a[1] = (a[1] + b[1]) >> 1;
      }
#pragma unroll
      for (int k = 0; k < 2; k++) {
        const int s = lane + 32 * k;
        if (s < cw * ch) dst.p[pl][(size_t)(cy + s / cw) * dst.stride[pl] + cx + (s % cw)] = (uint8_t)a[k];
      }
    }
  }
}

}  // namespace thor

using namespace thor;

extern "C" {

int thorcuda_predict_list(thorcuda_list hl, int bipred_filter, int planes) {
  THOR_NEED_CTX();
  Context &c = ctx();
  List *L = get_list(hl);
  if (!L || L->kind != THORCUDA_LIST_PRED) return fail("predict: not a prediction list");
  PredParams P;
  P.pics = c.d_pics;
  P.map = L->map();
  P.tasks = (const thorcuda_pred_task *)L->tasks();
  P.tiles = (const PredTile *)L->tiles();
  P.ntiles = L->ntiles;
  P.bipred_filter = bipred_filter;
  P.planes = planes;
  k_predict<<<(unsigned)((L->ntiles + PW - 1) / PW), PW * 32, 0, c.stream>>>(P);
  THOR_LAUNCHED();
  return 0;
}

int thorcuda_predict_blocks(const thorcuda_pic *pics, int npics, const thorcuda_pred_task *tasks, int ntasks,
                            int bipred_filter, int planes) {
  if (ntasks <= 0) return 0;
  const thorcuda_list l = thorcuda_list_create(THORCUDA_LIST_PRED, pics, npics, tasks, ntasks);
  if (l < 0) return -1;
  const int rc = thorcuda_predict_list(l, bipred_filter, planes);
  thorcuda_list_destroy(l);
  return rc;
}

}  // extern "C"
