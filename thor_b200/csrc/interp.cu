// libthorcuda: batched sub-pel inter prediction (luma 6-tap quarter-pel, chroma 4-tap eighth-pel,
// bi-prediction average) for a whole block list in one launch.
//
// Restates reference common/inter_prediction.c:72-180 (get_inter_prediction_chroma/_luma) and the
// truncating bi-pred average of enc/encode_block.c:1569-1588 / dec/decode_block.c:272-283.
//
// Work decomposition: the host splits every prediction block into tiles of at most 16x16 luma samples; one warp
// owns one tile (luma + both chroma planes).  All reference windows of the tile (up to 2 luma + 4 chroma) are
// fetched first, shifted so that the leftmost tap sits at byte 0 of a shared-memory row.  The separable filter
// runs HORIZONTAL first on packed bytes -- DP4A of a 4-byte window with the int8 taps -- into a transposed int16
// intermediate, then VERTICAL with DP2A on pairs of rows; because the reference keeps full precision between
// its two passes and rounds once at the end ((sum + 2048) >> 12), the order of the passes does not change the
// result.  A lane produces 2 rows x 4 samples and stores 32-bit words.
#include "internal.cuh"
#include "lists.cuh"

namespace thor {

// reference common/inter_prediction.c:47-70
__constant__ int8_t c_luma_bi[4][6] = {{0, 0, 64, 0, 0, 0}, {2, -10, 59, 17, -5, 1}, {1, -8, 39, 39, -8, 1}, {1, -5, 17, 59, -10, 2}};
__constant__ int8_t c_luma_uni[4][6] = {{0, 0, 64, 0, 0, 0}, {1, -7, 55, 19, -5, 1}, {1, -7, 38, 38, -7, 1}, {1, -5, 19, 55, -7, 1}};
__constant__ int8_t c_chroma[8][4] = {{0, 64, 0, 0}, {-2, 58, 10, -2}, {-4, 54, 16, -2}, {-4, 44, 28, -4},
                                      {-4, 36, 36, -4}, {-4, 28, 44, -4}, {-2, 16, 54, -4}, {-2, 10, 58, -2}};

constexpr int PW = 8;                 // warps per CTA
constexpr int LW_PITCH = 24;          // luma window row: 16 + 5 samples, taps start at byte 0
constexpr int LW_ROWS = 21;
constexpr int CW_PITCH = 12;          // chroma window row: 8 + 3 samples
constexpr int CW_ROWS = 11;
constexpr int PT = 22;                // int16 per column of the transposed intermediate (rows 0..th+4, +1 slack)

struct PredParams {
  const PicView *pics;                // device picture table, indexed by handle
  const uint16_t *map;                // list-local picture index -> handle
  const thorcuda_pred_task *tasks;
  const PredTile *tiles;
  int ntiles;
  int bipred_filter;
  int planes;
};

struct WarpSmem {
  uint32_t lwin[2][LW_ROWS * LW_PITCH / 4 + 2];     // luma windows of ref0 / ref1
  uint32_t cwin[4][CW_ROWS * CW_PITCH / 4 + 2];     // chroma windows U0, V0, U1, V1
  int16_t tmpT[16 * PT + 2];                        // transposed horizontal-pass output [column][row]
};

// 4 unsigned bytes (samples) x 4 signed bytes (taps) + c
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ uint32_t pack2(int lo, int hi) { return ((uint32_t)lo & 0xffu) | (((uint32_t)hi & 0xffu) << 8); }
__device__ __forceinline__ uint32_t taps4(int a, int b, int c, int d) {
  return ((uint32_t)a & 0xffu) | (((uint32_t)b & 0xffu) << 8) | (((uint32_t)c & 0xffu) << 16) | (((uint32_t)d & 0xffu) << 24);
}

// Fetch rows x bytes starting at src (any alignment) into win (row pitch `pitch` bytes, byte 0 = src[0]).
// wpr = words per staged row.
// LW = log2 of the lane columns used per row (8 for luma: wpr <= 6; 4 for chroma: wpr <= 3).
template <int LW>
__device__ __forceinline__ void stage_window(uint32_t *win, const uint8_t *src, int stride, int rows, int wpr, int pitch, int lane) {
  const int o = (int)((uintptr_t)src & 3);
  const uint8_t *base = src - o;
  const int sh = o * 8;
  const int wd = lane & ((1 << LW) - 1);
  if (wd >= wpr) return;
  for (int r = lane >> LW; r < rows; r += 32 >> LW) {
    const uint32_t *g = reinterpret_cast<const uint32_t *>(base + (size_t)r * stride) + wd;
    const uint32_t lo = g[0], hi = o ? g[1] : 0u;
    win[r * (pitch >> 2) + wd] = __funnelshift_r(lo, hi, sh);
  }
}

// ---- luma: one tile from one staged window -> out[2] (rows 2*rp, 2*rp+1; samples 4*xq .. 4*xq+3) ----
__device__ __forceinline__ void luma_tile(const uint32_t *win, int16_t *tmpT, int tw, int th, int fx, int fy, int bipred, int lane,
                                          uint32_t out[2]) {
  const int qpr = tw >> 2;                       // quads per row
  const int xq = lane & 3, rp = lane >> 2;       // this lane's output quad / row pair
  const bool act = xq < qpr && rp < (th >> 1);
  if (fx == 2 && fy == 2) {
    // centre position: [0 1 1 0; 1 2 2 1; 1 2 2 1; 0 1 1 0], (sum+8)>>4   (inter_prediction.c:146-157)
    // window byte 0 = sample x-2, window row 0 = sample row y-2; taps cover columns x-1..x+2, rows y-1..y+2
    if (act) {
#pragma unroll
      for (int k = 0; k < 2; k++) {
        const int r = 2 * rp + k;
        int s[4] = {0, 0, 0, 0};
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const uint32_t *w = win + (r + 1 + j) * (LW_PITCH / 4) + xq;
          const uint32_t a0 = w[0], a1 = w[1];
          const uint32_t t = (j == 0 || j == 3) ? 0x00010100u : 0x01020201u;
          s[0] = dp4a_us(__funnelshift_r(a0, a1, 8), t, s[0]);
          s[1] = dp4a_us(__funnelshift_r(a0, a1, 16), t, s[1]);
          s[2] = dp4a_us(__funnelshift_r(a0, a1, 24), t, s[2]);
          s[3] = dp4a_us(a1, t, s[3]);
        }
        out[k] = taps4(clip255((s[0] + 8) >> 4), clip255((s[1] + 8) >> 4), clip255((s[2] + 8) >> 4), clip255((s[3] + 8) >> 4));
      }
    }
    return;
  }
  const int8_t *kh = bipred ? c_luma_bi[fx] : c_luma_uni[fx];
  const int8_t *kv = bipred ? c_luma_bi[fy] : c_luma_uni[fy];
  // ---- horizontal pass: rows 0..th+4 of the window, 4 outputs per item, into tmpT[column][row] ----
  {
    const uint32_t hA = taps4(kh[0], kh[1], kh[2], kh[3]), hB = taps4(kh[4], kh[5], 0, 0);
    const int q = lane & 3;
    for (int r = lane >> 2; r < th + 5 && q < qpr; r += 8) {
      const uint32_t *w = win + r * (LW_PITCH / 4) + q;
      const uint32_t a0 = w[0], a1 = w[1], a2 = w[2];
      const int s0 = dp4a_us(a1, hB, dp4a_us(a0, hA, 0));
      const int s1 = dp4a_us(__funnelshift_r(a1, a2, 8), hB, dp4a_us(__funnelshift_r(a0, a1, 8), hA, 0));
      const int s2 = dp4a_us(__funnelshift_r(a1, a2, 16), hB, dp4a_us(__funnelshift_r(a0, a1, 16), hA, 0));
      const int s3 = dp4a_us(__funnelshift_r(a1, a2, 24), hB, dp4a_us(__funnelshift_r(a0, a1, 24), hA, 0));
      int16_t *t = tmpT + (4 * q) * PT + r;
      t[0] = (int16_t)s0; t[PT] = (int16_t)s1; t[2 * PT] = (int16_t)s2; t[3 * PT] = (int16_t)s3;
    }
  }
  __syncwarp();
  // ---- vertical pass: rows r = 2rp (even) and r+1; intermediate rows r .. r+7 are words 0..3 of the column ----
  if (act) {
    const uint32_t e0 = pack2(kv[0], kv[1]), e1 = pack2(kv[2], kv[3]), e2 = pack2(kv[4], kv[5]);
    const uint32_t o0 = pack2(0, kv[0]), o1 = pack2(kv[1], kv[2]), o2 = pack2(kv[3], kv[4]), o3 = pack2(kv[5], 0);
    int ve[4], vo[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const uint32_t *c = reinterpret_cast<const uint32_t *>(tmpT + (4 * xq + j) * PT) + rp;
      const uint32_t w0 = c[0], w1 = c[1], w2 = c[2], w3 = c[3];
      ve[j] = __dp2a_lo((int)w2, (int)e2, __dp2a_lo((int)w1, (int)e1, __dp2a_lo((int)w0, (int)e0, 0)));
      vo[j] = __dp2a_lo((int)w3, (int)o3, __dp2a_lo((int)w2, (int)o2, __dp2a_lo((int)w1, (int)o1, __dp2a_lo((int)w0, (int)o0, 0))));
    }
    out[0] = taps4(clip255((ve[0] + 2048) >> 12), clip255((ve[1] + 2048) >> 12), clip255((ve[2] + 2048) >> 12), clip255((ve[3] + 2048) >> 12));
    out[1] = taps4(clip255((vo[0] + 2048) >> 12), clip255((vo[1] + 2048) >> 12), clip255((vo[2] + 2048) >> 12), clip255((vo[3] + 2048) >> 12));
  }
  __syncwarp();
}

// ---- chroma: one tile (cw x ch <= 8x8) of one plane -> out[2] as for luma (rows 2rp, 2rp+1; samples 4xq..4xq+3) ----
__device__ __forceinline__ void chroma_tile(const uint32_t *win, int16_t *tmpT, int cw, int ch, int fx, int fy, int lane, uint32_t out[2]) {
  const int qpr = (cw + 3) >> 2;                 // 1 or 2
  const int xq = lane & 1, rp = lane >> 1;
  const bool act = xq < qpr && rp < (ch >> 1);
  const int8_t *kh = c_chroma[fx], *kv = c_chroma[fy];
  // horizontal pass over window rows 0..ch+2 (sample rows -1..ch+1); window byte 0 = sample x-1  (inter_prediction.c:97-106)
  {
    const uint32_t hA = taps4(kh[0], kh[1], kh[2], kh[3]);
    const int q = lane & 1;
    for (int r = lane >> 1; r < ch + 3 && q < qpr; r += 16) {
      const uint32_t *w = win + r * (CW_PITCH / 4) + q;
      const uint32_t a0 = w[0], a1 = w[1];
      int16_t *t = tmpT + (4 * q) * PT + r;
      t[0] = (int16_t)dp4a_us(a0, hA, 0);
      t[PT] = (int16_t)dp4a_us(__funnelshift_r(a0, a1, 8), hA, 0);
      t[2 * PT] = (int16_t)dp4a_us(__funnelshift_r(a0, a1, 16), hA, 0);
      t[3 * PT] = (int16_t)dp4a_us(__funnelshift_r(a0, a1, 24), hA, 0);
    }
  }
  __syncwarp();
  if (act) {
    const uint32_t e0 = pack2(kv[0], kv[1]), e1 = pack2(kv[2], kv[3]);
    const uint32_t o0 = pack2(0, kv[0]), o1 = pack2(kv[1], kv[2]), o2 = pack2(kv[3], 0);
    int ve[4], vo[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const uint32_t *c = reinterpret_cast<const uint32_t *>(tmpT + (4 * xq + j) * PT) + rp;
      const uint32_t w0 = c[0], w1 = c[1], w2 = c[2];
      ve[j] = __dp2a_lo((int)w1, (int)e1, __dp2a_lo((int)w0, (int)e0, 0));
      vo[j] = __dp2a_lo((int)w2, (int)o2, __dp2a_lo((int)w1, (int)o1, __dp2a_lo((int)w0, (int)o0, 0)));
    }
    out[0] = taps4(clip255((ve[0] + 2048) >> 12), clip255((ve[1] + 2048) >> 12), clip255((ve[2] + 2048) >> 12), clip255((ve[3] + 2048) >> 12));
    out[1] = taps4(clip255((vo[0] + 2048) >> 12), clip255((vo[1] + 2048) >> 12), clip255((vo[2] + 2048) >> 12), clip255((vo[3] + 2048) >> 12));
  }
  __syncwarp();
}

__device__ __forceinline__ uint32_t avg_trunc4(uint32_t a, uint32_t b) {   // per-byte (a+b)>>1
  return (a & b) + (((a ^ b) & 0xfefefefeu) >> 1);
}

__global__ void __launch_bounds__(PW * 32) k_predict(const PredParams P) {
  __shared__ __align__(16) WarpSmem sm[PW];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile = blockIdx.x * PW + warp;
  if (tile >= P.ntiles) return;
  WarpSmem &S = sm[warp];
  const uint4 *tp = reinterpret_cast<const uint4 *>(P.tiles + tile);
  const uint4 q0 = tp[0], q1 = tp[1];
  PredTile t;
  *reinterpret_cast<uint4 *>(&t) = q0;
  *(reinterpret_cast<uint4 *>(&t) + 1) = q1;
  const int x = t.x, y = t.y, tw = t.tw, th = t.th;
  const int mvx[2] = {t.mv0x, t.mv1x};
  const int mvy[2] = {t.mv0y, t.mv1y};
  const int nref = t.ref1 != 0xFFFF ? 2 : 1;
  const PicView dst = P.pics[t.dst];
  const int cx = x >> 1, cy = y >> 1, cw = tw >> 1, ch = th >> 1;

  // ---- fetch every window of the tile first (independent loads: their latencies overlap) ----
#pragma unroll
  for (int k = 0; k < 2; k++) {
    if (k >= nref) break;
    const PicView &r = P.pics[k ? t.ref1 : t.ref0];
    if (P.planes & 1) {
      const uint8_t *src = r.p[0] + (ptrdiff_t)(y + (mvy[k] >> 2) - 2) * r.stride[0] + (x + (mvx[k] >> 2) - 2);
      stage_window<3>(S.lwin[k], src, r.stride[0], th + 5, (tw + 5 + 3) >> 2, LW_PITCH, lane);
    }
    if (P.planes & 2) {
#pragma unroll
      for (int pl = 1; pl <= 2; pl++) {
        const uint8_t *src = r.p[pl] + (ptrdiff_t)(cy + (mvy[k] >> 3) - 1) * r.stride[pl] + (cx + (mvx[k] >> 3) - 1);
        stage_window<2>(S.cwin[2 * k + pl - 1], src, r.stride[pl], ch + 3, (cw + 3 + 3) >> 2, CW_PITCH, lane);
      }
    }
  }
  __syncwarp();

  if (P.planes & 1) {
    uint32_t a[2] = {0, 0}, b[2] = {0, 0};
    luma_tile(S.lwin[0], S.tmpT, tw, th, mvx[0] & 3, mvy[0] & 3, P.bipred_filter, lane, a);
    if (nref == 2) {
      luma_tile(S.lwin[1], S.tmpT, tw, th, mvx[1] & 3, mvy[1] & 3, P.bipred_filter, lane, b);
      a[0] = avg_trunc4(a[0], b[0]);
      a[1] = avg_trunc4(a[1], b[1]);
    }
    const int qpr = tw >> 2, xq = lane & 3, rp = lane >> 2;
    if (xq < qpr && rp < (th >> 1)) {
      uint8_t *d = dst.p[0] + (size_t)(y + 2 * rp) * dst.stride[0] + x + 4 * xq;
      *reinterpret_cast<uint32_t *>(d) = a[0];
      *reinterpret_cast<uint32_t *>(d + dst.stride[0]) = a[1];
    }
  }
  if (P.planes & 2) {
    const int qpr = (cw + 3) >> 2, xq = lane & 1, rp = lane >> 1;
#pragma unroll
    for (int pl = 1; pl <= 2; pl++) {
      uint32_t a[2] = {0, 0}, b[2] = {0, 0};
      chroma_tile(S.cwin[pl - 1], S.tmpT, cw, ch, mvx[0] & 7, mvy[0] & 7, lane, a);
      if (nref == 2) {
        chroma_tile(S.cwin[2 + pl - 1], S.tmpT, cw, ch, mvx[1] & 7, mvy[1] & 7, lane, b);
        a[0] = avg_trunc4(a[0], b[0]);
        a[1] = avg_trunc4(a[1], b[1]);
      }
      if (xq < qpr && rp < (ch >> 1)) {
        uint8_t *d = dst.p[pl] + (size_t)(cy + 2 * rp) * dst.stride[pl] + cx + 4 * xq;
        // chroma x is only 2-aligned and chroma widths are multiples of 2 (the last quad may be half): 16-bit stores
        *reinterpret_cast<uint16_t *>(d) = (uint16_t)a[0];
        *reinterpret_cast<uint16_t *>(d + dst.stride[pl]) = (uint16_t)a[1];
        if (4 * xq + 4 <= cw) {
          *reinterpret_cast<uint16_t *>(d + 2) = (uint16_t)(a[0] >> 16);
          *reinterpret_cast<uint16_t *>(d + dst.stride[pl] + 2) = (uint16_t)(a[1] >> 16);
        }
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
