// libthorcuda: batched sub-pel inter prediction (luma 6-tap quarter-pel, chroma 4-tap eighth-pel,
// bi-prediction average) for a whole block list in one launch.
//
// Restates reference common/inter_prediction.c:72-180 (get_inter_prediction_chroma/_luma) and the
// truncating bi-pred average of enc/encode_block.c:1569-1588 / dec/decode_block.c:272-283.
//
// Work decomposition: the host splits every prediction block into tiles of at most 16x16 luma samples and sorts
// them into two classes.  A tile larger than 8x8 is owned by one warp; tiles of at most 8x8 are owned by a group
// of 8 lanes, four to a warp (an 8x8 block would otherwise keep a quarter of the lanes busy).  All reference
// windows of the tile (up to 2 luma + 4 chroma) are brought in with 4-byte cp.async copies of the ALIGNED words that
// cover them -- every copy of the tile is in flight at once and no register is tied up -- and the byte offset of
// the window inside its first word is applied with a funnel shift when the words are read back.  The separable
// filter runs HORIZONTAL first on packed bytes -- DP4A of a 4-byte window with the int8 taps -- into a transposed
// int16 intermediate, then VERTICAL with DP2A on pairs of rows; because the reference keeps full precision between
// its two passes and rounds once at the end ((sum + 2048) >> 12), the order of the passes does not change the
// result.  A lane produces 2 rows x 4 samples.  The four chroma windows (U/V x ref0/ref1) are filtered side by
// side by the four quarters of the lane group; the bi-prediction average is one shuffle.
#include "internal.cuh"
#include "lists.cuh"

namespace thor {

// reference common/inter_prediction.c:47-70
__constant__ int8_t c_luma_bi[4][6] = {{0, 0, 64, 0, 0, 0}, {2, -10, 59, 17, -5, 1}, {1, -8, 39, 39, -8, 1}, {1, -5, 17, 59, -10, 2}};
__constant__ int8_t c_luma_uni[4][6] = {{0, 0, 64, 0, 0, 0}, {1, -7, 55, 19, -5, 1}, {1, -7, 38, 38, -7, 1}, {1, -5, 19, 55, -7, 1}};
__constant__ int8_t c_chroma[8][4] = {{0, 64, 0, 0}, {-2, 58, 10, -2}, {-4, 54, 16, -2}, {-4, 44, 28, -4},
                                      {-4, 36, 36, -4}, {-4, 28, 44, -4}, {-2, 16, 54, -4}, {-2, 10, 58, -2}};

constexpr int PW = 8;                 // warps per CTA

constexpr int imax(int a, int b) { return a > b ? a : b; }

// Geometry of a lane group of G lanes that owns one tile of at most T x T luma samples.
template <int G>
struct PredCfg {
  static constexpr int T = G == 32 ? 16 : 8;
  static constexpr int LR = T + 5;                      // luma window rows
  static constexpr int LW = (T + 5 + 3 + 3) / 4;        // aligned words that can hold a luma window row (offset <= 3)
  static constexpr int LP = LW + 1;                     // row pitch in words (+1: the shift reads one word further)
  static constexpr int CR = T / 2 + 3;                  // chroma window rows
  static constexpr int CW = (T / 2 + 3 + 3 + 3) / 4;
  static constexpr int CP = CW + 1;
  static constexpr int PT = LR + 1;                     // int16 per column of the transposed luma intermediate (even)
  static constexpr int PTC = CR + 1;                    // same for chroma
  static constexpr int QL = T / 4;                      // luma quads per row
  static constexpr int QC = T / 8;                      // chroma quads per row
  static constexpr int CG = G / 4;                      // lanes per chroma window
  static constexpr int LWIN = LR * LP, CWIN = CR * CP;  // words per window
  static constexpr int TMP = imax(T * PT, 4 * (T / 2) * PTC) / 2;   // words of the intermediate (luma and chroma alias)
  static constexpr int WORDS = 2 * LWIN + 4 * CWIN + TMP;
  static_assert(PT % 2 == 0 && PTC % 2 == 0, "columns of the intermediate are read as 32-bit words");
};
constexpr int WARP_WORDS = imax(PredCfg<32>::WORDS, 4 * PredCfg<8>::WORDS);

struct PredParams {
  const PicView *pics;                // device picture table, indexed by handle
  const PredTile *tiles;              // tiles larger than 8x8 first (nbig of them), then the small ones
  int ntiles, nbig;
  int bipred_filter;
  int planes;
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
__device__ __forceinline__ uint32_t round12x4(const int v[4]) {
  return taps4(clip255((v[0] + 2048) >> 12), clip255((v[1] + 2048) >> 12), clip255((v[2] + 2048) >> 12), clip255((v[3] + 2048) >> 12));
}
__device__ __forceinline__ void cp_async4(uint32_t *smem_dst, const void *gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}

// Queue the copies of the aligned words that cover rows x bytes starting at src (any alignment) into win (row pitch
// PITCH words; word 0 = the aligned word that holds src[0]).  NW = most words a row can need.  Returns the shift (bits).
template <int G, int NW, int PITCH>
__device__ __forceinline__ int stage_window(uint32_t *win, const uint8_t *src, int stride, int rows, int bytes, int gl) {
  const int o = (int)((uintptr_t)src & 3);
  const uint8_t *base = src - o;
  const int need = o + bytes;                     // bytes of the aligned row that are wanted
  for (int i = gl; i < rows * NW; i += G) {
    const int r = i / NW, wd = i - r * NW;
    if (4 * wd < need) cp_async4(win + r * PITCH + wd, base + (size_t)r * stride + 4 * wd);
  }
  return o * 8;
}

// ---- luma: one tile from one staged window -> out[2] (rows 2*rp, 2*rp+1; samples 4*xq .. 4*xq+3) ----
template <int G>
__device__ __forceinline__ void luma_tile(const uint32_t *win, int sh, int16_t *tmpT, int tw, int th, int fx, int fy, int bipred, int gl,
                                          unsigned gmask, uint32_t out[2]) {
  using C = PredCfg<G>;
  constexpr int LP = C::LP, PT = C::PT, QL = C::QL;
  const int qpr = tw >> 2;                       // quads per row
  const int xq = gl % QL, rp = gl / QL;          // this lane's output quad / row pair
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
          const uint32_t *w = win + (r + 1 + j) * LP + xq;
          const uint32_t r0 = w[0], r1 = w[1], r2 = w[2];
          const uint32_t a0 = __funnelshift_r(r0, r1, sh), a1 = __funnelshift_r(r1, r2, sh);
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
    for (int i = gl; i < (th + 5) * QL; i += G) {
      const int r = i / QL, q = i - r * QL;
      if (q >= qpr) continue;
      const uint32_t *w = win + r * LP + q;
      const uint32_t r0 = w[0], r1 = w[1], r2 = w[2], r3 = w[3];
      const uint32_t a0 = __funnelshift_r(r0, r1, sh), a1 = __funnelshift_r(r1, r2, sh), a2 = __funnelshift_r(r2, r3, sh);
      const int s0 = dp4a_us(a1, hB, dp4a_us(a0, hA, 0));
      const int s1 = dp4a_us(__funnelshift_r(a1, a2, 8), hB, dp4a_us(__funnelshift_r(a0, a1, 8), hA, 0));
      const int s2 = dp4a_us(__funnelshift_r(a1, a2, 16), hB, dp4a_us(__funnelshift_r(a0, a1, 16), hA, 0));
      const int s3 = dp4a_us(__funnelshift_r(a1, a2, 24), hB, dp4a_us(__funnelshift_r(a0, a1, 24), hA, 0));
      int16_t *t = tmpT + (4 * q) * PT + r;
      t[0] = (int16_t)s0; t[PT] = (int16_t)s1; t[2 * PT] = (int16_t)s2; t[3 * PT] = (int16_t)s3;
    }
  }
  __syncwarp(gmask);
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
    out[0] = round12x4(ve);
    out[1] = round12x4(vo);
  }
  __syncwarp(gmask);
}

// ---- chroma: the CG lanes cl = 0..CG-1 filter one window (cw x ch <= T/2 x T/2) -> out[2] as for luma ----
template <int G>
__device__ __forceinline__ void chroma_tile(const uint32_t *win, int sh, int16_t *tmpT, int cw, int ch, int fx, int fy, int cl, bool on,
                                            unsigned gmask, uint32_t out[2]) {
  using C = PredCfg<G>;
  constexpr int CP = C::CP, PTC = C::PTC, QC = C::QC, CG = C::CG;
  const int qpr = (cw + 3) >> 2;
  const int xq = cl % QC, rp = cl / QC;
  const bool act = on && xq < qpr && rp < (ch >> 1);
  const int8_t *kh = c_chroma[fx], *kv = c_chroma[fy];
  // horizontal pass over window rows 0..ch+2 (sample rows -1..ch+1); window byte 0 = sample x-1  (inter_prediction.c:97-106)
  if (on) {
    const uint32_t hA = taps4(kh[0], kh[1], kh[2], kh[3]);
    for (int i = cl; i < (ch + 3) * QC; i += CG) {
      const int r = i / QC, q = i - r * QC;
      if (q >= qpr) continue;
      const uint32_t *w = win + r * CP + q;
      const uint32_t r0 = w[0], r1 = w[1], r2 = w[2];
      const uint32_t a0 = __funnelshift_r(r0, r1, sh), a1 = __funnelshift_r(r1, r2, sh);
      int16_t *t = tmpT + (4 * q) * PTC + r;
      t[0] = (int16_t)dp4a_us(a0, hA, 0);
      t[PTC] = (int16_t)dp4a_us(__funnelshift_r(a0, a1, 8), hA, 0);
      t[2 * PTC] = (int16_t)dp4a_us(__funnelshift_r(a0, a1, 16), hA, 0);
      t[3 * PTC] = (int16_t)dp4a_us(__funnelshift_r(a0, a1, 24), hA, 0);
    }
  }
  __syncwarp(gmask);
  if (act) {
    const uint32_t e0 = pack2(kv[0], kv[1]), e1 = pack2(kv[2], kv[3]);
    const uint32_t o0 = pack2(0, kv[0]), o1 = pack2(kv[1], kv[2]), o2 = pack2(kv[3], 0);
    int ve[4], vo[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const uint32_t *c = reinterpret_cast<const uint32_t *>(tmpT + (4 * xq + j) * PTC) + rp;
      const uint32_t w0 = c[0], w1 = c[1], w2 = c[2];
      ve[j] = __dp2a_lo((int)w1, (int)e1, __dp2a_lo((int)w0, (int)e0, 0));
      vo[j] = __dp2a_lo((int)w2, (int)o2, __dp2a_lo((int)w1, (int)o1, __dp2a_lo((int)w0, (int)o0, 0)));
    }
    out[0] = round12x4(ve);
    out[1] = round12x4(vo);
  }
  __syncwarp(gmask);
}

__device__ __forceinline__ uint32_t avg_trunc4(uint32_t a, uint32_t b) {   // per-byte (a+b)>>1
  return (a & b) + (((a ^ b) & 0xfefefefeu) >> 1);
}

// One tile by the G lanes gl = 0..G-1 of a group; `valid` gates the stores (a padding group repeats the last tile).
template <int G>
__device__ __forceinline__ void predict_tile(const PredParams &P, int tile, bool valid, uint32_t *S, int gl) {
  using C = PredCfg<G>;
  // groups of a warp work on different tiles and diverge (phase, uni/bi): every barrier and shuffle names its group only
  const unsigned gmask = G == 32 ? 0xffffffffu : 0xffu << ((threadIdx.x & 31) & 24);
  uint32_t *lwin = S, *cwin = S + 2 * C::LWIN;
  int16_t *tmp = reinterpret_cast<int16_t *>(S + 2 * C::LWIN + 4 * C::CWIN);
  const uint4 *tp = reinterpret_cast<const uint4 *>(P.tiles + tile);
  const uint4 q0 = tp[0], q1 = tp[1];
  PredTile t;
  *reinterpret_cast<uint4 *>(&t) = q0;
  *(reinterpret_cast<uint4 *>(&t) + 1) = q1;
  const int x = t.x, y = t.y, tw = t.tw, th = t.th;
  const int mvx[2] = {t.mv0x, t.mv1x};
  const int mvy[2] = {t.mv0y, t.mv1y};
  const int nref = t.ref1 != 0xFFFF ? 2 : 1;
  const int cx = x >> 1, cy = y >> 1, cw = tw >> 1, ch = th >> 1;

  // ---- queue every window of the tile ----
  int lsh[2] = {0, 0}, csh[2] = {0, 0};
#pragma unroll
  for (int k = 0; k < 2; k++) {
    if (k >= nref) break;
    const PicView &r = P.pics[k ? t.ref1 : t.ref0];
    if (P.planes & 1) {
      const uint8_t *src = r.p[0] + (ptrdiff_t)(y + (mvy[k] >> 2) - 2) * r.stride[0] + (x + (mvx[k] >> 2) - 2);
      lsh[k] = stage_window<G, C::LW, C::LP>(lwin + k * C::LWIN, src, r.stride[0], th + 5, tw + 5, gl);
    }
    if (P.planes & 2) {
#pragma unroll
      for (int pl = 1; pl <= 2; pl++) {
        const uint8_t *src = r.p[pl] + (ptrdiff_t)(cy + (mvy[k] >> 3) - 1) * r.stride[pl] + (cx + (mvx[k] >> 3) - 1);
        // U and V planes share geometry (same stride and alignment), so one shift per reference serves both
        csh[k] = stage_window<G, C::CW, C::CP>(cwin + (2 * k + pl - 1) * C::CWIN, src, r.stride[pl], ch + 3, cw + 3, gl);
      }
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  const PicView &dst = P.pics[t.dst];
  uint8_t *const dp0 = dst.p[0], *const dp1 = dst.p[1], *const dp2 = dst.p[2];
  const int ds0 = dst.stride[0], ds1 = dst.stride[1];
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncwarp(gmask);

  if (P.planes & 1) {
    uint32_t a[2] = {0, 0}, b[2] = {0, 0};
    luma_tile<G>(lwin, lsh[0], tmp, tw, th, mvx[0] & 3, mvy[0] & 3, P.bipred_filter, gl, gmask, a);
    if (nref == 2) {
      luma_tile<G>(lwin + C::LWIN, lsh[1], tmp, tw, th, mvx[1] & 3, mvy[1] & 3, P.bipred_filter, gl, gmask, b);
      a[0] = avg_trunc4(a[0], b[0]);
      a[1] = avg_trunc4(a[1], b[1]);
    }
    const int qpr = tw >> 2, xq = gl % C::QL, rp = gl / C::QL;
    if (valid && xq < qpr && rp < (th >> 1)) {
      uint8_t *d = dp0 + (size_t)(y + 2 * rp) * ds0 + x + 4 * xq;
      *reinterpret_cast<uint32_t *>(d) = a[0];
      *reinterpret_cast<uint32_t *>(d + ds0) = a[1];
    }
  }
  if (P.planes & 2) {
    // quarter s of the group: plane 1 + (s & 1) of reference s >> 1
    const int sgrp = gl / C::CG, cl = gl % C::CG;
    const int pl = 1 + (sgrp & 1), k = sgrp >> 1;
    uint32_t a[2] = {0, 0};
    const int cmx = k ? mvx[1] : mvx[0], cmy = k ? mvy[1] : mvy[0];
    chroma_tile<G>(cwin + sgrp * C::CWIN, k ? csh[1] : csh[0], tmp + sgrp * (C::T / 2) * C::PTC, cw, ch, cmx & 7, cmy & 7, cl, k < nref, gmask, a);
    const uint32_t b0 = __shfl_xor_sync(gmask, a[0], 2 * C::CG), b1 = __shfl_xor_sync(gmask, a[1], 2 * C::CG);
    if (nref == 2) {
      a[0] = avg_trunc4(a[0], b0);
      a[1] = avg_trunc4(a[1], b1);
    }
    const int qpr = (cw + 3) >> 2, xq = cl % C::QC, rp = cl / C::QC;
    if (valid && k == 0 && xq < qpr && rp < (ch >> 1)) {
      uint8_t *d = (pl == 1 ? dp1 : dp2) + (size_t)(cy + 2 * rp) * ds1 + cx + 4 * xq;
      // chroma x is only 2-aligned and chroma widths are multiples of 2 (the last quad may be half): 16-bit stores
      *reinterpret_cast<uint16_t *>(d) = (uint16_t)a[0];
      *reinterpret_cast<uint16_t *>(d + ds1) = (uint16_t)a[1];
      if (4 * xq + 4 <= cw) {
        *reinterpret_cast<uint16_t *>(d + 2) = (uint16_t)(a[0] >> 16);
        *reinterpret_cast<uint16_t *>(d + ds1 + 2) = (uint16_t)(a[1] >> 16);
      }
    }
  }
}

__global__ void __launch_bounds__(PW * 32) k_predict(const PredParams P) {
  __shared__ __align__(16) uint32_t sm[PW][WARP_WORDS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gw = blockIdx.x * PW + warp;
  if (gw < P.nbig) {
    predict_tile<32>(P, gw, true, sm[warp], lane);
  } else {
    const int first = P.nbig + (gw - P.nbig) * 4;
    if (first >= P.ntiles) return;
    const int tile = first + (lane >> 3);
    predict_tile<8>(P, min(tile, P.ntiles - 1), tile < P.ntiles, sm[warp] + (lane >> 3) * PredCfg<8>::WORDS, lane & 7);
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
  if (list_stale(L)) return -1;
  PredParams P;
  P.pics = c.d_pics;
  P.tiles = (const PredTile *)L->tiles();
  P.ntiles = L->ntiles;
  P.nbig = L->nbig;
  P.bipred_filter = bipred_filter;
  P.planes = planes;
  const int warps = L->nbig + (L->ntiles - L->nbig + 3) / 4;
  static bool carveout_set = false;
  if (!carveout_set) {   // 39 KB per CTA: ask for the large shared-memory split so that registers bound the occupancy
    cudaFuncSetAttribute(k_predict, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    carveout_set = true;
  }
  k_predict<<<(unsigned)((warps + PW - 1) / PW), PW * 32, 0, c.stream>>>(P);
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
