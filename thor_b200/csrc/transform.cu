// libthorcuda: batched residual -> integer DCT -> quantise -> dequantise -> inverse DCT -> reconstruct
// (encoder leg) and dequantise -> inverse DCT -> reconstruct (decoder leg) for whole transform-block lists.
//
// Restates (bit-exactly) reference common/transform.c:249-330 (transform, incl. the `fast` box-sum path and
// the 64->32 down-sampling), :432-518 (inverse_transform, incl. 2x up-sampling for 64), enc/encode_block.c:75-178
// (quantize with rdoq == 0), :484-493 (get_residual), :782-797 (ssd_calc), common/common_block.c:132-156
// (dequantize, reconstruct_block), chained as encode_and_reconstruct_block_inter (enc/encode_block.c:1469-1532)
// / decode_and_reconstruct_block_inter (dec/decode_block.c:90-120).
//
// Work decomposition: ONE launch for a whole list.  Tasks are sorted by transform size at list creation; the
// grid is cut into per-size ranges, largest blocks first (their long units start early, the small ones fill
// in).  A unit (one transform block) is processed by a group of G lanes: 32 for N >= 16, 8 for N = 8, 4 for
// N = 4, so a warp handles 1 / 4 / 8 blocks at a time; everything between the pixel loads and the pixel stores
// stays in shared memory.
// The transforms are integer matrix products against the HEVC bases, evaluated with DP2A (two 16-bit x 8-bit
// MACs per instruction: every basis coefficient fits in int8, every operand in int16), 32-bit accumulate, then
// round, shift and store to int16 after each pass exactly like the reference; the reference's butterflies are
// an exact factorisation of the same products.
#include <string.h>
#include "internal.cuh"

namespace thor {

// Bases as int8.  Forward: c_fw[off(n)] = rows k < min(n,16) of the n-point basis, row pitch n + 4 bytes.
// Inverse: c_iv[off(n)] = TRANSPOSED basis, row r (sample) holds frequencies v < min(n,16), pitch min(n,16) + 4.
constexpr int FW_OFF4 = 0, FW_OFF8 = FW_OFF4 + 4 * 8, FW_OFF16 = FW_OFF8 + 8 * 12, FW_OFF32 = FW_OFF16 + 16 * 20, FW_BYTES = FW_OFF32 + 16 * 36;
constexpr int IV_OFF4 = 0, IV_OFF8 = IV_OFF4 + 4 * 8, IV_OFF16 = IV_OFF8 + 8 * 12, IV_OFF32 = IV_OFF16 + 16 * 20, IV_BYTES = IV_OFF32 + 32 * 20;
__constant__ __align__(16) int8_t c_fw[FW_BYTES];
__constant__ __align__(16) int8_t c_iv[IV_BYTES];
__constant__ __align__(16) uint8_t c_zz[16 + 64 + 256]; // zig-zag scan position of coefficient (i,j), for q = 4, 8, 16
static bool g_tables_ready = false;

__host__ __device__ constexpr int fw_off(int n) { return n == 4 ? FW_OFF4 : n == 8 ? FW_OFF8 : n == 16 ? FW_OFF16 : FW_OFF32; }
__host__ __device__ constexpr int iv_off(int n) { return n == 4 ? IV_OFF4 : n == 8 ? IV_OFF8 : n == 16 ? IV_OFF16 : IV_OFF32; }

// common/common_block.c:97-98
__constant__ uint16_t c_qscale[6] = {26214, 23302, 20560, 18396, 16384, 14564};
__constant__ uint16_t c_dqscale[6] = {40, 45, 51, 57, 64, 72};

static int upload_tables() {
  if (g_tables_ready) return 0;
  // cos(pi*a/64) ladder of the HEVC integer DCT (reference common/transform.c:41-245 lists the expanded matrices);
  // the n-point basis is every (32/n)-th row of the 32-point one restricted to its first n columns
  static const int8_t cos32[33] = {64, 90, 90, 90, 89, 88, 87, 85, 83, 82, 80, 78, 75, 73, 70, 67, 64, 61, 57, 54, 50, 46, 43, 38, 36, 31, 25, 22, 18, 13, 9, 4, 0};
  auto m32 = [&](int k, int n) {
    int a = ((2 * n + 1) * k) & 127;
    if (a > 64) a = 128 - a;
    return (int8_t)(a > 32 ? -cos32[64 - a] : cos32[a]);
  };
  int8_t fw[FW_BYTES], iv[IV_BYTES];
  memset(fw, 0, sizeof(fw));
  memset(iv, 0, sizeof(iv));
  for (int n = 4; n <= 32; n <<= 1) {
    const int q = n < 16 ? n : 16, step = 32 / n;
    for (int k = 0; k < q; k++)
      for (int c = 0; c < n; c++) {
        fw[fw_off(n) + k * (n + 4) + c] = m32(k * step, c);
        iv[iv_off(n) + c * (q + 4) + k] = m32(k * step, c);
      }
  }
  uint8_t zz[16 + 64 + 256];
  int base = 0;
  for (int q = 4; q <= 16; q <<= 1) {     // reference common/common_block.c:38-73
    int idx = 0;
    for (int s = 0; s <= 2 * q - 2; s++)
      for (int t = 0; t <= s; t++) {
        const int i = (s & 1) ? t : s - t, j = s - i;
        if (i < q && j < q) zz[base + i * q + j] = (uint8_t)idx++;
      }
    base += q * q;
  }
  THOR_CUDA(cudaMemcpyToSymbolAsync(c_fw, fw, sizeof(fw), 0, cudaMemcpyHostToDevice, ctx().stream));
  THOR_CUDA(cudaMemcpyToSymbolAsync(c_iv, iv, sizeof(iv), 0, cudaMemcpyHostToDevice, ctx().stream));
  THOR_CUDA(cudaMemcpyToSymbolAsync(c_zz, zz, sizeof(zz), 0, cudaMemcpyHostToDevice, ctx().stream));
  THOR_CUDA(cudaStreamSynchronize(ctx().stream));   // the host arrays are stack temporaries
  g_tables_ready = true;
  return 0;
}

struct TbParams {
  const PicView *pics;
  const uint16_t *map;
  const thorcuda_tb_task *tasks;
  const uint32_t *order;
  int16_t *pool;
  uint8_t *cbp;
  uint32_t *ssd;
  int decode;
  // launch layout: size class c (0: 64, 1: 32, 2: 16, 3: 8, 4: 4) owns CTAs [cta_first[c], cta_first[c+1]) and the
  // tasks order[first[c] .. first[c] + count[c])
  int cta_first[6];
  int first[5], count[5];
};

constexpr int TW = 4;                 // warps per CTA
constexpr int TB_ITERS = 4;           // units a lane group processes one after the other (amortises the CTA's table loads)

__device__ __forceinline__ int clip16(int v) { return min(max(v, -32768), 32767); }

// OUT[a][b] = sum_j A[a][j] * D[b][j]   (a < NA, b < NB, j < J; A int8 with row pitch J + 4 bytes, D int16 with
// row pitch dpitch).  A group of G lanes tiles the output 4 x TB per lane; per 4-j step a lane loads 4 basis words
// and 2 TB data words and issues 8 TB DP2A.
template <int G, int NA, int NB, int J, typename Store>
__device__ __forceinline__ void matpass(const int8_t *A, const int16_t *D, int dpitch, int gl, Store store) {
  constexpr int TA = 4;
  constexpr int PER = (NA * NB) / (TA * G);
  constexpr int TBL = PER >= 4 ? 4 : PER >= 2 ? 2 : 1;
  constexpr int NTA = NA / TA, NTB = NB / TBL, T = NTA * NTB;
  constexpr int APW = (J + 4) / 4;
  const int dpw = dpitch >> 1;
#pragma unroll 1
  for (int t = gl; t < T; t += G) {
    const int ta = t % NTA, tb = t / NTA;
    int acc[TA][TBL];
#pragma unroll
    for (int i = 0; i < TA; i++)
#pragma unroll
      for (int b = 0; b < TBL; b++) acc[i][b] = 0;
    const uint32_t *a = reinterpret_cast<const uint32_t *>(A) + ta * TA * APW;
    const uint32_t *d = reinterpret_cast<const uint32_t *>(D) + tb * TBL * dpw;
#pragma unroll
    for (int j = 0; j < J / 4; j++) {
      uint32_t aw[TA], d0[TBL], d1[TBL];
#pragma unroll
      for (int i = 0; i < TA; i++) aw[i] = a[i * APW + j];
#pragma unroll
      for (int b = 0; b < TBL; b++) { d0[b] = d[b * dpw + 2 * j]; d1[b] = d[b * dpw + 2 * j + 1]; }
#pragma unroll
      for (int i = 0; i < TA; i++)
#pragma unroll
        for (int b = 0; b < TBL; b++) {
          acc[i][b] = __dp2a_lo((int)d0[b], (int)aw[i], acc[i][b]);
          acc[i][b] = __dp2a_hi((int)d1[b], (int)aw[i], acc[i][b]);
        }
    }
#pragma unroll
    for (int i = 0; i < TA; i++)
#pragma unroll
      for (int b = 0; b < TBL; b++) store(ta * TA + i, tb * TBL + b, acc[i][b]);
  }
}

// Shared-memory scratch of one unit (int16 counts) for size class N; NI = transform length of the inverse.
template <int N>
struct TbLayout {
  static constexpr int NI = N == 64 ? 32 : N;
  static constexpr int Q = N < 16 ? N : 16;
  static constexpr int PI = NI + 2;                       // pitch of the spatial block
  static constexpr int PQ = Q + 2;                        // pitch of Q-long rows
  static constexpr int in_off = 0;                        // [NI][PI]  residual, later the inverse-transformed block
  static constexpr int mid_sz = (Q * PI > NI * PQ ? Q * PI : NI * PQ);
  static constexpr int mid_off = in_off + NI * PI;        // forward: [Q][PI] (k, r); inverse: [NI][PQ] (r, k)
  static constexpr int coef_off = mid_off + mid_sz;       // [Q][PQ] coefficients (v, k); later dequantised, TRANSPOSED (k, v)
  static constexpr int lin_off = coef_off + Q * PQ;       // [Q*Q] coefficients in scan order
  static constexpr int lev_off = lin_off + Q * Q;         // [Q*Q] levels in scan order
  static constexpr int total = (lev_off + Q * Q + 1) & ~1;
  static constexpr int G = N >= 16 ? 32 : N == 8 ? 8 : 4; // lanes per unit
  static constexpr int UPW = 32 / G;                      // units per warp
};

// forward transform of the n x n block in S.in (n = 4..32) -> coef[v][k], v,k < min(n,16), pitch PQ
template <int G, int n, int PI, int PQ>
__device__ __forceinline__ void fwd_passes(int16_t *in, int16_t *mid, int16_t *coef, const int8_t *sFw, int sh1, int sh2, int gl) {
  constexpr int keep = n < 16 ? n : 16;
  const int8_t *A = sFw + fw_off(n);
  const int r1 = 1 << (sh1 - 1), r2 = 1 << (sh2 - 1);
  // pass 1: mid[k][r] = (sum_c M[k][c] in[r][c] + rnd) >> sh1
  matpass<G, keep, n, n>(A, in, PI, gl, [&](int k, int r, int acc) { mid[k * PI + r] = (int16_t)((acc + r1) >> sh1); });
  __syncwarp();
  // pass 2: coef[v][k] = (sum_r M[v][r] mid[k][r] + rnd) >> sh2
  matpass<G, keep, keep, n>(A, mid, PI, gl, [&](int v, int k, int acc) { coef[v * PQ + k] = (int16_t)((acc + r2) >> sh2); });
  __syncwarp();
}

// inverse transform: cT[k][v] (dequantised coefficients, transposed, pitch PQ) -> in[r][c] (NI x NI, pitch PI)
template <int G, int NI, int Q, int PI, int PQ>
__device__ __forceinline__ void inv_passes(int16_t *in, int16_t *mid, const int16_t *cT, const int8_t *sIv, int gl) {
  const int8_t *A = sIv + iv_off(NI);
  // pass 1: midT[r][k] = clip16((sum_v M[v][r] c[v][k] + 64) >> 7)
  matpass<G, NI, Q, Q>(A, cT, PQ, gl, [&](int r, int k, int acc) { mid[r * PQ + k] = (int16_t)clip16((acc + 64) >> 7); });
  __syncwarp();
  // pass 2: block[r][c] = clip16((sum_k M[k][c] midT[r][k] + 2048) >> 12)
  matpass<G, NI, NI, Q>(A, mid, PQ, gl, [&](int c, int r, int acc) { in[r * PI + c] = (int16_t)clip16((acc + 2048) >> 12); });
  __syncwarp();
}

template <int G>
__device__ __forceinline__ unsigned group_ballot(bool pred, int gbase) {
  const unsigned b = __ballot_sync(0xffffffffu, pred);
  return G == 32 ? b : (b >> gbase) & ((1u << (G & 31)) - 1u);
}

// One transform block of size class N, processed by the G lanes gl = 0..G-1 of a group (all groups of a warp run
// this in lock step; `valid` only gates the global stores).  S: the unit's scratch (TbLayout<N>).
template <int N>
__device__ __forceinline__ void tb_unit(const TbParams &P, uint32_t ti, const thorcuda_tb_task &t, bool valid, int16_t *S, const int8_t *sFw,
                                        const int8_t *sIv, const uint8_t *sZ, int gl, int gbase) {
  using L = TbLayout<N>;
  constexpr int G = L::G, NI = L::NI, Q = L::Q, PI = L::PI, PQ = L::PQ;
  constexpr int LG = N == 4 ? 2 : N == 8 ? 3 : N == 16 ? 4 : N == 32 ? 5 : 6;
  int16_t *in = S + L::in_off, *mid = S + L::mid_off, *coef = S + L::coef_off, *lin = S + L::lin_off, *lev = S + L::lev_off;
  const PicView &vp = P.pics[P.map[t.pred]], &vr = P.pics[P.map[t.rec]];
  const int pl = t.plane;
  const uint8_t *pred = vp.p[pl] + (size_t)t.y * vp.stride[pl] + t.x;
  uint8_t *rec = vr.p[pl] + (size_t)t.y * vr.stride[pl] + t.x;
  const int pstride = vp.stride[pl], rstride = vr.stride[pl];
  const int qp = t.qp, qper = qp / 6, qrem = qp - 6 * qper;
  int16_t *pool = P.pool + t.coeff_offset;
  int any = 0;

  if (!P.decode) {
    const PicView &vo = P.pics[P.map[t.org]];
    const uint8_t *org = vo.p[pl] + (size_t)t.y * vo.stride[pl] + t.x;
    const int ostride = vo.stride[pl];
    // ---- residual, box-summed where the reference does so (transform.c:273-307) ----
    const bool fast = N > 16 && t.fast;
    const int f = fast ? N / 16 : (N == 64 ? 2 : 1);   // box size
    const int n = N / f;                               // forward transform length: 4..32
    const int sh1 = fast ? (N == 64 ? 8 : 6) : (N == 64 ? 7 : LG), sh2 = fast ? 9 : (N == 64 ? 10 : LG + 5);
    if (N <= 32 && f == 1) {
      for (int i = gl; i < N * (N / 4); i += G) {
        const int r = i / (N / 4), c = 4 * (i % (N / 4));
        const uint32_t o = *reinterpret_cast<const uint32_t *>(org + (size_t)r * ostride + c);
        const uint32_t p = *reinterpret_cast<const uint32_t *>(pred + (size_t)r * pstride + c);
        // packed (o - p) per byte pair: bytes 0,1 -> word 0; bytes 2,3 -> word 1
        const uint32_t w0 = ((o & 0xffu) - (p & 0xffu)) & 0xffffu | ((((o >> 8) & 0xffu) - ((p >> 8) & 0xffu)) << 16);
        const uint32_t w1 = (((o >> 16) & 0xffu) - ((p >> 16) & 0xffu)) & 0xffffu | (((o >> 24) - (p >> 24)) << 16);
        uint32_t *dst = reinterpret_cast<uint32_t *>(in + r * PI + c);
        dst[0] = w0; dst[1] = w1;
      }
    } else if (N > 16) {
      // one item = two horizontally adjacent output samples: f x f boxes of residuals (f = 2 or 4; 4-byte aligned rows)
      for (int i = gl; i < n * (n / 2); i += G) {
        const int r = i / (n / 2), c = 2 * (i - r * (n / 2));
        int s0 = 0, s1 = 0;
        for (int a = 0; a < f; a++) {
          const uint8_t *po = org + (size_t)(r * f + a) * ostride + c * f, *pp = pred + (size_t)(r * f + a) * pstride + c * f;
          if (f == 2) {
            const uint32_t o = *reinterpret_cast<const uint32_t *>(po), p = *reinterpret_cast<const uint32_t *>(pp);
            s0 += (int)(o & 255) + (int)((o >> 8) & 255) - (int)(p & 255) - (int)((p >> 8) & 255);
            s1 += (int)((o >> 16) & 255) + (int)(o >> 24) - (int)((p >> 16) & 255) - (int)(p >> 24);
          } else {
            const uint2 o = *reinterpret_cast<const uint2 *>(po), p = *reinterpret_cast<const uint2 *>(pp);
            s0 += (int)__dp4a(o.x, 0x01010101u, 0u) - (int)__dp4a(p.x, 0x01010101u, 0u);
            s1 += (int)__dp4a(o.y, 0x01010101u, 0u) - (int)__dp4a(p.y, 0x01010101u, 0u);
          }
        }
        *reinterpret_cast<uint32_t *>(in + r * PI + c) = ((uint32_t)s0 & 0xffffu) | ((uint32_t)s1 << 16);
      }
    }
    __syncwarp();
    if constexpr (N <= 16) {
      fwd_passes<G, N, PI, PQ>(in, mid, coef, sFw, sh1, sh2, gl);
    } else {
      if (n == 16) fwd_passes<G, 16, PI, PQ>(in, mid, coef, sFw, sh1, sh2, gl);
      else fwd_passes<G, 32, PI, PQ>(in, mid, coef, sFw, sh1, sh2, gl);
    }
    // ---- quantise (enc/encode_block.c:75-178, rdoq == 0) ----
    const uint8_t *scan = sZ + (Q == 4 ? 0 : Q == 8 ? 16 : 80);
    const int intra = (t.coeff_type >> 1) & 1, chroma = t.coeff_type & 1;
    const int scale = c_qscale[qrem], shift = 21 - LG + qper;
    for (int i = gl; i < Q * Q; i += G) { lin[scan[i]] = coef[(i / Q) * PQ + (i % Q)]; lev[i] = 0; }
    __syncwarp();
    const int tail_off = (intra ? 38 : -26) << (shift - 8);
    int last = -1;
    for (int p = gl; p < Q * Q; p += G)
      if ((abs(abs((int)lin[p]) * scale + tail_off) >> shift) != 0) last = p;   // ascending p: keeps the largest
#pragma unroll
    for (int m = G / 2; m > 0; m >>= 1) last = max(last, __shfl_xor_sync(0xffffffffu, last, m));
    const int off_small = (intra ? 102 : 51) << (shift - 8), off_big = (intra ? 115 : 90) << (shift - 8);
    constexpr int NW = (Q * Q + 31) / 32;
    constexpr int NIT = Q * Q / G;
    int lv[NIT];                                       // this lane's levels (magnitudes)
#pragma unroll
    for (int it = 0; it < NIT; it++) {
      const int p = it * G + gl;
      int l = 0;
      if (p <= last) {
        const int c = lin[p], mag = scale * abs(c);
        const int off = ((mag >> shift) == 0 || chroma) ? off_small : off_big;
        l = (mag + off) >> shift;
        lev[p] = (int16_t)(c < 0 ? -l : l);
      }
      any |= (l != 0);
      lv[it] = l;
    }
    any = group_ballot<G>(any != 0, gbase) != 0;
    __syncwarp();
    // "RDOQ light" (:135-168): a level > 1 that follows two zeros is traded for +-1s.  The pass only ever writes
    // +-1 (onto the level itself or onto one of the two zeros before it), so neither "|level| > 1" nor the two
    // zeros of a LATER position can come into being during the pass: the positions that can fire are known up
    // front -- found here in parallel -- and only those are walked in order (the walk re-tests everything on the
    // current values, as the conditions on positions p-3 / p-4 do depend on earlier outcomes).
    unsigned trig_mask[NW];
#pragma unroll
    for (int w = 0; w < NW; w++) trig_mask[w] = 0;
#pragma unroll
    for (int it = 0; it < NIT; it++) {                 // uniform trip count: every lane reaches the ballot
      const int p = it * G + gl;
      const bool trig = lv[it] > 1 && p >= 2 && lev[p - 1] == 0 && lev[p - 2] == 0;
      trig_mask[(it * G) / 32] |= group_ballot<G>(trig, gbase) << ((it * G) & 31);
    }
    if (any && gl == 0) {
      const int NN = chroma ? last + 1 : Q * Q;
      const int thr = (73 * (int)c_dqscale[qrem] << qper) >> (4 + LG);
#pragma unroll
      for (int w = 0; w < NW; w++) {
        unsigned m = trig_mask[w];
        while (m) {
          const int p = w * 32 + __ffs(m) - 1;
          m &= m - 1;
          if (p < 2 || p >= NN) continue;
          bool ok = true;
          if (p > 2 && abs((int)lev[p - 3]) > 1) ok = false;
          if (p > 3 && abs((int)lev[p - 4]) > 1 && abs((int)lev[p - 3]) > 0) ok = false;
          if (p == 2 && (!chroma || last >= 6)) ok = false;
          if (!(ok && lev[p - 2] == 0 && lev[p - 1] == 0 && abs((int)lev[p]) > 1)) continue;
          const int k1 = abs((int)lin[p]), k2 = abs((int)lin[p - 1]), k3 = abs((int)lin[p - 2]);
          if (k1 + max(k2, k3) < thr) lev[p] = lin[p] < 0 ? -1 : 1;
          else if (k2 > k3) lev[p - 1] = lin[p - 1] < 0 ? -1 : 1;
          else lev[p - 2] = lin[p - 2] < 0 ? -1 : 1;
        }
      }
    }
    __syncwarp();
    // levels back to raster order: to the pool, and (dequantised below) transposed into coef
    for (int i = gl; i < Q * Q; i += G) {
      const int16_t l = lev[scan[i]];
      lin[i] = l;                                       // lin now holds the raster-order levels
      if (valid) pool[i] = l;
    }
    if (gl == 0 && valid) P.cbp[ti] = (uint8_t)any;
    __syncwarp();
  } else {
    for (int i = gl; i < Q * Q; i += G) lin[i] = pool[i];
    any = 1;
    __syncwarp();
  }

  uint32_t ssd = 0;
  // `any` is uniform inside a group but may differ between the groups of a warp: the inverse runs for the whole
  // warp when any group needs it (its result is simply not used by the others)
  if (__any_sync(0xffffffffu, any)) {
    // ---- dequantise (common_block.c:132-146), stored transposed: cT[k][v] ----
    const int dq = c_dqscale[qrem], rshift = LG - 1, add = 1 << (rshift - 1);
    for (int i = gl; i < Q * Q; i += G) {
      const int v = i / Q, k = i - v * Q;
      coef[k * PQ + v] = (int16_t)((((int)lin[i] * dq << qper) + add) >> rshift);
    }
    __syncwarp();
    inv_passes<G, NI, Q, PI, PQ>(in, mid, coef, sIv, gl);
  }
  // ---- reconstruct (common_block.c:148-156; 64: 2x nearest-neighbour up-sampling, transform.c:507-514) + SSD ----
  const uint8_t *org = nullptr;
  int ostride = 0;
  if (!P.decode && P.ssd) {
    const PicView &vo = P.pics[P.map[t.org]];
    org = vo.p[pl] + (size_t)t.y * vo.stride[pl] + t.x;
    ostride = vo.stride[pl];
  }
  if (valid && (any || rec != pred || org)) {
    for (int i = gl; i < N * (N / 4); i += G) {
      const int r = i / (N / 4), c = 4 * (i % (N / 4));
      const uint32_t p = *reinterpret_cast<const uint32_t *>(pred + (size_t)r * pstride + c);
      uint32_t out = p;
      if (any) {
        out = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const int res = N == 64 ? in[(r >> 1) * PI + ((c + k) >> 1)] : in[r * PI + c + k];
          out |= (uint32_t)clip255((int)((p >> (8 * k)) & 255) + res) << (8 * k);
        }
      }
      if (any || rec != pred) *reinterpret_cast<uint32_t *>(rec + (size_t)r * rstride + c) = out;
      if (org) {
        const uint32_t o = *reinterpret_cast<const uint32_t *>(org + (size_t)r * ostride + c);
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const int d = (int)((o >> (8 * k)) & 255) - (int)((out >> (8 * k)) & 255);
          ssd += d * d;
        }
      }
    }
  }
  if (org) {     // uniform per launch
#pragma unroll
    for (int m = G / 2; m > 0; m >>= 1) ssd += __shfl_xor_sync(0xffffffffu, ssd, m);
    if (gl == 0 && valid) P.ssd[ti] = ssd;
  }
  __syncwarp();
}

// Pull the pixel rows of a unit (org and pred) towards the SM ahead of their use.
template <int N>
__device__ __forceinline__ void tb_prefetch(const TbParams &P, const thorcuda_tb_task &t, int gl) {
  constexpr int G = TbLayout<N>::G;
  const int pl = t.plane;
  const PicView &vp = P.pics[P.map[t.pred]];
  const uint8_t *pred = vp.p[pl] + (size_t)t.y * vp.stride[pl] + t.x;
  for (int r = gl; r < N; r += G) asm volatile("prefetch.global.L1 [%0];" ::"l"(pred + (size_t)r * vp.stride[pl]));
  if (!P.decode) {
    const PicView &vo = P.pics[P.map[t.org]];
    const uint8_t *org = vo.p[pl] + (size_t)t.y * vo.stride[pl] + t.x;
    for (int r = gl; r < N; r += G) asm volatile("prefetch.global.L1 [%0];" ::"l"(org + (size_t)r * vo.stride[pl]));
  }
}

constexpr int cmax(int a, int b) { return a > b ? a : b; }
// scratch of one warp: the largest of the five classes
constexpr int TB_SCRATCH = cmax(cmax(cmax(TbLayout<64>::total * TbLayout<64>::UPW, TbLayout<32>::total * TbLayout<32>::UPW),
                                     cmax(TbLayout<16>::total * TbLayout<16>::UPW, TbLayout<8>::total * TbLayout<8>::UPW)),
                                TbLayout<4>::total * TbLayout<4>::UPW);

template <int N>
__device__ __forceinline__ void tb_class(const TbParams &P, int cls, int16_t *sScratch, const int8_t *sFw, const int8_t *sIv, const uint8_t *sZ) {
  using L = TbLayout<N>;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gl = lane % L::G, gbase = lane - gl, grp = lane / L::G;
  int16_t *S = sScratch + warp * TB_SCRATCH + grp * L::total;
  const int cta = blockIdx.x - P.cta_first[cls];
  const int count = P.count[cls];
  // unit index: iteration-major inside the CTA so that the groups of a warp work on neighbouring tasks
  const int per_iter = TW * L::UPW;
  const int slot0 = cta * TB_ITERS * per_iter + warp * L::UPW + grp;
  if (cta * TB_ITERS * per_iter >= count) return;
  uint32_t ti = P.order[P.first[cls] + min(slot0, count - 1)];
  thorcuda_tb_task t = P.tasks[ti];
#pragma unroll 1
  for (int it = 0; it < TB_ITERS; it++) {
    const int base = (cta * TB_ITERS + it) * per_iter;
    if (base >= count) break;                              // uniform over the CTA
    const int slot = base + warp * L::UPW + grp;
    // the next unit's task record is fetched, and its pixel rows are pulled towards the SM, while this one is worked on
    uint32_t ti_n = ti;
    thorcuda_tb_task t_n = t;
    if (it + 1 < TB_ITERS && base + per_iter < count) {
      ti_n = P.order[P.first[cls] + min(slot + per_iter, count - 1)];
      t_n = P.tasks[ti_n];
      tb_prefetch<N>(P, t_n, gl);
    }
    tb_unit<N>(P, ti, t, slot < count, S, sFw, sIv, sZ, gl, gbase);
    ti = ti_n;
    t = t_n;
  }
}

__global__ void __launch_bounds__(TW * 32, 6) k_tb(const TbParams P) {
  __shared__ __align__(16) int8_t sFw[FW_BYTES];
  __shared__ __align__(16) int8_t sIv[IV_BYTES];
  __shared__ __align__(16) uint8_t sZ[16 + 64 + 256];
  __shared__ __align__(16) int16_t sScratch[TW * TB_SCRATCH];
  for (int i = threadIdx.x; i < FW_BYTES / 4; i += TW * 32) reinterpret_cast<uint32_t *>(sFw)[i] = reinterpret_cast<const uint32_t *>(c_fw)[i];
  for (int i = threadIdx.x; i < IV_BYTES / 4; i += TW * 32) reinterpret_cast<uint32_t *>(sIv)[i] = reinterpret_cast<const uint32_t *>(c_iv)[i];
  for (int i = threadIdx.x; i < (16 + 64 + 256) / 4; i += TW * 32) reinterpret_cast<uint32_t *>(sZ)[i] = reinterpret_cast<const uint32_t *>(c_zz)[i];
  __syncthreads();
  const int b = blockIdx.x;
  if (b < P.cta_first[1]) tb_class<64>(P, 0, sScratch, sFw, sIv, sZ);
  else if (b < P.cta_first[2]) tb_class<32>(P, 1, sScratch, sFw, sIv, sZ);
  else if (b < P.cta_first[3]) tb_class<16>(P, 2, sScratch, sFw, sIv, sZ);
  else if (b < P.cta_first[4]) tb_class<8>(P, 3, sScratch, sFw, sIv, sZ);
  else tb_class<4>(P, 4, sScratch, sFw, sIv, sZ);
}

// Stand-alone transform of ONE int16 block (the reference's transform()/inverse_transform() signatures,
// common/transform.h:36-37): forward N x N (stride N) -> low Q x Q corner written at stride N; inverse reads
// that corner at stride N -> N x N block.  One warp.  Used by the per-block compatibility entry points.
template <int N>
__global__ void __launch_bounds__(32) k_tb_raw(const int16_t *__restrict__ in, int16_t *__restrict__ out, int fastflag, int inverse) {
  using L = TbLayout<N>;
  constexpr int NI = L::NI, Q = L::Q, PI = L::PI, PQ = L::PQ;
  __shared__ __align__(16) int8_t sFw[FW_BYTES];
  __shared__ __align__(16) int8_t sIv[IV_BYTES];
  __shared__ __align__(16) int16_t S[L::total];
  const int lane = threadIdx.x;
  for (int i = lane; i < FW_BYTES / 4; i += 32) reinterpret_cast<uint32_t *>(sFw)[i] = reinterpret_cast<const uint32_t *>(c_fw)[i];
  for (int i = lane; i < IV_BYTES / 4; i += 32) reinterpret_cast<uint32_t *>(sIv)[i] = reinterpret_cast<const uint32_t *>(c_iv)[i];
  __syncwarp();
  constexpr int LG = N == 4 ? 2 : N == 8 ? 3 : N == 16 ? 4 : N == 32 ? 5 : 6;
  int16_t *sin = S + L::in_off, *mid = S + L::mid_off, *coef = S + L::coef_off;
  if (!inverse) {
    const bool fast = N > 16 && fastflag;
    const int f = fast ? N / 16 : (N == 64 ? 2 : 1);
    const int n = N / f;
    const int sh1 = fast ? (N == 64 ? 8 : 6) : (N == 64 ? 7 : LG), sh2 = fast ? 9 : (N == 64 ? 10 : LG + 5);
    for (int i = lane; i < n * n; i += 32) {
      const int r = i / n, c = i - r * n;
      int s = 0;
      for (int a = 0; a < f; a++)
        for (int b = 0; b < f; b++) s += in[(r * f + a) * N + c * f + b];
      sin[r * PI + c] = (int16_t)s;
    }
    __syncwarp();
    if constexpr (N <= 16) {
      fwd_passes<32, N, PI, PQ>(sin, mid, coef, sFw, sh1, sh2, lane);
    } else {
      if (n == 16) fwd_passes<32, 16, PI, PQ>(sin, mid, coef, sFw, sh1, sh2, lane);
      else fwd_passes<32, 32, PI, PQ>(sin, mid, coef, sFw, sh1, sh2, lane);
    }
    for (int i = lane; i < Q * Q; i += 32) out[(i / Q) * N + (i % Q)] = coef[(i / Q) * PQ + (i % Q)];
  } else {
    for (int i = lane; i < Q * Q; i += 32) coef[(i % Q) * PQ + (i / Q)] = in[(i / Q) * N + (i % Q)];   // transposed: cT[k][v]
    __syncwarp();
    inv_passes<32, NI, Q, PI, PQ>(sin, mid, coef, sIv, lane);
    for (int i = lane; i < N * N; i += 32) {
      const int r = i / N, c = i - r * N;
      out[i] = N == 64 ? sin[(r >> 1) * PI + (c >> 1)] : sin[r * PI + c];
    }
  }
}

int tb_raw(int size, int fast, int inverse, const int16_t *d_in, int16_t *d_out) {
  Context &c = ctx();
  if (upload_tables()) return -1;
  switch (size) {
    case 4: k_tb_raw<4><<<1, 32, 0, c.stream>>>(d_in, d_out, fast, inverse); break;
    case 8: k_tb_raw<8><<<1, 32, 0, c.stream>>>(d_in, d_out, fast, inverse); break;
    case 16: k_tb_raw<16><<<1, 32, 0, c.stream>>>(d_in, d_out, fast, inverse); break;
    case 32: k_tb_raw<32><<<1, 32, 0, c.stream>>>(d_in, d_out, fast, inverse); break;
    case 64: k_tb_raw<64><<<1, 32, 0, c.stream>>>(d_in, d_out, fast, inverse); break;
    default: return fail("transform: size %d", size);
  }
  THOR_LAUNCHED();
  return 0;
}

// ---- early-skip significance test (enc/encode_block.c:2507-2610): one warp per sub-block ----
struct SkipParams {
  PicView pics[8];                     // the batch's pictures (by value: no resident list needed)
  const thorcuda_skip_task *tasks;
  uint8_t *flags;
  int ntasks;
};

template <int n>   // n = size / 2: length of the transform applied to the 2x2-averaged residual
__device__ __forceinline__ bool skip_luma(const PicView &vo, const PicView &vp, const thorcuda_skip_task &t, int16_t *S, const int8_t *sFw, int lane) {
  using L = TbLayout<8>;
  constexpr int PI = L::PI, PQ = L::PQ, LG = n == 4 ? 2 : 3;
  int16_t *in = S + L::in_off, *mid = S + L::mid_off, *coef = S + L::coef_off;
  const uint8_t *org = vo.p[0] + (size_t)t.y * vo.stride[0] + t.x, *pred = vp.p[0] + (size_t)t.y * vp.stride[0] + t.x;
  for (int i = lane; i < n * n; i += 32) {
    const int r = i / n, c = i - r * n;
    int s = 2;
#pragma unroll
    for (int a = 0; a < 2; a++) {
      const uint8_t *po = org + (size_t)(2 * r + a) * vo.stride[0] + 2 * c, *pp = pred + (size_t)(2 * r + a) * vp.stride[0] + 2 * c;
      s += (int)po[0] + (int)po[1] - (int)pp[0] - (int)pp[1];
    }
    in[r * PI + c] = (int16_t)(s >> 2);
  }
  __syncwarp();
  fwd_passes<32, n, PI, PQ>(in, mid, coef, sFw, LG, LG + 5, lane);
  bool sig = false;
  for (int i = lane; i < n * n; i += 32) sig |= abs((int)coef[(i / n) * PQ + (i % n)]) > t.thr_y;
  return __any_sync(0xffffffffu, sig);
}

__device__ __forceinline__ bool skip_chroma(const PicView &vo, const PicView &vp, const thorcuda_skip_task &t, int pl, int lane) {
  const int cs = t.size >> 1, cx = t.x >> 1, cy = t.y >> 1;
  const uint8_t *org = vo.p[pl] + (size_t)cy * vo.stride[pl] + cx, *pred = vp.p[pl] + (size_t)cy * vp.stride[pl] + cx;
  bool sig = false;
  if (lane < cs) {
    int sum = 0;
    for (int i = 0; i < cs; i++) sum += (int)org[(size_t)i * vo.stride[pl] + lane] - (int)pred[(size_t)i * vp.stride[pl] + lane];
    if (cs == 8) {
      sig = (int16_t)sum > (int16_t)t.thr_c;            // v128_cmpgt_s16 against v128_dup_16(threshold)
    } else {
      sum += __shfl_xor_sync(0x0000000fu, sum, 1);     // pairs of columns
      sig = sum > t.thr_c;
    }
  }
  return __any_sync(0xffffffffu, sig);
}

__global__ void __launch_bounds__(TW * 32) k_skip_test(const SkipParams P) {
  __shared__ __align__(16) int8_t sFw[FW_BYTES];
  __shared__ __align__(16) int16_t sScratch[TW * TbLayout<8>::total];
  for (int i = threadIdx.x; i < FW_BYTES / 4; i += TW * 32) reinterpret_cast<uint32_t *>(sFw)[i] = reinterpret_cast<const uint32_t *>(c_fw)[i];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ti = blockIdx.x * TW + warp;
  if (ti >= P.ntasks) return;
  const thorcuda_skip_task t = P.tasks[ti];
  const PicView &vo = P.pics[t.org], &vp = P.pics[t.pred];
  int16_t *S = sScratch + warp * TbLayout<8>::total;
  const bool y = t.size == 16 ? skip_luma<8>(vo, vp, t, S, sFw, lane) : skip_luma<4>(vo, vp, t, S, sFw, lane);
  const bool u = skip_chroma(vo, vp, t, 1, lane), v = skip_chroma(vo, vp, t, 2, lane);
  if (lane == 0) P.flags[ti] = (uint8_t)((y ? 1 : 0) | (u ? 2 : 0) | (v ? 4 : 0));
}

// One launch over a sub-range of every size class of the list: class_first / class_count are indexed like List::ntb
// (4, 8, 16, 32, 64) and address the class-sorted `order` array.  nullptr: the whole list.
int run_tb_classes(List *L, int decode, bool want_ssd, const int *class_first, const int *class_count) {
  Context &c = ctx();
  if (upload_tables()) return -1;
  TbParams P;
  P.pics = c.d_pics; P.map = L->map();
  P.tasks = (const thorcuda_tb_task *)L->tasks();
  P.order = L->order;
  P.pool = L->pool; P.cbp = L->cbp; P.ssd = want_ssd ? L->ssd : nullptr;
  P.decode = decode;
  // list classes are ordered 4,8,16,32,64 (lists.cu); the launch runs them largest first
  const int upw[5] = {TbLayout<64>::UPW, TbLayout<32>::UPW, TbLayout<16>::UPW, TbLayout<8>::UPW, TbLayout<4>::UPW};
  int cta = 0;
  for (int k = 0; k < 5; k++) {
    const int lk = 4 - k;
    const int n = class_count ? class_count[lk] : L->ntb[lk];
    P.first[k] = class_first ? class_first[lk] : L->tb_first[lk]; P.count[k] = n;
    P.cta_first[k] = cta;
    const int per_cta = TW * upw[k] * TB_ITERS;
    cta += (n + per_cta - 1) / per_cta;
  }
  P.cta_first[5] = cta;
  if (!cta) return 0;
  static bool carveout_set = false;
  if (!carveout_set) {   // 6 CTAs x 22 KB per SM: ask for the large shared-memory split
    cudaFuncSetAttribute(k_tb, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    carveout_set = true;
  }
  k_tb<<<(unsigned)cta, TW * 32, 0, c.stream>>>(P);
  THOR_LAUNCHED();
  return 0;
}
static int run_tb(List *L, int decode, bool want_ssd) { return run_tb_classes(L, decode, want_ssd, nullptr, nullptr); }

}  // namespace thor

using namespace thor;

extern "C" {

int thorcuda_encode_list(thorcuda_list hl, int16_t *host_coeffq, uint8_t *host_cbp, uint32_t *host_ssd) {
  THOR_NEED_CTX();
  List *L = get_list(hl);
  if (!L || L->kind != THORCUDA_LIST_TB) return fail("encode: not a transform-block list");
  if (list_stale(L)) return -1;
  if (list_fence(L)) return -1;                 // a previous call's output copies must not be overwritten
  if (run_tb(L, 0, true)) return -1;
  if (host_coeffq && d2h_bulk(host_coeffq, L->pool, L->pool_entries * sizeof(int16_t))) return -1;
  if (host_cbp && d2h_bulk(host_cbp, L->cbp, (size_t)L->ntasks)) return -1;
  if (host_ssd && d2h_bulk(host_ssd, L->ssd, (size_t)L->ntasks * 4)) return -1;
  L->late_pool = host_coeffq || host_cbp || host_ssd;
  return 0;
}

int thorcuda_reconstruct_list(thorcuda_list hl, const int16_t *host_coeffq) {
  THOR_NEED_CTX();
  List *L = get_list(hl);
  if (!L || L->kind != THORCUDA_LIST_TB) return fail("reconstruct: not a transform-block list");
  if (list_stale(L)) return -1;
  if (list_fence(L)) return -1;
  if (host_coeffq && h2d(L->pool, host_coeffq, L->pool_entries * sizeof(int16_t))) return -1;
  return run_tb(L, 1, false);
}

void thorcuda_early_skip_thresholds(int qp_y, int size, float early_skip_thr, int *thr_y, int *thr_c) {
  static const int scale[6] = {26214, 23302, 20560, 18396, 16384, 14564};     // gquant_table, common/common_block.c:97
  const int n = size > 4 ? size / 2 : size;                                    // transform length after the 2x2 mean
  const int lg = n == 4 ? 2 : n == 8 ? 3 : n == 16 ? 4 : 5;
  // check_early_skip_sub_block passes 0.5*early_skip_threshold (double) for size > 4 (enc/encode_block.c:2533)
  const double rel = size > 4 ? 0.5 * early_skip_thr : (double)early_skip_thr;
  const double fql_y = (double)(1 << (21 - lg + qp_y / 6)) / (double)scale[qp_y % 6];
  *thr_y = (int)(rel * fql_y);
  const int qpc = h_chroma_qp[qp_y < 0 ? 0 : qp_y > 51 ? 51 : qp_y];
  const double fql_c = (double)(1 << (21 - 5 + qpc / 6)) / (double)scale[qpc % 6];
  *thr_c = (int)(early_skip_thr * fql_c);
}

int thorcuda_early_skip_batch(const thorcuda_pic *pics, int npics, const thorcuda_skip_task *tasks, int ntasks, uint8_t *host_flags) {
  THOR_NEED_CTX();
  Context &c = ctx();
  if (ntasks <= 0) return 0;
  if (npics <= 0 || npics > 8) return fail("early_skip: npics %d out of range (1..8)", npics);
  if (upload_tables()) return -1;
  SkipParams P;
  memset(&P, 0, sizeof(P));
  for (int i = 0; i < npics; i++) {
    Pic *p = get_pic(pics[i]);
    if (!p) return fail("early_skip: bad picture handle %d", pics[i]);
    P.pics[i] = p->view();
  }
  for (int i = 0; i < ntasks; i++) {
    const thorcuda_skip_task &t = tasks[i];
    if (t.org >= npics || t.pred >= npics) return fail("early_skip: task %d names a picture outside pics[%d]", i, npics);
    const Pic *o = get_pic(pics[t.org]), *p = get_pic(pics[t.pred]);
    if ((t.size != 8 && t.size != 16) || (t.x & 3) || (t.y & 3) || t.x + t.size > o->w || t.y + t.size > o->h || o->w != p->w || o->h != p->h)
      return fail("early_skip: task %d has an invalid sub-block (%d,%d size %d)", i, t.x, t.y, t.size);
  }
  const size_t tbytes = (size_t)ntasks * sizeof(thorcuda_skip_task);
  uint8_t *d = (uint8_t *)c.scratch.reserve(tbytes + ntasks + 64);
  if (!d) return fail("early_skip: device allocation failed");
  if (h2d(d, tasks, tbytes)) return -1;
  P.tasks = (const thorcuda_skip_task *)d;
  P.flags = d + tbytes;
  P.ntasks = ntasks;
  k_skip_test<<<(ntasks + TW - 1) / TW, TW * 32, 0, c.stream>>>(P);
  THOR_LAUNCHED();
  if (host_flags) return d2h(host_flags, P.flags, ntasks);
  return 0;
}

int thorcuda_encode_blocks(const thorcuda_pic *pics, int npics, const thorcuda_tb_task *tasks, int ntasks,
                           int16_t *host_coeffq, uint8_t *host_cbp, uint32_t *host_ssd) {
  if (ntasks <= 0) return 0;
  const thorcuda_list l = thorcuda_list_create(THORCUDA_LIST_TB, pics, npics, tasks, ntasks);
  if (l < 0) return -1;
  const int rc = thorcuda_encode_list(l, host_coeffq, host_cbp, host_ssd);
  thorcuda_list_destroy(l);
  return rc;
}

int thorcuda_reconstruct_blocks(const thorcuda_pic *pics, int npics, const thorcuda_tb_task *tasks, int ntasks,
                                const int16_t *host_coeffq) {
  if (ntasks <= 0) return 0;
  const thorcuda_list l = thorcuda_list_create(THORCUDA_LIST_TB, pics, npics, tasks, ntasks);
  if (l < 0) return -1;
  const int rc = thorcuda_reconstruct_list(l, host_coeffq);
  thorcuda_list_destroy(l);
  return rc;
}

}  // extern "C"
