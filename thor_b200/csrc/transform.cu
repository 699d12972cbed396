// libthorcuda: batched residual -> integer DCT -> quantise -> dequantise -> inverse DCT -> reconstruct
// (encoder leg) and dequantise -> inverse DCT -> reconstruct (decoder leg) for whole transform-block lists.
//
// Restates (bit-exactly) reference common/transform.c:249-330 (transform, incl. the `fast` box-sum path and
// the 64->32 down-sampling), :432-518 (inverse_transform, incl. 2x up-sampling for 64), enc/encode_block.c:75-178
// (quantize with rdoq == 0), :484-493 (get_residual), :782-797 (ssd_calc), common/common_block.c:132-156
// (dequantize, reconstruct_block), chained as encode_and_reconstruct_block_inter (enc/encode_block.c:1469-1532)
// / decode_and_reconstruct_block_inter (dec/decode_block.c:90-120).
//
// Work decomposition: tasks are sorted by transform size at list creation; one launch per size class, one
// warp per transform block, everything between the pixel loads and the pixel stores stays in shared memory.
// The transforms are integer matrix products against the 32-point HEVC basis (smaller bases are its even
// rows), 32-bit accumulate, round, shift and store to int16 after each pass exactly like the reference; the
// reference's butterflies are an exact factorisation of the same products.
#include <string.h>
#include "internal.cuh"

namespace thor {

__constant__ __align__(16) int16_t c_m32[32 * 32];      // 32-point basis, row k = frequency
__constant__ __align__(16) uint8_t c_zz[16 + 64 + 256]; // zig-zag scan position of coefficient (i,j), for q = 4, 8, 16
static bool g_tables_ready = false;

// common/common_block.c:97-98
__constant__ uint16_t c_qscale[6] = {26214, 23302, 20560, 18396, 16384, 14564};
__constant__ uint16_t c_dqscale[6] = {40, 45, 51, 57, 64, 72};

static int upload_tables() {
  if (g_tables_ready) return 0;
  // cos(pi*a/64) ladder of the HEVC integer DCT (reference common/transform.c:41-245 lists the expanded matrices)
  static const int8_t cos32[33] = {64, 90, 90, 90, 89, 88, 87, 85, 83, 82, 80, 78, 75, 73, 70, 67, 64, 61, 57, 54, 50, 46, 43, 38, 36, 31, 25, 22, 18, 13, 9, 4, 0};
  int16_t m[32 * 32];
  for (int k = 0; k < 32; k++)
    for (int n = 0; n < 32; n++) {
      int a = ((2 * n + 1) * k) & 127;
      if (a > 64) a = 128 - a;
      m[k * 32 + n] = (int16_t)(a > 32 ? -cos32[64 - a] : cos32[a]);
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
  THOR_CUDA(cudaMemcpyToSymbolAsync(c_m32, m, sizeof(m), 0, cudaMemcpyHostToDevice, ctx().stream));
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
  int first, count;
  int16_t *pool;
  uint8_t *cbp;
  uint32_t *ssd;
  int decode;
};

constexpr int TW = 4;                 // warps (= transform blocks) per CTA
constexpr int IN_PITCH = 34;          // int16 per row of the spatial block (conflict-free column access)

struct WarpScratch {
  int16_t in[32 * IN_PITCH];          // residual (possibly box-summed), later the inverse-transformed block
  int16_t mid[16 * IN_PITCH];         // first-pass output, [frequency k][row r]
  int16_t coef[256];                  // q x q coefficients, later dequantised coefficients
  int16_t lin[256];                   // coefficients in scan order
  int16_t lev[256];                   // levels in scan order
};

__device__ __forceinline__ int clip16(int v) { return min(max(v, -32768), 32767); }
__device__ __forceinline__ int warp_or(int v) { return __any_sync(0xffffffffu, v); }

// forward passes on the n x n block in S.in (n = 4..32; row k of the n-point basis = row k << msk of the
// 32-point one): S.coef[v*Q + k], v,k < min(n,16)
template <int Q>
__device__ __forceinline__ void fwd_passes(WarpScratch &S, const int16_t *sM, int n, int msk, int sh1, int sh2, int lane) {
  // ---- forward pass 1: mid[k][r] = (sum_c M[k][c] in[r][c] + rnd) >> sh1, k < keep ----
  const int keep = n < 16 ? n : 16;
  for (int i = lane; i < keep * n; i += 32) {
    const int k = i / n, r = i - k * n;
    const int16_t *m = sM + ((k << msk) << 5), *x = S.in + r * IN_PITCH;
    int acc = 0;
    for (int c = 0; c < n; c++) acc += m[c] * x[c];
    S.mid[k * IN_PITCH + r] = (int16_t)((acc + (1 << (sh1 - 1))) >> sh1);
  }
  __syncwarp();
  // ---- forward pass 2: coef[v][k] = (sum_r M[v][r] mid[k][r] + rnd) >> sh2 ----
  for (int i = lane; i < keep * keep; i += 32) {
    const int v = i / keep, k = i - v * keep;
    const int16_t *m = sM + ((v << msk) << 5), *x = S.mid + k * IN_PITCH;
    int acc = 0;
    for (int r = 0; r < n; r++) acc += m[r] * x[r];
    S.coef[v * Q + k] = (int16_t)((acc + (1 << (sh2 - 1))) >> sh2);
  }
  __syncwarp();
}

// inverse passes: S.coef (Q x Q, low-frequency corner) -> S.in (NI x NI)
template <int NI, int Q>
__device__ __forceinline__ void inv_passes(WarpScratch &S, const int16_t *sM, int lane) {
  constexpr int MSK = NI == 4 ? 3 : NI == 8 ? 2 : NI == 16 ? 1 : 0;
  // ---- inverse pass 1: mid[k][r] = clip16((sum_v M[v][r] c[v][k] + 64) >> 7) ----
  for (int i = lane; i < Q * NI; i += 32) {
    const int k = i / NI, r = i - k * NI;
    int acc = 0;
#pragma unroll 4
    for (int v = 0; v < Q; v++) acc += sM[((v << MSK) << 5) + r] * S.coef[v * Q + k];
    S.mid[k * IN_PITCH + r] = (int16_t)clip16((acc + 64) >> 7);
  }
  __syncwarp();
  // ---- inverse pass 2: block[r][c] = clip16((sum_k M[k][c] mid[k][r] + 2048) >> 12) ----
  for (int i = lane; i < NI * NI; i += 32) {
    const int r = i / NI, c = i - r * NI;
    int acc = 0;
#pragma unroll 4
    for (int k = 0; k < Q; k++) acc += sM[((k << MSK) << 5) + c] * S.mid[k * IN_PITCH + r];
    S.in[r * IN_PITCH + c] = (int16_t)clip16((acc + 2048) >> 12);
  }
  __syncwarp();
}

template <int N>
__global__ void __launch_bounds__(TW * 32) k_tb(const TbParams P) {
  __shared__ __align__(16) int16_t sM[32 * 32];
  __shared__ __align__(16) WarpScratch sW[TW];
  __shared__ __align__(16) uint8_t sZ[16 + 64 + 256];
  for (int i = threadIdx.x; i < 32 * 32 / 2; i += TW * 32) reinterpret_cast<uint32_t *>(sM)[i] = reinterpret_cast<const uint32_t *>(c_m32)[i];
  for (int i = threadIdx.x; i < (16 + 64 + 256) / 4; i += TW * 32) reinterpret_cast<uint32_t *>(sZ)[i] = reinterpret_cast<const uint32_t *>(c_zz)[i];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slot = blockIdx.x * TW + warp;
  if (slot >= P.count) return;
  const uint32_t ti = P.order[P.first + slot];
  const thorcuda_tb_task t = P.tasks[ti];
  WarpScratch &S = sW[warp];
  constexpr int LG = N == 4 ? 2 : N == 8 ? 3 : N == 16 ? 4 : N == 32 ? 5 : 6;
  constexpr int Q = N < 16 ? N : 16;                 // coefficients kept per dimension
  constexpr int NI = N == 64 ? 32 : N;               // inverse transform length
  const PicView vp = P.pics[P.map[t.pred]], vr = P.pics[P.map[t.rec]];
  const int pl = t.plane;
  const uint8_t *pred = vp.p[pl] + (size_t)t.y * vp.stride[pl] + t.x;
  uint8_t *rec = vr.p[pl] + (size_t)t.y * vr.stride[pl] + t.x;
  const int pstride = vp.stride[pl], rstride = vr.stride[pl];
  const int qp = t.qp, qper = qp / 6, qrem = qp - 6 * qper;
  int16_t *pool = P.pool + t.coeff_offset;
  int any = 0;

  if (!P.decode) {
    const PicView vo = P.pics[P.map[t.org]];
    const uint8_t *org = vo.p[pl] + (size_t)t.y * vo.stride[pl] + t.x;
    const int ostride = vo.stride[pl];
    // ---- residual, box-summed where the reference does so (transform.c:273-307) ----
    const bool fast = N > 16 && t.fast;
    const int f = fast ? N / 16 : (N == 64 ? 2 : 1);   // box size
    const int n = N / f;                               // forward transform length: 4..32
    const int sh1 = fast ? (N == 64 ? 8 : 6) : (N == 64 ? 7 : LG), sh2 = fast ? 9 : (N == 64 ? 10 : LG + 5);
    const int msk = 5 - (31 - __clz(n));               // row k of the n-point basis = row k << msk of the 32-point one
    if (f == 1) {
      for (int i = lane; i < N * (N / 4); i += 32) {
        const int r = i / (N / 4), c = 4 * (i % (N / 4));
        const uint32_t o = *reinterpret_cast<const uint32_t *>(org + (size_t)r * ostride + c);
        const uint32_t p = *reinterpret_cast<const uint32_t *>(pred + (size_t)r * pstride + c);
#pragma unroll
        for (int k = 0; k < 4; k++) S.in[r * IN_PITCH + c + k] = (int16_t)((int)((o >> (8 * k)) & 255) - (int)((p >> (8 * k)) & 255));
      }
    } else {
      // one item = one output sample: f x f box of residuals (f = 2 or 4; rows are 4-byte aligned)
      for (int i = lane; i < n * n; i += 32) {
        const int r = i / n, c = i - r * n;
        int s = 0;
        for (int a = 0; a < f; a++) {
          const uint8_t *po = org + (size_t)(r * f + a) * ostride + c * f, *pp = pred + (size_t)(r * f + a) * pstride + c * f;
          for (int b = 0; b < f; b++) s += (int)po[b] - (int)pp[b];
        }
        S.in[r * IN_PITCH + c] = (int16_t)s;
      }
    }
    __syncwarp();
    fwd_passes<Q>(S, sM, n, msk, sh1, sh2, lane);
    // ---- quantise (enc/encode_block.c:75-178, rdoq == 0) ----
    const uint8_t *scan = sZ + (Q == 4 ? 0 : Q == 8 ? 16 : 80);
    const int intra = (t.coeff_type >> 1) & 1, chroma = t.coeff_type & 1;
    const int scale = c_qscale[qrem], shift = 21 - LG + qper;
    for (int i = lane; i < Q * Q; i += 32) { S.lin[scan[i]] = S.coef[i]; S.lev[i] = 0; }
    __syncwarp();
    const int tail_off = (intra ? 38 : -26) << (shift - 8);
    int last = -1;
    for (int p = lane; p < Q * Q; p += 32)
      if ((abs(abs((int)S.lin[p]) * scale + tail_off) >> shift) != 0) last = p;   // ascending p: keeps the largest
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) last = max(last, __shfl_xor_sync(0xffffffffu, last, m));
    const int off_small = (intra ? 102 : 51) << (shift - 8), off_big = (intra ? 115 : 90) << (shift - 8);
    unsigned big_mask[8] = {0, 0, 0, 0, 0, 0, 0, 0};    // positions with |level| > 1 (bit = lane, word = p / 32)
#pragma unroll
    for (int wdi = 0; wdi < (Q * Q + 31) / 32; wdi++) {   // uniform trip count: every lane reaches the ballot
      const int p = wdi * 32 + lane;
      int l = 0;
      if (p < Q * Q && p <= last) {
        const int c = S.lin[p], mag = scale * abs(c);
        const int off = ((mag >> shift) == 0 || chroma) ? off_small : off_big;
        l = (mag + off) >> shift;
        S.lev[p] = (int16_t)(c < 0 ? -l : l);
      }
      any |= (l != 0);
      big_mask[wdi] = __ballot_sync(0xffffffffu, l > 1);
    }
    any = warp_or(any);
    __syncwarp();
    if (any && lane == 0) {
      // "RDOQ light" (:135-168): an isolated level > 1 after two zeros is traded for +-1s.  Only positions
      // that hold |level| > 1 can trigger, and the pass never creates new ones, so walk that bit set.
      const int NN = chroma ? last + 1 : Q * Q;
      const int lg = LG;
      const int thr = (73 * (int)c_dqscale[qrem] << qper) >> (4 + lg);
      for (int wdi = 0; wdi < (Q * Q + 31) / 32; wdi++) {
        unsigned m = big_mask[wdi];
        while (m) {
          const int p = wdi * 32 + __ffs(m) - 1;
          m &= m - 1;
          if (p < 2 || p >= NN) continue;
          bool ok = true;
          if (p > 2 && abs((int)S.lev[p - 3]) > 1) ok = false;
          if (p > 3 && abs((int)S.lev[p - 4]) > 1 && abs((int)S.lev[p - 3]) > 0) ok = false;
          if (p == 2 && (!chroma || last >= 6)) ok = false;
          if (!(ok && S.lev[p - 2] == 0 && S.lev[p - 1] == 0 && abs((int)S.lev[p]) > 1)) continue;
          const int k1 = abs((int)S.lin[p]), k2 = abs((int)S.lin[p - 1]), k3 = abs((int)S.lin[p - 2]);
          if (k1 + max(k2, k3) < thr) S.lev[p] = S.lin[p] < 0 ? -1 : 1;
          else if (k2 > k3) S.lev[p - 1] = S.lin[p - 1] < 0 ? -1 : 1;
          else S.lev[p - 2] = S.lin[p - 2] < 0 ? -1 : 1;
        }
      }
    }
    __syncwarp();
    for (int i = lane; i < Q * Q; i += 32) {
      const int16_t l = S.lev[scan[i]];
      S.coef[i] = l;
      pool[i] = l;
    }
    if (lane == 0) P.cbp[ti] = (uint8_t)any;
    __syncwarp();
  } else {
    for (int i = lane; i < Q * Q; i += 32) S.coef[i] = pool[i];
    any = 1;
    __syncwarp();
  }

  uint32_t ssd = 0;
  if (any) {
    // ---- dequantise (common_block.c:132-146) ----
    const int dq = c_dqscale[qrem], rshift = LG - 1, add = 1 << (rshift - 1);
    for (int i = lane; i < Q * Q; i += 32) S.coef[i] = (int16_t)((((int)S.coef[i] * dq << qper) + add) >> rshift);
    __syncwarp();
    inv_passes<NI, Q>(S, sM, lane);
  }
  // ---- reconstruct (common_block.c:148-156; 64: 2x nearest-neighbour up-sampling, transform.c:507-514) + SSD ----
  const uint8_t *org = nullptr;
  int ostride = 0;
  if (!P.decode && P.ssd) {
    const PicView vo = P.pics[P.map[t.org]];
    org = vo.p[pl] + (size_t)t.y * vo.stride[pl] + t.x;
    ostride = vo.stride[pl];
  }
  if (any || rec != pred || org) {
    for (int i = lane; i < N * (N / 4); i += 32) {
      const int r = i / (N / 4), c = 4 * (i % (N / 4));
      const uint32_t p = *reinterpret_cast<const uint32_t *>(pred + (size_t)r * pstride + c);
      uint32_t out = p;
      if (any) {
        out = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const int res = N == 64 ? S.in[(r >> 1) * IN_PITCH + ((c + k) >> 1)] : S.in[r * IN_PITCH + c + k];
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
  if (org) {
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) ssd += __shfl_xor_sync(0xffffffffu, ssd, m);
    if (lane == 0) P.ssd[ti] = ssd;
  }
}

// Stand-alone transform of ONE int16 block (the reference's transform()/inverse_transform() signatures,
// common/transform.h:36-37): forward N x N (stride N) -> low Q x Q corner written at stride N; inverse reads
// that corner at stride N -> N x N block.  One warp.  Used by the per-block compatibility entry points.
template <int N>
__global__ void __launch_bounds__(32) k_tb_raw(const int16_t *__restrict__ in, int16_t *__restrict__ out, int fastflag, int inverse) {
  __shared__ __align__(16) int16_t sM[32 * 32];
  __shared__ __align__(16) WarpScratch S;
  const int lane = threadIdx.x;
  for (int i = lane; i < 32 * 32 / 2; i += 32) reinterpret_cast<uint32_t *>(sM)[i] = reinterpret_cast<const uint32_t *>(c_m32)[i];
  __syncwarp();
  constexpr int LG = N == 4 ? 2 : N == 8 ? 3 : N == 16 ? 4 : N == 32 ? 5 : 6;
  constexpr int Q = N < 16 ? N : 16;
  constexpr int NI = N == 64 ? 32 : N;
  if (!inverse) {
    const bool fast = N > 16 && fastflag;
    const int f = fast ? N / 16 : (N == 64 ? 2 : 1);
    const int n = N / f;
    const int sh1 = fast ? (N == 64 ? 8 : 6) : (N == 64 ? 7 : LG), sh2 = fast ? 9 : (N == 64 ? 10 : LG + 5);
    const int msk = 5 - (31 - __clz(n));
    for (int i = lane; i < n * n; i += 32) {
      const int r = i / n, c = i - r * n;
      int s = 0;
      for (int a = 0; a < f; a++)
        for (int b = 0; b < f; b++) s += in[(r * f + a) * N + c * f + b];
      S.in[r * IN_PITCH + c] = (int16_t)s;
    }
    __syncwarp();
    fwd_passes<Q>(S, sM, n, msk, sh1, sh2, lane);
    for (int i = lane; i < Q * Q; i += 32) out[(i / Q) * N + (i % Q)] = S.coef[i];
  } else {
    for (int i = lane; i < Q * Q; i += 32) S.coef[i] = in[(i / Q) * N + (i % Q)];
    __syncwarp();
    inv_passes<NI, Q>(S, sM, lane);
    for (int i = lane; i < N * N; i += 32) {
      const int r = i / N, c = i - r * N;
      out[i] = N == 64 ? S.in[(r >> 1) * IN_PITCH + (c >> 1)] : S.in[r * IN_PITCH + c];
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

static int run_tb(List *L, int decode, bool want_ssd) {
  Context &c = ctx();
  if (upload_tables()) return -1;
  TbParams P;
  P.pics = c.d_pics; P.map = L->map();
  P.tasks = (const thorcuda_tb_task *)L->tasks();
  P.order = L->order;
  P.pool = L->pool; P.cbp = L->cbp; P.ssd = want_ssd ? L->ssd : nullptr;
  P.decode = decode;
  for (int k = 0; k < 5; k++) {
    if (!L->ntb[k]) continue;
    P.first = L->tb_first[k]; P.count = L->ntb[k];
    const unsigned grid = (unsigned)((P.count + TW - 1) / TW);
    switch (k) {
      case 0: k_tb<4><<<grid, TW * 32, 0, c.stream>>>(P); break;
      case 1: k_tb<8><<<grid, TW * 32, 0, c.stream>>>(P); break;
      case 2: k_tb<16><<<grid, TW * 32, 0, c.stream>>>(P); break;
      case 3: k_tb<32><<<grid, TW * 32, 0, c.stream>>>(P); break;
      default: k_tb<64><<<grid, TW * 32, 0, c.stream>>>(P); break;
    }
    THOR_LAUNCHED();
  }
  return 0;
}

}  // namespace thor

using namespace thor;

extern "C" {

int thorcuda_encode_list(thorcuda_list hl, int16_t *host_coeffq, uint8_t *host_cbp, uint32_t *host_ssd) {
  THOR_NEED_CTX();
  List *L = get_list(hl);
  if (!L || L->kind != THORCUDA_LIST_TB) return fail("encode: not a transform-block list");
  if (run_tb(L, 0, true)) return -1;
  if (host_coeffq && d2h(host_coeffq, L->pool, L->pool_entries * sizeof(int16_t))) return -1;
  if (host_cbp && d2h(host_cbp, L->cbp, (size_t)L->ntasks)) return -1;
  if (host_ssd && d2h(host_ssd, L->ssd, (size_t)L->ntasks * 4)) return -1;
  return 0;
}

int thorcuda_reconstruct_list(thorcuda_list hl, const int16_t *host_coeffq) {
  THOR_NEED_CTX();
  List *L = get_list(hl);
  if (!L || L->kind != THORCUDA_LIST_TB) return fail("reconstruct: not a transform-block list");
  if (host_coeffq && h2d(L->pool, host_coeffq, L->pool_entries * sizeof(int16_t))) return -1;
  return run_tb(L, 1, false);
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
