// libthorcuda: bit-cost model on the device, first slice -- the exact number of bits the reference's
// entropy coder spends on the quantised coefficients of a transform block.
//
// Restates reference enc/write_bits.c:110-253 (write_coeff), :71-108 (find_code) and the code lengths of
// enc/putvlc.c:133-229 (quote_vlc, tables 0, 1, 2 and 10 -- the ones write_coeff uses) as a pure length computation
// over the coefficient pool of a transform-block list, i.e. on the levels thorcuda_encode_list has just produced: the
// encoder's cost_calc (enc/encode_block.c:1218-1228) needs `nbits` of every trial encode, and the coefficients are
// the part of it that scales with the pixels (mode, vector and cbp bits are a few table look-ups per block).
//
// One warp per transform block.  The low q x q corner (q = min(size, 16)) is read with one vector load per lane (a block
// without levels ends there) and scattered into scan order in shared memory.  write_coeff's two-mode scan looks
// sequential, but the mode, the run start and the level table of a position are functions of bit masks of the
// EARLIER positions (level present, level > 1, level > 3), so after three ballots per 32 positions every lane prices
// the symbol at its own position and the warp adds up: the cost is the same for sparse and dense blocks.
#include <string.h>
#include "internal.cuh"

namespace thor {

__constant__ __align__(16) uint8_t c_scan[16 + 64 + 256];      // zig-zag scan position of coefficient (i, j) for q = 4, 8, 16

__device__ __forceinline__ int ilog2u(uint32_t v) { return 31 - __clz(v); }
// quote_vlc for tables 0..5 (enc/putvlc.c:141-157) and table 10 (:210-213)
__device__ __forceinline__ int vlc_len(int n, uint32_t cn) {
  const uint32_t lim = 6u << n;
  return cn < lim ? 1 + n + (int)(cn >> n) : 7 - n + 2 * ilog2u(cn - lim + (1u << n));
}
__device__ __forceinline__ int expgolomb_len(uint32_t cn) { return 1 + 2 * ilog2u(cn + 1); }

// find_code (enc/write_bits.c:71-108) for a run-mode event
__device__ __forceinline__ int find_code(int run, int level, int maxrun, bool chroma) {
  const int m = max(maxrun, 4);
  const int index = run + (level > 1 ? m + 1 : 0);
  if (!chroma && index < 2) return index;
  if (index <= 4) return index + 1;
  if (index <= m) return index + 3;
  if (index == m + 1) return 6;
  if (index == m + 2) return 7;
  return index + 1;
}
// bits of a run-mode code number / of the end-of-block code (:196-206, :238-249)
__device__ __forceinline__ int event_len(int cn, bool small_chroma, bool eob) {
  if (small_chroma) return eob ? vlc_len(0, (uint32_t)cn) : expgolomb_len((uint32_t)cn);
  return cn == 0 ? 2 : vlc_len(2, (uint32_t)cn + 1);
}

constexpr int BW = 8;      // warps (= blocks in flight) per CTA

__global__ void __launch_bounds__(BW * 32) k_coeff_bits(const thorcuda_tb_task *__restrict__ tasks, int ntasks,
                                                        const int16_t *__restrict__ pool, uint32_t *__restrict__ bits) {
  __shared__ int16_t S[BW][256];          // the block's levels in scan order
  __shared__ uint8_t sScan[16 + 64 + 256];   // c_scan is indexed per lane: from constant memory that access would serialise
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < (16 + 64 + 256) / 4; i += BW * 32) reinterpret_cast<uint32_t *>(sScan)[i] = reinterpret_cast<const uint32_t *>(c_scan)[i];
  __syncthreads();
  int16_t *s = S[warp];
  for (int t = blockIdx.x * BW + warp; t < ntasks; t += gridDim.x * BW) {
    const thorcuda_tb_task task = tasks[t];
    const int size = task.size, q = min(size, 16), n = q * q;
    const bool chroma = task.coeff_type & 1, intra = (task.coeff_type >> 1) & 1;
    const uint8_t *scan = sScan + (q == 4 ? 0 : q == 8 ? 16 : 80);
    const int16_t *c = pool + task.coeff_offset;
    // the corner as one vector load per lane (8 levels; 4 for a 4x4 block, whose pool slot is only 8-byte aligned);
    // a block without levels -- nearly half of an encoder's blocks -- ends here
    const int per = q == 4 ? 4 : 8;
    uint32_t w[4] = {0, 0, 0, 0};
    if (lane * per < n) {
      if (q == 4) { const uint2 v = *reinterpret_cast<const uint2 *>(c + 4 * lane); w[0] = v.x; w[1] = v.y; }
      else { const uint4 v = *reinterpret_cast<const uint4 *>(c + 8 * lane); w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w; }
    }
    if (!__any_sync(0xffffffffu, (w[0] | w[1] | w[2] | w[3]) != 0)) {
      if (lane == 0) bits[t] = 0;
      continue;
    }
    if (lane * per < n) {
#pragma unroll
      for (int j = 0; j < 8; j++)
        if (j < per) s[scan[lane * per + j]] = (int16_t)(w[j >> 1] >> (16 * (j & 1)));
    }
    __syncwarp();
    // ---- pass 1: per 32-position word, which positions hold a level, a level > 1, a level > 3 ----
    const int nw = n >= 32 ? n >> 5 : 1;
    int lv[8];                               // |level| of position 32k + lane (bit 16: negative)
    uint32_t nz[8], big[8], gt3[8];
    int last = -1;
#pragma unroll
    for (int k = 0; k < 8; k++) {
      nz[k] = big[k] = gt3[k] = 0;
      lv[k] = 0;
      if (k < nw) {
        const int p = 32 * k + lane, v = p < n ? (int)s[p] : 0, a = abs(v);
        lv[k] = a | (v < 0 ? 0x10000 : 0);
        nz[k] = __ballot_sync(0xffffffffu, a != 0);
        big[k] = __ballot_sync(0xffffffffu, a > 1);
        gt3[k] = __ballot_sync(0xffffffffu, a > 3);
        if (nz[k]) last = 32 * k + ilog2u(nz[k]);
      }
    }
    // ---- pass 2: every position prices its own symbol ----
    // write_coeff's mode before position p follows from the nearest earlier position that is not +-1: level mode after a
    // level > 1 (or at the start), run mode after a zero (:176-178, :216).  A run starts behind the nearest earlier CODED
    // position (a level, or a zero met in level mode).  The level table is adaptive only for intra luma: > 3 of the
    // previous symbol if that was coded in level mode, else table 0 (the zero that opened the run reset it).
    const bool small_chroma = chroma && size <= 8;
    const uint32_t below = (1u << lane) - 1;
    bool carry_level = true;                 // mode at the first position of the word
    int carry_start = 0;                     // run start at the first position of the word
    uint32_t prev_level_msb = 0, prev_gt3_msb = 0;   // bit 31 of the previous word's level-mode / > 3 masks
    int acc = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
      if (k < nw && 32 * k <= last) {
        const uint32_t dec = ~nz[k] | big[k];                       // positions that decide the mode of what follows
        const uint32_t md = dec & below;
        const bool level_mode = md ? (big[k] >> ilog2u(md)) & 1 : carry_level;
        const uint32_t lm = __ballot_sync(0xffffffffu, level_mode);
        const uint32_t coded = nz[k] | (~nz[k] & lm);
        const uint32_t mc = coded & below;
        const int start = mc ? 32 * k + ilog2u(mc) + 1 : carry_start;
        const int p = 32 * k + lane, a = lv[k] & 0xffff, neg = lv[k] >> 16;
        if (p <= last) {
          if (level_mode) {                                         // :166-178
            int table = 0;
            if (!chroma) {
              if (p == 0) table = intra;
              else {
                const uint32_t pl = lane ? (lm >> (lane - 1)) & 1 : prev_level_msb, pg = lane ? (gt3[k] >> (lane - 1)) & 1 : prev_gt3_msb;
                table = pl & pg;
              }
            }
            acc += vlc_len(table, (uint32_t)a) + (a > 0);
          } else if (a) {                                           // :191-218
            acc += event_len(find_code(p - start, a, n - start - 1, chroma), small_chroma, false);
            acc += a > 1 ? vlc_len(0, (uint32_t)(2 * (a - 2) + neg)) : 1;
          }
          if (p == last) {                                          // what follows the last level (:222-250)
            int pos = p + 1;
            if (pos < n && (a > 1 || level_mode)) {                 // still in level mode: one zero level first
              acc += 1 + ((!chroma && level_mode && a > 3) ? 1 : 0);
              pos++;
            }
            if (pos < n) acc += small_chroma ? 1 : chroma ? 2 : 3;  // end of block: table 0 cn 0 / 2 bits / table 2 cn 3
          }
        }
        carry_level = dec ? (big[k] >> ilog2u(dec)) & 1 : carry_level;
        carry_start = coded ? 32 * k + ilog2u(coded) + 1 : carry_start;
        prev_level_msb = lm >> 31;
        prev_gt3_msb = gt3[k] >> 31;
      }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    int total = acc + (chroma ? 1 : 0);                              // :147-158: the chroma flag bit
    if (chroma && last == 0 && (nz[0] & 1) && !(big[0] & 1)) total = 2;   // a lone +-1: flag + sign, nothing else
    if (lane == 0) bits[t] = (uint32_t)total;
    __syncwarp();
  }
}

static bool scan_ready = false;
static int upload_scan() {
  if (scan_ready) return 0;
  uint8_t zz[16 + 64 + 256];
  int off = 0;
  for (int q = 4; q <= 16; q *= 2) {              // classic zig-zag (reference common/common_block.c:38-73), generated
    int idx = 0;
    for (int d = 0; d <= 2 * q - 2; d++)
      for (int m = 0; m <= d; m++) {
        const int i = (d & 1) ? m : d - m, j = d - i;
        if (i < q && j < q) zz[off + i * q + j] = (uint8_t)idx++;
      }
    off += q * q;
  }
  THOR_CUDA(cudaMemcpyToSymbolAsync(c_scan, zz, sizeof(zz), 0, cudaMemcpyHostToDevice, ctx().stream));
  THOR_CUDA(cudaStreamSynchronize(ctx().stream));            // zz lives on this stack frame
  scan_ready = true;
  return 0;
}
void bits_reset() { scan_ready = false; }

}  // namespace thor

using namespace thor;

extern "C" {

int thorcuda_coeff_bits_list(thorcuda_list hl, const int16_t *host_coeffq, uint32_t *host_bits) {
  THOR_NEED_CTX();
  Context &c = ctx();
  List *L = get_list(hl);
  if (!L || L->kind != THORCUDA_LIST_TB) return fail("coeff_bits: not a transform-block list");
  if (list_fence(L)) return -1;
  if (upload_scan()) return -1;
  if (host_coeffq && h2d(L->pool, host_coeffq, L->pool_entries * sizeof(int16_t))) return -1;
  const int ctas = std::min((L->ntasks + BW - 1) / BW, c.sm_count * 8);
  k_coeff_bits<<<ctas, BW * 32, 0, c.stream>>>((const thorcuda_tb_task *)L->tasks(), L->ntasks, L->pool, L->bits);
  THOR_LAUNCHED();
  if (host_bits) {
    if (d2h_bulk(host_bits, L->bits, (size_t)L->ntasks * 4)) return -1;
    L->late_pool = true;
  }
  return 0;
}

// pure host arithmetic, like thorcuda_clip_mv: the scalar glue of the bit-cost model
int thorcuda_vlc_len(unsigned n, unsigned cn) {                 // quote_vlc (enc/putvlc.c:133-229); -1 where it aborts
  auto lg = [](unsigned v) { int k = 0; while (v >>= 1) k++; return k; };
  if (n <= 5) { const unsigned lim = 6u << n; return (int)cn < (int)lim ? 1 + (int)n + (int)(cn >> n) : 7 - (int)n + 2 * lg(cn - lim + (1u << n)); }
  if (n <= 7) return 1 + (int)n - 4 + (int)(cn >> (n - 4));
  if (n == 8) return cn == 0 ? 1 : cn <= 2 ? 2 : -1;
  if (n == 9) return cn == 0 ? 3 : cn <= 2 ? 4 : cn < 11 ? 5 : 5 + (int)((cn + 5) >> 4);
  if (n == 10) return 1 + 2 * lg(cn + 1);
  if (n == 11) return cn < 2 ? (int)cn + 1 : (int)(cn / 2) + 3;
  if (n == 12) return cn + 1 < 4 ? (int)cn + 1 : 4;
  if (n == 13) return cn + 1 < 6 ? (int)cn + 1 : 6;
  return -1;
}
int thorcuda_mv_bits(int mvx, int mvy, int mvpx, int mvpy) {     // write_mv (enc/write_bits.c:50-69) = quote_mv_bits (enc/encode_block.c:799-814)
  const int dx = mvx - mvpx, dy = mvy - mvpy;
  return thorcuda_vlc_len(10, (unsigned)(2 * abs(dx) - (dx < 0))) + thorcuda_vlc_len(10, (unsigned)(2 * abs(dy) - (dy < 0)));
}

}  // extern "C"
