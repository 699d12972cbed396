// libthorcuda: bit-cost model on the device, first slice -- the exact number of bits the reference's
// entropy coder spends on the quantised coefficients of a transform block.
//
// Restates reference enc/write_bits.c:110-253 (write_coeff), :71-108 (find_code) and the code lengths of
// enc/putvlc.c:133-229 (quote_vlc, tables 0, 1, 2 and 10 -- the ones write_coeff uses) as a pure length computation
// over the coefficient pool of a transform-block list, i.e. on the levels thorcuda_encode_list has just produced: the
// encoder's cost_calc (enc/encode_block.c:1218-1228) needs `nbits` of every trial encode, and the coefficients are
// the part of it that scales with the pixels (mode, vector and cbp bits are a few table look-ups per block).
//
// One warp per transform block.  The low q x q corner (q = min(size, 16)) is read coalesced and scattered into scan
// order in shared memory; ballots give the non-zero map and the last position.  write_coeff's two-mode scan (level
// mode after a level > 1, run mode after a zero or a +-1) is inherently sequential in the number of CODED symbols, so
// it is replayed warp-uniformly, but run mode jumps from one non-zero level to the next through the map instead of
// stepping over the zeros.
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
  __shared__ uint32_t M[BW][8];           // non-zero map, bit p = scan position p
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int16_t *s = S[warp];
  uint32_t *nzmap = M[warp];
  for (int t = blockIdx.x * BW + warp; t < ntasks; t += gridDim.x * BW) {
    const thorcuda_tb_task task = tasks[t];
    const int size = task.size, q = min(size, 16), n = q * q;
    const bool chroma = task.coeff_type & 1, intra = (task.coeff_type >> 1) & 1;
    const uint8_t *scan = c_scan + (q == 4 ? 0 : q == 8 ? 16 : 80);
    const int16_t *c = pool + task.coeff_offset;
    for (int r = lane; r < n; r += 32) s[scan[r]] = c[r];
    __syncwarp();
    int last = -1;
    for (int k = 0; k < (n + 31) / 32; k++) {
      const int p = 32 * k + lane;
      const uint32_t m = __ballot_sync(0xffffffffu, p < n && s[p] != 0);
      if (lane == 0) nzmap[k] = m;
      if (m) last = 32 * k + ilog2u(m);
    }
    __syncwarp();
    int total = 0;
    if (last >= 0) {
      const int first = abs((int)s[0]);
      if (chroma && last == 0 && first == 1) total = 2;                   // :148-153: flag + sign, nothing else
      else {
        total = chroma ? 1 : 0;
        int adaptive = intra && !chroma;                                  // table of the next level-mode symbol
        bool level_mode = true;
        int pos = 0, maxrun = 0;
        const bool small_chroma = chroma && size <= 8;
        while (pos <= last) {
          if (level_mode) {                                               // :166-178
            const int level = abs((int)s[pos]);
            total += vlc_len(adaptive, (uint32_t)level) + (level > 0);
            if (!chroma) adaptive = level > 3;
            pos++;
            if (level == 0) { level_mode = false; maxrun = n - pos - 1; }
          } else {                                                        // :181-218, zeros skipped through the map
            int w = pos >> 5;
            uint32_t m = nzmap[w] & (0xffffffffu << (pos & 31));
            while (!m) m = nzmap[++w];
            const int nz = 32 * w + __ffs(m) - 1, run = nz - pos, v = s[nz], level = abs(v);
            total += event_len(find_code(run, level, maxrun, chroma), small_chroma, false);
            total += level > 1 ? vlc_len(0, (uint32_t)(2 * (level - 2) + (v < 0))) : 1;
            pos = nz + 1;
            level_mode = level > 1;
            maxrun = n - pos - 1;
          }
        }
        if (pos < n && level_mode) { total += vlc_len(adaptive, 0); pos++; }   // :222-234
        if (pos < n) total += event_len(chroma ? 0 : 2, small_chroma, true);      // :237-250
      }
    }
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
