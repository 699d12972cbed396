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

// ---- write_block (enc/write_bits.c:364-650) as a length: the syntax around the coefficients ----
// One thread per block record: a few table look-ups and exp-Golomb lengths, then the coefficient bits of the block's
// transform blocks (k_coeff_bits' output) where the cbp bit says write_block codes them.
__device__ __forceinline__ int mvd_len(short2 mv, short2 p) {                      // write_mv :50-69
  const int dx = mv.x - p.x, dy = mv.y - p.y;
  return expgolomb_len((uint32_t)(2 * abs(dx) - (dx < 0))) + expgolomb_len((uint32_t)(2 * abs(dy) - (dy < 0)));
}
__device__ __forceinline__ int mv_list_len(const short2 *mv, short2 p, int part) {   // :421-444, :468-490
  int n = mvd_len(mv[0], p);
  if (part == 1) n += mvd_len(mv[2], mv[0]);
  else if (part == 2) n += mvd_len(mv[1], mv[0]);
  else if (part == 3) n += mvd_len(mv[1], mv[0]) + mvd_len(mv[2], mv[0]) + mvd_len(mv[3], mv[0]);
  return n;
}
__device__ __forceinline__ int idx_len(int nvec, int idx) { return nvec == 4 ? 2 : nvec == 3 ? (idx == 0 ? 1 : 2) : nvec == 2 ? 1 : 0; }   // :507-530
__device__ __forceinline__ int cbp_code(int y, int u, int v, int ctx_cbp) {        // :382, :561-566
  const int code = (0x47362501u >> (4 * (y + 2 * u + 4 * v))) & 7;               // {1,0,5,2,6,3,7,4}
  return ctx_cbp == 0 && code < 2 ? 1 - code : code;
}
__device__ int super_mode_len(const thorcuda_block_syntax &b) {                    // write_super_mode :268-362, split_flag 0
  if (b.frame_type == 0) return b.size > 8;
  const int bip = b.num_ref > 1 && b.enable_bipred, spl = b.size > 8, maxbit = 2 + b.num_ref + spl + bip;
  const bool demote = (b.ctx_index == 2 || b.ctx_index > 3) && spl;
  const int m = b.mode, r0 = b.ref_idx0;
  int code;
  if (b.interp_ref) {
    code = m == 0 ? 0 : m == 4 ? 2 : m == 3 ? 3 : m == 1 ? 4 : r0 > 0 ? 4 + r0 : 4 + b.num_ref;
    if (!bip && code > 3) code--;
    if (!spl && code > 1) code--;
    if (demote && code < 3) code = (code + 2) % 3;
  } else {
    code = m == 0 ? 0 : m == 4 ? 3 : m == 3 ? 4 : m == 1 ? 5 : r0 > 0 ? 5 + r0 : 2;
    if (!bip && code > 4) code--;
    if (!spl && code > 1) code--;
    if (demote && code < 4) code = (code + 3) & 3;
  }
  return code == maxbit ? maxbit : code + 1;
}

static_assert(sizeof(thorcuda_block_syntax) == 64, "block records are read as four 16-byte words");
__global__ void __launch_bounds__(128) k_block_bits(const thorcuda_block_syntax *__restrict__ blocks, int n,
                                                    const uint32_t *__restrict__ tb_bits, int ntb, uint32_t *__restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  __align__(16) thorcuda_block_syntax b;
  {
    const uint4 *src = reinterpret_cast<const uint4 *>(blocks + i);
    uint4 *dst = reinterpret_cast<uint4 *>(&b);
#pragma unroll
    for (int k = 0; k < 4; k++) dst[k] = src[k];
  }
  const short2 mvp = make_short2(b.mvp.x, b.mvp.y);
  const short2 *mv0 = reinterpret_cast<const short2 *>(b.mv0), *mv1 = reinterpret_cast<const short2 *>(b.mv1);
  const int mode = b.mode, size = b.size, part = b.pb_part;
  int bits = super_mode_len(b);
  if (size == 64 && mode != 0 && b.max_delta_qp) bits += vlc_len(0, (uint32_t)abs((int)b.delta_qp)) + (b.delta_qp != 0);   // :389-391
  if (mode == 1) {                                                                 // intra mode :393-413
    const int im = b.intra_mode;
    if (b.num_intra_modes <= 4) bits += 2;
    else if (b.num_intra_modes <= 8) bits += im == 0 || im == 2 || im == 3 ? 2 : im == 4 || im == 8 || im == 9 ? 4 : 5;   // map {2,8,1,0,5,9,7,6,4,3}, len {2,2,2,4,4,4,5,5}
    else bits += im < 2 ? 3 : im < 4 ? 2 : im >= 8 ? 4 : 5;                       // map {2,3,1,0,6,9,8,7,5,4}, len {2,2,3,3,4,4,5,5,5,5}
  } else if (mode == 2) {
    if (b.max_num_pb_part > 1) bits += part == 0 ? 1 : part == 1 ? 2 : 3;
    bits += mv_list_len(mv0, mvp, part);
  } else if (mode == 3) {                                                          // :446-505 (BIPRED_PART 0)
    short2 p = mvp;
    if (part == 0) bits += mvd_len(mv0[0], p);
    if (b.frame_type == 2) p = mv0[0];
    bits += mv_list_len(mv1, p, part);
    if (b.frame_type == 1) {
      if (b.num_ref == 2) { const int c = 2 * b.ref_idx0 + b.ref_idx1; bits += c == 3 ? 3 : c + 1; }
      else bits += expgolomb_len((uint32_t)(4 * b.ref_idx0 + b.ref_idx1));
    }
  } else if (mode == 0) {
    bits += idx_len(b.num_skip_vec, b.skip_idx);
  } else {
    bits += idx_len(b.num_merge_vec, b.skip_idx);
  }
  if (mode != 0) {
    const int cy = b.cbp_y, cu = b.cbp_u, cv = b.cbp_v, split = b.tb_split;
    const int max_tb = mode == 2 ? (b.max_num_tb_part > 1 ? 2 : 1) : mode == 1 ? b.max_num_tb_part : 1;
    int code;
    if (max_tb > 1) {
      if (split) code = 2;
      else { code = cbp_code(cy, cu, cv, b.ctx_cbp); if (code > 1) code++; }
    } else if (mode == 4) {
      code = (0x47362501u >> (4 * (cy + 2 * cu + 4 * cv))) & 7;
      code = code == 1 ? 7 : code > 1 ? code - 1 : code;
    } else code = cbp_code(cy, cu, cv, b.ctx_cbp);
    bits += vlc_len(0, (uint32_t)code);
    const uint32_t *tb = tb_bits ? tb_bits + b.tb_first : nullptr;
    const bool have = tb_bits && b.tb_first != 0xffffffffu;
    if (!split) {                                                                  // :572-582  Y U V
      if (have && (size_t)b.tb_first + 3 <= (size_t)ntb) bits += (cy ? tb[0] : 0) + (cu ? tb[1] : 0) + (cv ? tb[2] : 0);
    } else if (size > 8) {                                                         // :585-612  (Y U V) per quadrant
      for (int q = 0; q < 4; q++) {
        const int y = (cy >> (3 - q)) & 1, u = (cu >> (3 - q)) & 1, v = (cv >> (3 - q)) & 1;
        bits += vlc_len(0, (uint32_t)cbp_code(y, u, v, b.ctx_cbp));
        if (have && (size_t)b.tb_first + 12 <= (size_t)ntb) bits += (y ? tb[3 * q] : 0) + (u ? tb[3 * q + 1] : 0) + (v ? tb[3 * q + 2] : 0);
      }
    } else {                                                                       // :613-641  Y0..Y3 U V
      const int c = cu + 2 * cv;
      bits += 4 + (c == 0 ? 1 : c == 1 ? 2 : 3);
      if (have && (size_t)b.tb_first + 6 <= (size_t)ntb) {
        for (int q = 0; q < 4; q++) bits += ((cy >> (3 - q)) & 1) ? tb[q] : 0;
        bits += (cu ? tb[4] : 0) + (cv ? tb[5] : 0);
      }
    }
  }
  out[i] = (uint32_t)bits;
}

static DevBuf g_blk_in, g_blk_out;

// Scratch of thorcuda_block_bits_static: the records of one call and their results, in a file-scope device array -- the
// entry point must not allocate.  Every GPU that loads the library holds these 272 KiB until exit.  In host code the name
// below is only a stand-in for the device symbol: its device address comes from cudaGetSymbolAddress, once per device.
static __device__ __align__(256) uint8_t g_bb_static[THORCUDA_BLOCK_BITS_STATIC_MAX * (sizeof(thorcuda_block_syntax) + sizeof(uint32_t))];
static uint8_t *g_bb_static_ptr[64];          // device address of g_bb_static, per device ordinal (nullptr: not looked up yet)
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
void bits_reset() { scan_ready = false; g_blk_in.release(); g_blk_out.release(); }

// what k_block_bits would otherwise have to guard per thread; L: the list tb_first indexes, or nullptr
static int validate_blocks(const char *who, const thorcuda_block_syntax *host_blocks, int nblocks, const List *L) {
  for (int i = 0; i < nblocks; i++) {
    const thorcuda_block_syntax &b = host_blocks[i];
    if (b.size != 8 && b.size != 16 && b.size != 32 && b.size != 64) return fail("%s: block %d has size %d", who, i, b.size);
    if (b.mode > 4 || b.frame_type > 2 || b.pb_part > 3 || b.intra_mode > 9) return fail("%s: block %d is malformed", who, i);
    if (!b.tb_split && (b.cbp_y | b.cbp_u | b.cbp_v) > 1) return fail("%s: block %d has cbp masks without tb_split", who, i);
    if (L && b.mode != 0 && b.tb_first != 0xffffffffu) {
      const uint32_t cnt = !b.tb_split ? 3 : b.size > 8 ? 12 : 6;
      if ((uint64_t)b.tb_first + cnt > (uint64_t)L->ntasks) return fail("%s: block %d: transform blocks %u..%u outside the list", who, i, b.tb_first, b.tb_first + cnt - 1);
    }
  }
  return 0;
}

}  // namespace thor

using namespace thor;

extern "C" {

int thorcuda_coeff_bits_list(thorcuda_list hl, const int16_t *host_coeffq, uint32_t *host_bits) {
  THOR_NEED_CTX();
  Context &c = ctx();
  List *L = get_list(hl);
  if (!L || L->kind != THORCUDA_LIST_TB) return fail("coeff_bits: not a transform-block list");
  if (list_stale(L)) return -1;
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

int thorcuda_block_bits(thorcuda_list hl, const thorcuda_block_syntax *host_blocks, int nblocks, uint32_t *host_bits) {
  THOR_NEED_CTX();
  Context &c = ctx();
  if (nblocks <= 0) return 0;
  if (!host_blocks || !host_bits) return fail("block_bits: null argument");
  List *L = nullptr;
  if (hl >= 0) {
    L = get_list(hl);
    if (!L || L->kind != THORCUDA_LIST_TB) return fail("block_bits: not a transform-block list");
  }
  if (validate_blocks("block_bits", host_blocks, nblocks, L)) return -1;
  // the previous call's output copy may still be reading g_blk_out
  if (copies_fence()) return -1;
  auto *d_in = (thorcuda_block_syntax *)g_blk_in.reserve((size_t)nblocks * sizeof(thorcuda_block_syntax));
  auto *d_out = (uint32_t *)g_blk_out.reserve((size_t)nblocks * 4);
  if (!d_in || !d_out) return fail("block_bits: device allocation failed");
  if (h2d(d_in, host_blocks, (size_t)nblocks * sizeof(thorcuda_block_syntax))) return -1;
  k_block_bits<<<(nblocks + 127) / 128, 128, 0, c.stream>>>(d_in, nblocks, L ? L->bits : nullptr, L ? L->ntasks : 0, d_out);
  THOR_LAUNCHED();
  return d2h(host_bits, d_out, (size_t)nblocks * 4);
}

// The syntax bits of write_block without any allocation by the library: see thorcuda.h.  Synchronous.
int thorcuda_block_bits_static(const thorcuda_block_syntax *host_blocks, int nblocks, uint32_t *host_bits) {
  THOR_NEED_CTX();
  Context &c = ctx();
  if (nblocks <= 0) return 0;
  if (!host_blocks || !host_bits) return fail("block_bits_static: null argument");
  if (nblocks > THORCUDA_BLOCK_BITS_STATIC_MAX)
    return fail("block_bits_static: %d blocks exceed the static scratch of %d (THORCUDA_BLOCK_BITS_STATIC_MAX); split the call or use thorcuda_block_bits", nblocks,
                THORCUDA_BLOCK_BITS_STATIC_MAX);
  if (c.device < 0 || c.device >= 64) return fail("block_bits_static: device ordinal %d", c.device);
  if (validate_blocks("block_bits_static", host_blocks, nblocks, nullptr)) return -1;
  if (!g_bb_static_ptr[c.device]) {             // set-up, once per device, with that device current (thorcuda_init made it so)
    void *p = nullptr;
    THOR_CUDA(cudaGetSymbolAddress(&p, g_bb_static));
    g_bb_static_ptr[c.device] = (uint8_t *)p;
  }
  auto *d_in = (thorcuda_block_syntax *)g_bb_static_ptr[c.device];
  auto *d_out = (uint32_t *)(g_bb_static_ptr[c.device] + (size_t)THORCUDA_BLOCK_BITS_STATIC_MAX * sizeof(thorcuda_block_syntax));
  // the array keeps the previous call's contents; nothing in it has to start at zero: the first nblocks records and results are
  // overwritten before they are read.  Straight copies from / to the caller's buffers (no staging ring, which may allocate).
  THOR_CUDA(cudaMemcpyAsync(d_in, host_blocks, (size_t)nblocks * sizeof(thorcuda_block_syntax), cudaMemcpyHostToDevice, c.stream));
  k_block_bits<<<(nblocks + 127) / 128, 128, 0, c.stream>>>(d_in, nblocks, nullptr, 0, d_out);
  THOR_LAUNCHED();
  THOR_CUDA(cudaMemcpyAsync(host_bits, d_out, (size_t)nblocks * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
  THOR_CUDA(cudaStreamSynchronize(c.stream));
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
